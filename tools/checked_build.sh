#!/bin/bash
# the library with our own bounds assertions compiled in (-DEWN_CHECKS, see EWN_CHECK in csrc/ewn_kernels.cu) -> build/checked/libewn_b200.so
# run the GPU tests against it:   EWN_B200_LIB=$PWD/build/checked/libewn_b200.so python -m pytest tests -m gpu -q
set -e
cd "$(dirname "$0")/.."
mkdir -p build/checked
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -Xlinker --exclude-libs,ALL \
     -DEWN_CHECKS -shared -o build/checked/libewn_b200.so ewn_gym_b200/csrc/ewn_kernels.cu
