#!/bin/bash
# build an experiment variant of the library: tools/build_variant.sh NAME -DFLAG=VALUE ...  -> build/variants/NAME.so
set -e
cd "$(dirname "$0")/.."
mkdir -p build/variants
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -Xlinker --exclude-libs,ALL \
     "$@" -shared -o build/variants/$name.so ewn_gym_b200/csrc/ewn_kernels.cu
