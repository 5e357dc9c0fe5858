#!/bin/bash
# one GPU iteration on k_vstep: parity tests, short sweep, ncu full capture with source counters (run under gpurun)
mkdir -p gpurun_out
tag=${1:-iter}
(timeout 400 python -m pytest tests/test_vec_env.py -m gpu -x -q 2>&1 | tail -3) | tee gpurun_out/pytest_$tag.log
EWN_SWEEP_SHORT=1 timeout 120 python tools/step_sweep.py 2>&1 | tail -6 | tee gpurun_out/sweep_$tag.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_vstep -s 12 -c 1 -o gpurun_out/k_vstep_$tag -f python bench.py --steps 20 --warmup 10 --skip-cpu --skip-clocks --skip-e2e --skip-playouts --skip-search --skip-rollout > gpurun_out/ncu_$tag.log 2>&1
