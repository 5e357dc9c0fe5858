#!/bin/bash
# n-sweep of K1: T(n) for product / noflush / nostep (fit T0 + slope)
for n in 524288 1048576 2097152 4194304; do
  for v in "" build/variants/noflush.so build/variants/nostep.so; do
    echo -n "n=$n "; EWN_K1_N=$n EWN_K1_GRAPH=0 timeout 300 python tools/k1_sweep.py $v
  done
done
