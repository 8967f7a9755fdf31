#!/bin/bash
# per-layer sweep of the depthwise kernels at b5/bs32/512^2 stage shapes
for cfg in "256 24 3 1" "256 48 3 1" "256 144 3 2" "128 240 3 1" "128 240 5 2" "64 384 5 1" "64 384 3 2" "32 768 3 1" "32 1056 5 1" "32 1056 5 2" "16 1824 5 1" "16 3072 3 1"; do
  set -- $cfg
  for what in dwfwd dwbwdf; do
    python tools/bench_kernel.py $what --n 32 --h $1 --w $1 --cin $2 --k $3 --s $4 2>&1 | tail -1
  done
done
