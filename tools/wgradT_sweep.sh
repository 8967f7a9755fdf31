#!/bin/bash
# the five transposed-conv weight gradients of b5 / bs=32 / 512x512 tiles: streaming engine vs slab engine
for cfg in "256 256 64 32" "128 128 128 64" "64 64 256 128" "32 32 512 256" "16 16 2048 512"; do
  set -- $cfg
  for ns in 1 0; do
    echo -n "H=$1 W=$2 $3->$4 S2LC_NO_STREAM=$ns: "
    S2LC_NO_STREAM=$ns python tools/bench_kernel.py wgradT --n 32 --h $1 --w $2 --cin $3 --cout $4 --reps 10 2>&1 | grep "ms " | tail -1
  done
done
