#!/bin/bash
# deep 1x1 forward convs of b5 / bs=32 / 512x512 tiles (project convs with SE gate: cin = expanded width)
for cfg in "64 64 384 64" "64 64 64 384" "32 32 768 128" "32 32 128 768" "32 32 1056 176" "32 32 176 1056" "16 16 1824 304" "16 16 304 1824" "16 16 3072 512" "128 128 240 40"; do
  set -- $cfg
  echo -n "H=$1 W=$2 $3->$4: "
  python tools/bench_kernel.py conv1x1 --n 32 --h $1 --w $2 --cin $3 --cout $4 --reps 10 2>&1 | grep "ms " | tail -1
done
