#!/bin/bash
# representative 1x1 weight gradients of b5 / bs=32 / 512x512 tiles on the streaming engine
for cfg in "256 256 24 144" "256 256 144 24" "256 256 48 24" "128 128 240 40" "128 128 40 240" "64 64 384 64" "32 32 768 128" "32 32 128 768" "16 16 1824 304"; do
  set -- $cfg
  echo -n "H=$1 W=$2 $3->$4: "
  python tools/bench_kernel.py wgrad1x1 --n 32 --h $1 --w $2 --cin $3 --cout $4 --reps 10 2>&1 | grep "ms " | tail -1
done
