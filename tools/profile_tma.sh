#!/bin/bash
# ncu --set full captures of the TMA-fed engines on production shapes (profiles/r02*_ncu_*.txt are the summaries).
# Usage (GPU box): bash tools/profile_tma.sh r02e
tag=${1:-rXX}
cap() {  # name kernel-regex bench_kernel args...
  local name=$1 rx=$2; shift 2
  python tools/bench_kernel.py "$@" --plain --reps 2 > gpurun_out/${tag}_${name}.log 2>&1 \
   && ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/${tag}_${name} \
      python tools/bench_kernel.py "$@" --plain --reps 2 >> gpurun_out/${tag}_${name}.log 2>&1
}
cap tconv3x3_32 tconv_kernel conv3x3 --n 32 --h 32 --w 32 --cin 640 --cout 512
cap tconv3x3_128 tconv_kernel conv3x3 --n 32 --h 128 --w 128 --cin 128 --cout 128
cap tconv1x1_16 tconv_kernel conv1x1 --n 32 --h 16 --w 16 --cin 1824 --cout 304
cap twgrad3x3_32 twgrad_kernel wgrad3x3 --n 32 --h 32 --w 32 --cin 640 --cout 512
ls -la gpurun_out/ | grep ${tag}
