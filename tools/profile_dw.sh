#!/bin/bash
# ncu --set full captures of the depthwise kernels on their production shapes.  Usage (GPU box): bash tools/profile_dw.sh r02m
tag=${1:-rXX}
cap() {  # name kernel-regex bench_kernel args...
  local name=$1 rx=$2; shift 2
  python tools/bench_kernel.py "$@" --reps 2 > gpurun_out/${tag}_${name}.log 2>&1 \
   && ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/${tag}_${name} \
      python tools/bench_kernel.py "$@" --reps 2 >> gpurun_out/${tag}_${name}.log 2>&1
}
cap dw2_bwd_k5s1 dw2_bwd dwbwdf --n 32 --h 64 --w 64 --cin 384 --k 5 --s 1
cap dw2_bwd_k3s1 dw2_bwd dwbwdf --n 32 --h 128 --w 128 --cin 240 --k 3 --s 1
cap dw2_bwd_k3s2 dw2_bwd dwbwdf --n 32 --h 256 --w 256 --cin 144 --k 3 --s 2
cap dw2_fwd_k5s1 dw2_fwd dwfwd --n 32 --h 64 --w 64 --cin 384 --k 5 --s 1
ls -la gpurun_out/ | grep ${tag}
