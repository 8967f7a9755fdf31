#!/bin/bash
# ncu --set full captures of the conv-family kernels on representative b5/bs32 shapes.  Usage: bash tools/profile_convs.sh r01d
tag=${1:-rXX}
cap() {  # name, kernel regex, bench_kernel args...
  local name=$1 rx=$2; shift 2
  python tools/bench_kernel.py "$@" --reps 2 > gpurun_out/${tag}_${name}.log 2>&1 \
   && ncu --set full --clock-control none --import-source on -k regex:$rx -s 1 -c 1 -f -o gpurun_out/${tag}_${name} \
      python tools/bench_kernel.py "$@" --reps 2 >> gpurun_out/${tag}_${name}.log 2>&1
}
cap wgrad3x3_64 wgrad_tc wgrad3x3 --n 32 --h 64 --w 64 --cin 256 --cout 256
cap conv1x1_16 conv_tc conv1x1 --n 32 --h 16 --w 16 --cin 1824 --cout 304
cap wgrad1x1_128 "swgrad|wgrad_tc" wgrad1x1 --n 32 --h 128 --w 128 --cin 240 --cout 40
cap conv1x1_128 "sconv|conv_tc" conv1x1 --n 32 --h 128 --w 128 --cin 240 --cout 40
ncu --set full --clock-control none --import-source on -k regex:se_bwd_apply -s 8 -c 1 -f -o gpurun_out/${tag}_se_bwd_apply \
   python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-breakdown > gpurun_out/${tag}_se.log 2>&1
grep -h "ms " gpurun_out/${tag}_*.log | grep -v "^==" | head -20
