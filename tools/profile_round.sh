#!/bin/bash
# One profiling pass for profiles/: per-launch times, ncu launch list of a bench step, and ncu --set full captures of
# the depthwise kernels on their biggest layer.  Usage (on the GPU box): bash tools/profile_round.sh r01c
tag=${1:-rXX}
python tools/launch_times.py --top 90 --out gpurun_out/${tag}_launch_times.txt > /dev/null 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-breakdown > gpurun_out/${tag}_bench_short.json 2>/dev/null \
 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-breakdown > gpurun_out/${tag}_ncu_bench.log 2>&1
python tools/bench_kernel.py dwbwdf --n 32 --h 256 --w 256 --cin 144 --k 3 --s 2 --reps 2 > gpurun_out/${tag}_k1.log 2>&1 \
 && ncu --set full --clock-control none --import-source on -k regex:dw2_bwd -s 1 -c 1 -f -o gpurun_out/${tag}_dw2_bwd_k3s2 \
    python tools/bench_kernel.py dwbwdf --n 32 --h 256 --w 256 --cin 144 --k 3 --s 2 --reps 2 >> gpurun_out/${tag}_k1.log 2>&1
python tools/bench_kernel.py dwfwd --n 32 --h 128 --w 128 --cin 240 --k 3 --s 1 --reps 2 > gpurun_out/${tag}_k2.log 2>&1 \
 && ncu --set full --clock-control none --import-source on -k regex:dw2_fwd -s 1 -c 1 -f -o gpurun_out/${tag}_dw2_fwd_k3s1 \
    python tools/bench_kernel.py dwfwd --n 32 --h 128 --w 128 --cin 240 --k 3 --s 1 --reps 2 >> gpurun_out/${tag}_k2.log 2>&1
python tools/bench_kernel.py dwbwdf --n 32 --h 64 --w 64 --cin 384 --k 5 --s 1 --reps 2 > gpurun_out/${tag}_k3.log 2>&1 \
 && ncu --set full --clock-control none --import-source on -k regex:dw2_bwd -s 1 -c 1 -f -o gpurun_out/${tag}_dw2_bwd_k5s1 \
    python tools/bench_kernel.py dwbwdf --n 32 --h 64 --w 64 --cin 384 --k 5 --s 1 --reps 2 >> gpurun_out/${tag}_k3.log 2>&1
ls -la gpurun_out/ | tail -12
