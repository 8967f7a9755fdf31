#!/bin/bash
# TMA weight-gradient engine (twgrad_tc.cu) vs the loader-warp engines on plain operands, production shapes
run() { echo "--- $*"; S2LC_NO_TMA=1 python tools/bench_kernel.py "$@" --plain | sed 's/^/noTMA  /'; S2LC_TMA_ALL=1 python tools/bench_kernel.py "$@" --plain | sed 's/^/TMA    /'; }
run wgrad3x3 --n 32 --h 32 --w 32 --cin 640 --cout 512
run wgrad3x3 --n 32 --h 32 --w 32 --cin 512 --cout 512
run wgrad3x3 --n 32 --h 64 --w 64 --cin 320 --cout 256
run wgrad3x3 --n 32 --h 64 --w 64 --cin 256 --cout 256
run wgrad3x3 --n 32 --h 128 --w 128 --cin 168 --cout 128
run wgrad3x3 --n 32 --h 128 --w 128 --cin 128 --cout 128
run wgrad3x3 --n 32 --h 256 --w 256 --cin 64 --cout 64
run wgrad1x1 --n 32 --h 16 --w 16 --cin 304 --cout 1824
run wgrad1x1 --n 32 --h 16 --w 16 --cin 1824 --cout 304
run wgrad1x1 --n 32 --h 16 --w 16 --cin 512 --cout 3072
run wgrad1x1 --n 32 --h 16 --w 16 --cin 3072 --cout 512
run wgrad1x1 --n 32 --h 16 --w 16 --cin 512 --cout 2048
run wgrad1x1 --n 32 --h 32 --w 32 --cin 176 --cout 1056
run wgrad1x1 --n 32 --h 32 --w 32 --cin 1056 --cout 176
run wgrad1x1 --n 32 --h 32 --w 32 --cin 128 --cout 768
run wgrad1x1 --n 32 --h 32 --w 32 --cin 768 --cout 128
run wgrad1x1 --n 32 --h 64 --w 64 --cin 64 --cout 384
run wgrad1x1 --n 32 --h 64 --w 64 --cin 384 --cout 64
run wgradT --n 32 --h 16 --w 16 --cin 2048 --cout 512
run wgradT --n 32 --h 32 --w 32 --cin 512 --cout 256
run wgradT --n 32 --h 64 --w 64 --cin 256 --cout 128
run wgradT --n 32 --h 128 --w 128 --cin 128 --cout 64
