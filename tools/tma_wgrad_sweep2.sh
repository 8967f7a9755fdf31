#!/bin/bash
# rolling-row weight-gradient engine: TMA producer vs loader warps on plain operands (wide 3x3 layers)
run() { echo "--- $*"; S2LC_NO_TMA=1 python tools/bench_kernel.py "$@" --plain | sed 's/^/noTMA  /'; python tools/bench_kernel.py "$@" --plain | sed 's/^/TMA    /'; }
run wgrad3x3 --n 32 --h 128 --w 128 --cin 168 --cout 128
run wgrad3x3 --n 32 --h 128 --w 128 --cin 128 --cout 128
run wgrad3x3 --n 32 --h 256 --w 256 --cin 88 --cout 64
run wgrad3x3 --n 32 --h 256 --w 256 --cin 64 --cout 64
run wgrad3x3 --n 32 --h 512 --w 512 --cin 38 --cout 32
run wgrad3x3 --n 32 --h 512 --w 512 --cin 32 --cout 32
