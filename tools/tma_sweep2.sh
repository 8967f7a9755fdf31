#!/bin/bash
# short A/B list for epilogue work on the TMA engine
run() { echo "--- $*"; S2LC_TMA_ALL=1 python tools/bench_kernel.py "$@" --plain | sed 's/^/TMA    /'; }
run conv3x3 --n 32 --h 32 --w 32 --cin 512 --cout 512
run conv3x3 --n 32 --h 64 --w 64 --cin 256 --cout 256
run conv3x3 --n 32 --h 128 --w 128 --cin 128 --cout 128
run conv3x3 --n 32 --h 256 --w 256 --cin 64 --cout 64
run conv1x1 --n 32 --h 16 --w 16 --cin 304 --cout 1824
run conv1x1 --n 32 --h 32 --w 32 --cin 176 --cout 1056
run conv1x1 --n 32 --h 32 --w 32 --cin 128 --cout 768
run conv1x1 --n 32 --h 64 --w 64 --cin 64 --cout 384
run conv1x1 --n 32 --h 128 --w 128 --cin 40 --cout 240
run conv1x1 --n 32 --h 256 --w 256 --cin 24 --cout 144
run conv1x1 --n 32 --h 256 --w 256 --cin 144 --cout 24
