#!/bin/bash
# TEST INFRASTRUCTURE: every parity case through the host build of the kernel functors + engine (tests/hostsim) compiled
# with AddressSanitizer: out-of-bounds gathers / staging writes show up here without a GPU (compute-sanitizer is not
# available on the GPU pool).  usage: bash tools/asan_hostsim.sh [case ...]
set -e
cd "$(dirname "$0")/.."
make -C oracle >/dev/null
g++ -std=c++17 -O1 -g -fsanitize=address -fno-omit-frame-pointer -ffp-contract=off -fPIC -shared -x c++ \
    -o /tmp/libhostsim_asan.so tests/hostsim/hostsim.cpp
ASAN_OPTIONS=detect_leaks=0 LD_PRELOAD=$(g++ -print-file-name=libasan.so) python tools/asan_hostsim.py "$@"
