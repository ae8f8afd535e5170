#!/bin/bash
# Host-side stand-in for the reference's compute-sanitizer matrix (components/eamxx/scripts/test_factory.py:155-235),
# which cannot run here (compute-sanitizer is closed on this GPU pool, profiles/r02_compute_sanitizer_unavailable.txt):
# the column driver and level physics the sm_100a kernels instantiate (p3_column.cuh, p3_physics.cuh) compiled for the
# host (tests/hostsim, one lane per warp) under AddressSanitizer + UndefinedBehaviorSanitizer, run on the random and
# mixed columns of the CPU test-suite.  Catches out-of-bounds level/window indexing, uninitialised-read-driven UB,
# signed overflow and misaligned access in the shared source; it cannot see races between lanes.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/build/libp3hostsim_asan.so
g++ -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined -march=x86-64-v3 -std=c++17 -fPIC \
    -ffp-contract=off -shared -I/usr/local/cuda/include -DP3B200_EXACT_MATH -o $OUT \
    $ROOT/tests/hostsim/p3_hostsim.cpp $ROOT/scream_b200/csrc/p3_host.cpp
cd $ROOT
P3B200_HOSTSIM_EXACT_LIB=$OUT LD_PRELOAD=$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so) \
  ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
  python -m pytest tests/test_cpu_hostsim.py -q -x -k "not product and not parity" 2>&1 | tail -15
