#!/bin/bash
# Both simulator cores (sim_core.cuh, par_core.cuh: the kernel bodies) built for the host with AddressSanitizer and
# UndefinedBehaviorSanitizer, then the host test-suites incl. the differential fuzzing (compute-sanitizer is not
# available on the GPU pool).  Restores the normal harness afterwards.
set -e
cd "$(dirname "$0")/../tests/host_harness"
g++ -O1 -g -std=c++17 -ffp-contract=off -fno-fast-math -fPIC -shared -fsanitize=address,undefined \
    -fno-sanitize-recover=undefined -fno-omit-frame-pointer -Wno-unused-function -Wno-unknown-pragmas -o libhost_sim.so host_sim.cpp
cd ../..
LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 python -m pytest tests/test_core_host.py \
    tests/test_par_host.py tests/test_fuzz_engines.py -x -q -p no:cacheprovider
touch tests/host_harness/host_sim.cpp
python -c "import sys; sys.path.insert(0, '.'); from tests.host_harness import host_sim; host_sim.build()"
