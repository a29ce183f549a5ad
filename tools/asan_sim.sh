#!/bin/bash
# compute-sanitizer is closed on this pool (profiles/r2_compute_sanitizer_closed.txt), so the memory checks run where they can:
# every kernel body (the *_body.h files) is compiled for the host by tests/hostsim, and this script rebuilds that simulation
# with AddressSanitizer + UndefinedBehaviorSanitizer (device memory = heap blocks poisoned with 0xA5, so out-of-bounds and
# use-after-free accesses of the bodies and of the engine trap) and runs the parity tests that drive it.
#   tools/asan_sim.sh > profiles/r2_sim_asan.txt 2>&1
set -u
cd "$(dirname "$0")/.."
ASAN=$(gcc -print-file-name=libasan.so)
UBSAN=$(gcc -print-file-name=libubsan.so)
export UVIO_SIM_SANITIZE=1
export ASAN_OPTIONS=detect_leaks=0:abort_on_error=1:halt_on_error=1
export UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
echo "LD_PRELOAD=$ASAN:$UBSAN  (gcc $(gcc -dumpversion))"
LD_PRELOAD="$ASAN:$UBSAN" python -m pytest -x -q -m "not gpu" -p no:cacheprovider \
  tests/test_desc.py tests/test_ahead.py tests/test_hostsim.py tests/test_featdb.py tests/test_tri.py tests/test_lmeds.py tests/test_gpu_cpp_shim.py 2>&1 | tail -15
echo "exit code ${PIPESTATUS[0]}"
