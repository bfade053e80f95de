#!/bin/bash
# The fuzzed, hand-written host parsers under AddressSanitizer + UndefinedBehaviorSanitizer (no GPU needed):
# json.cc, loader.cc, quantity.cc, printers.cc, pack.cc and the host build of pod_scan.cuh, driven by the
# existing fuzz / fixture tests through the C ABI.  Any report aborts the test process.
set -e
cd "$(dirname "$0")/.."
make -C arena_b200/csrc -s -j8 all asan        # (both: the test fixture's own `make` must find nothing to do under LD_PRELOAD)
make -C oracle -s
ASAN=$(gcc -print-file-name=libasan.so)
UBSAN=$(gcc -print-file-name=libubsan.so)
LD_PRELOAD="$ASAN:$UBSAN" ASAN_OPTIONS=detect_leaks=0:abort_on_error=1 UBSAN_OPTIONS=halt_on_error=1:print_stacktrace=1 \
ARENA_GPUACCT_LIB=$PWD/arena_b200/libarena_gpuacct_asan.so \
python -m pytest -x -q -m "not gpu" -p no:cacheprovider ${@:-tests/test_pod_scan.py tests/test_loader_scan.py tests/test_quantity.py \
    tests/test_fixtures.py tests/test_packed.py tests/test_multigpu_gloo.py tests/test_loader_full_scan.py}
