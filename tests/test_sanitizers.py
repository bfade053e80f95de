"""The hand-written host parsers (json.cc, loader.cc, quantity.cc, the host build of pod_scan.cuh) under
AddressSanitizer + UndefinedBehaviorSanitizer: tools/asan_tests.sh builds libarena_gpuacct_asan.so and
re-runs the parser fuzz tests against it in a child process.  (The full list the script runs by default --
fixtures, packing, sharding -- takes a minute longer; here the fuzzers.)"""

import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parser_fuzz_tests_are_clean_under_asan_and_ubsan(built):
    if not shutil.which("gcc") or not os.path.exists(subprocess.run(["gcc", "-print-file-name=libasan.so"], capture_output=True,
                                                                      text=True).stdout.strip()):
        pytest.skip("no libasan in this toolchain")
    if "ARENA_GPUACCT_LIB" in os.environ:
        pytest.skip("already running against an alternate build")
    p = subprocess.run(["bash", os.path.join(ROOT, "tools", "asan_tests.sh"), "tests/test_pod_scan.py", "tests/test_loader_scan.py",
                        "tests/test_loader_full_scan.py"],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-2000:])
    assert "passed" in p.stdout and "runtime error" not in p.stderr and "AddressSanitizer" not in p.stderr
