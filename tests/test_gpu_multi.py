"""Multi-GPU parity on the box: tools/multigpu_check.py under torchrun with as many ranks as there
are GPUs (at most 8).  On a one-GPU box it still runs, with a world of one (the exchange talks to
itself); the gloo tests in test_multigpu_gloo.py cover the host logic at world size 2."""

import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_rank_agrees_with_the_oracle(gpu):
    import torch
    n = min(8, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "multigpu_check.py"), "3000000"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, (p.stdout[-3000:], p.stderr[-3000:])
    assert "FAIL" not in p.stdout and p.stdout.count("OK") >= 12
