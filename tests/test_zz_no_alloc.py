"""The allocation-free `top job` entry point, through the C ABI.

arena_top_job_device_noalloc exists for callers that cannot allocate device
memory while querying (a polling agent on a fixed memory budget, a host whose
CUDA context forbids allocation after start-up): its work arrays sit in a fixed
device array inside the library.  Because every call reuses that array -- and
finds the previous call's data in it -- the check is made on TWO calls in a
row with different inputs, against the regular (allocating) entry point.
Integers: the outputs must be identical, as in the other GPU tests.

Kept in a file that sorts last so that a `-x` run has already executed every
other test when it gets here.
"""

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_cols(torch, host):
    return {k: torch.from_numpy(v.copy()).cuda() for k, v in host.cols.items() if v.size or k == "pod_ctr_begin"}


def _run_noalloc(gpu, torch, cols, J):
    d_req = torch.full((J,), -7, dtype=torch.int64, device="cuda")       # garbage in: the call must not rely on zeros
    d_alloc = torch.full((J,), -7, dtype=torch.int64, device="cuda")
    d_order = torch.zeros(J, dtype=torch.int32, device="cuda")
    d_latest = torch.zeros(J, dtype=torch.int32, device="cuda")
    d_totals = torch.full((2,), -7, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    gpu.top_job_device_noalloc(cols, J, d_req, d_alloc, d_order, d_latest, d_totals, st)
    torch.cuda.synchronize()
    return (d_req.cpu().numpy(), d_alloc.cpu().numpy(), d_order.cpu().numpy().view(np.uint32),
            d_latest.cpu().numpy().view(np.uint32), d_totals.cpu().numpy())


def _run_regular(gpu, host):
    dev = gpu.DeviceSnapshot.upload(host)
    r = gpu.top_job(dev)
    dev.close()
    return r.job_req, r.job_alloc, r.job_order, r.job_latest, r.totals


def test_noalloc_matches_regular_entry_point_twice(gpu):
    import torch
    # two different snapshots back to back: the second call sees the first one's leftovers
    for P, N, J, seed in [(60_000, 500, 2_048, 11), (9_000, 90, 37, 12), (60_000, 500, 2_048, 11)]:
        host = gpu.HostSnapshot.synth(P, N, seed=seed, n_jobs=J)
        want = _run_regular(gpu, host)
        got = _run_noalloc(gpu, torch, _device_cols(torch, host), J)
        for name, g, w in zip(("job_req", "job_alloc", "job_order", "job_latest", "totals"), got, want):
            assert np.array_equal(g, w), f"{name} differs (P={P}, J={J})"


def test_noalloc_rejects_more_jobs_than_its_fixed_array_holds(gpu):
    import torch
    J = 2_049
    host = gpu.HostSnapshot.synth(5_000, 50, seed=13, n_jobs=J)
    cols = _device_cols(torch, host)
    z64 = torch.zeros(J, dtype=torch.int64, device="cuda")
    z32 = torch.zeros(J, dtype=torch.int32, device="cuda")
    with pytest.raises(gpu.ArenaError) as ei:
        gpu.top_job_device_noalloc(cols, J, z64, z64.clone(), z32, z32.clone(),
                                   torch.zeros(2, dtype=torch.int64, device="cuda"))
    assert ei.value.code == -5 and "at most 2048 jobs" in str(ei.value)
