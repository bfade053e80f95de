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

from oracle import cref

pytestmark = pytest.mark.gpu


def _device_cols(torch, host):
    return {k: torch.from_numpy(v.copy()).cuda() for k, v in host.cols.items() if v.size or k == "pod_ctr_begin"}


def _run_noalloc(gpu, torch, cols, J, n_pods=None):
    d_req = torch.full((J,), -7, dtype=torch.int64, device="cuda")       # garbage in: the call must not rely on zeros
    d_alloc = torch.full((J,), -7, dtype=torch.int64, device="cuda")
    d_order = torch.zeros(J, dtype=torch.int32, device="cuda")
    d_latest = torch.zeros(J, dtype=torch.int32, device="cuda")
    d_totals = torch.full((2,), -7, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    gpu.top_job_device_noalloc(cols, J, d_req, d_alloc, d_order, d_latest, d_totals, st, n_pods=n_pods)
    torch.cuda.synchronize()
    return (d_req.cpu().numpy(), d_alloc.cpu().numpy(), d_order.cpu().numpy().view(np.uint32),
            d_latest.cpu().numpy().view(np.uint32), d_totals.cpu().numpy())


def _run_regular(gpu, host):
    dev = gpu.DeviceSnapshot.upload(host)
    r = gpu.top_job(dev)
    dev.close()
    return r.job_req, r.job_alloc, r.job_order, r.job_latest, r.totals


def _oracle(host, J):
    """The CPU restatement (oracle/acct_ref.c) plus the documented ordering: requested desc, id asc."""
    req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
    order = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
    return req, alloc, order, latest, totals


def test_noalloc_matches_regular_entry_point_twice(gpu):
    import torch
    # different snapshots back to back: every call after the first sees its predecessor's leftovers.
    # The third shape is above the regular entry point's single-launch limit (65 536 pods), so there
    # the two sides run different kernel chains end to end; all are also held against the CPU oracle.
    for P, N, J, seed in [(60_000, 500, 2_048, 11), (9_000, 90, 37, 12), (200_000, 1_500, 2_000, 14),
                          (60_000, 500, 2_048, 11)]:
        host = gpu.HostSnapshot.synth(P, N, seed=seed, n_jobs=J)
        want = _run_regular(gpu, host)
        got = _run_noalloc(gpu, torch, _device_cols(torch, host), J)
        ref = _oracle(host, J)
        for name, g, w, o in zip(("job_req", "job_alloc", "job_order", "job_latest", "totals"), got, want, ref):
            assert np.array_equal(g, w), f"{name} differs from arena_top_job (P={P}, J={J})"
            assert np.array_equal(g, o), f"{name} differs from the CPU oracle (P={P}, J={J})"


def test_noalloc_needs_only_the_job_columns(gpu):
    """A caller that holds nothing but what `top job` reads: no pod_node, no node capacities."""
    import torch
    P, N, J = 30_000, 300, 500
    host = gpu.HostSnapshot.synth(P, N, seed=15, n_jobs=J)
    cols = _device_cols(torch, host)
    for k in ("pod_node", "node_cap_val", "node_cap_scale"):
        cols[k] = torch.zeros(0, dtype=cols[k].dtype, device="cuda")       # numel 0 -> NULL in the view
    got = _run_noalloc(gpu, torch, cols, J, n_pods=P)
    for name, g, o in zip(("job_req", "job_alloc", "job_order", "job_latest", "totals"), got, _oracle(host, J)):
        assert np.array_equal(g, o), f"{name} differs from the CPU oracle"


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
