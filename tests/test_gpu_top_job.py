"""GPU parity for `top job`: per-job requested/allocated, latest-pod rule,
ordering and totals, through the C ABI, against oracle/acct_ref.c."""

import numpy as np
import pytest

from oracle import cref

pytestmark = pytest.mark.gpu


def _check_job(gpu, host):
    J = host.n_jobs
    dev = gpu.DeviceSnapshot.upload(host)
    res = gpu.top_job(dev)
    req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
    assert np.array_equal(res.job_req, req)
    assert np.array_equal(res.job_alloc, alloc)
    assert np.array_equal(res.job_latest, latest)
    assert np.array_equal(res.totals, totals)
    # ordering: requested desc, ties by ascending id (trainer.go:31-33 + documented tie-break)
    want = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
    assert np.array_equal(res.job_order, want)
    res2 = gpu.top_job(dev)          # repeatable on the same handle
    assert np.array_equal(res2.job_req, req) and np.array_equal(res2.job_order, want)
    dev.close()


@pytest.mark.parametrize("P,N,J,seed", [(10_000, 1_000, 1_000, 20180723), (500, 10, 7, 2), (300_000, 3_000, 5_000, 3),
                                        (50_000, 100, 1, 4), (2_000, 50, 4_000, 5)])
def test_top_job_synth(gpu, P, N, J, seed):
    host = gpu.HostSnapshot.synth(P, N, seed=seed, n_jobs=J)
    _check_job(gpu, host)


def test_top_job_negative_and_ties(gpu):
    """Many equal requested values (tie-break), negative sums (sort of signed keys)."""
    rng = np.random.default_rng(5)
    P, J = 6_000, 300
    nctr = rng.integers(0, 3, P)
    cb = np.zeros(P + 1, np.uint32)
    cb[1:] = np.cumsum(nctr)
    C = int(cb[-1])
    cols = dict(pod_node=np.zeros(P, np.uint32), pod_phase=rng.choice(np.array([2, 3, 4, 1], np.uint8), P),
                pod_ctr_begin=cb, ctr_val=rng.integers(-2, 3, C).astype(np.int64), ctr_scale=np.zeros(C, np.int8),
                node_cap_val=np.array([8], np.int64), node_cap_scale=np.zeros(1, np.int8),
                pod_job=rng.integers(0, J + 3, P).astype(np.uint32), pod_role=rng.integers(0, 2, P).astype(np.uint8),
                pod_created=rng.integers(-5, 5, P).astype(np.int64))
    host = gpu.HostSnapshot.from_arrays(cols, n_jobs=J)
    _check_job(gpu, host)


def test_top_job_large_sort(gpu):
    """More jobs than one shared-memory sort tile: exercises the global merge steps."""
    host = gpu.HostSnapshot.synth(400_000, 1_000, seed=8, n_jobs=70_000)
    _check_job(gpu, host)


def test_top_job_without_any_job(gpu):
    """Pods but no job (what an empty job range of a multi-GPU cut looks like): zero totals, no fault."""
    host = gpu.HostSnapshot.synth(5_000, 50, seed=9)
    dev = gpu.DeviceSnapshot.upload(host)
    res = gpu.top_job(dev)
    assert res.totals.tolist() == [0, 0] and len(res.job_req) == 0
    dev.close()


def test_top_job_radix_ordering_on_wide_and_sparse_keys(gpu):
    """From the second query on, an uploaded snapshot's jobs are radix-sorted on the key bits that vary.
    Huge, negative and bit-sparse `requested` values make many / scattered bits vary (several passes,
    non-contiguous digit masks); ties must stay in ascending id order."""
    rng = np.random.default_rng(21)
    P, J = 150_000, 9_000
    for values in (rng.integers(-2**62, 2**62, P, dtype=np.int64),                       # ~all 64 bits vary
                   (rng.integers(0, 4, P) << 40) | (rng.integers(0, 2, P) << 3),        # two separate groups of bits
                   np.full(P, 2, np.int64),                                              # few distinct sums
                   np.zeros(P, np.int64)):                                               # nothing varies: zero passes
        cb = np.arange(P + 1, dtype=np.uint32)                                           # one container per pod
        cols = dict(pod_node=np.zeros(P, np.uint32), pod_phase=rng.choice(np.array([2, 3, 1], np.uint8), P),
                    pod_ctr_begin=cb, ctr_val=values.astype(np.int64), ctr_scale=np.zeros(P, np.int8),
                    node_cap_val=np.array([8], np.int64), node_cap_scale=np.zeros(1, np.int8),
                    pod_job=rng.integers(0, J, P).astype(np.uint32), pod_role=np.zeros(P, np.uint8),
                    pod_created=np.zeros(P, np.int64))
        host = gpu.HostSnapshot.from_arrays(cols, n_jobs=J)
        dev = gpu.DeviceSnapshot.upload(host)
        req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
        want = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
        for call in range(3):                      # bitonic first, radix afterwards
            res = gpu.top_job(dev)
            assert np.array_equal(res.job_req, req) and np.array_equal(res.job_alloc, alloc), call
            assert np.array_equal(res.job_order, want), call
        dev.close()
