"""GPU parity: the CUDA path, called through the C ABI, against the plain-C
oracle (oracle/acct_ref.c) on the same seeded inputs.  Bit-exact (integers)."""

import numpy as np
import pytest

from oracle import cref

pytestmark = pytest.mark.gpu

NONE = 0xFFFFFFFF


def _check(gpu, host, details=True):
    from arena_b200 import DeviceSnapshot, top_node
    dev = DeviceSnapshot.upload(host)
    res = top_node(dev, details=details)
    tot, alloc, cl, pg = cref.top_node_columnar(host.cols, details)
    assert np.array_equal(res.node_total, tot)
    assert np.array_equal(res.node_alloc, alloc)
    assert np.array_equal(res.cluster, cl)
    if details:
        assert np.array_equal(res.pod_gpu, pg)
    # a second query on the same handle (scratch tickets must have been reset)
    res2 = top_node(dev, details=False)
    assert np.array_equal(res2.node_alloc, alloc) and np.array_equal(res2.cluster, cl)
    dev.close()
    return res


@pytest.mark.parametrize("P,N,seed", [(200, 16, 1), (1, 1, 2), (127, 3, 3), (128, 3, 4), (129, 7, 5),
                                      (10_000, 1_000, 6), (100_003, 977, 7), (1_000_000, 100_000, 1)])
def test_uniform_grouped(gpu, P, N, seed):
    host = gpu.HostSnapshot.synth(P, N, seed=seed)
    _check(gpu, host)


def test_ungrouped_list_order(gpu):
    """Pods in arbitrary list order: still exact (only slower)."""
    host = gpu.HostSnapshot.synth(300_000, 5_000, seed=11, grouped=False)
    _check(gpu, host)


def test_skewed_gpu_sharing(gpu):
    host = gpu.HostSnapshot.synth(2_000_000, 20_000, seed=5, skewed=True)
    res = _check(gpu, host)
    assert (res.node_alloc > res.node_total).any()      # over-subscribed "GPU sharing" nodes


def test_empty_inputs(gpu):
    z = lambda dt, n=0: np.zeros(n, dt)
    cols = dict(pod_node=z(np.uint32), pod_phase=z(np.uint8), pod_ctr_begin=z(np.uint32, 1), ctr_val=z(np.int64),
                ctr_scale=z(np.int8), node_cap_val=np.array([8, 4], np.int64), node_cap_scale=z(np.int8, 2))
    host = gpu.HostSnapshot.from_arrays(cols)
    res = _check(gpu, host, details=False)
    assert res.cluster.tolist() == [12, 0]
    cols["node_cap_val"] = z(np.int64)
    cols["node_cap_scale"] = z(np.int8)
    host = gpu.HostSnapshot.from_arrays(cols)
    res = _check(gpu, host, details=False)
    assert res.cluster.tolist() == [0, 0]


def _ragged(rng, P, N, max_ctr, big=False):
    nctr = rng.integers(0, max_ctr + 1, P)
    cb = np.zeros(P + 1, np.uint32)
    cb[1:] = np.cumsum(nctr)
    C = int(cb[-1])
    if big:
        val = rng.integers(-2**63, 2**63 - 1, C, dtype=np.int64)
        scale = rng.integers(-18, 19, C).astype(np.int8)
    else:
        val = rng.integers(-5, 2000, C).astype(np.int64)
        scale = rng.choice(np.array([0, 0, 0, -3, -1, 1, 3], np.int8), C)
    node = np.sort(rng.integers(0, N + 2, P).astype(np.uint32))
    node[node >= N] = NONE           # ids N, N+1 -> no such node; sorted => they are last
    return dict(pod_node=node, pod_phase=rng.choice(np.array([0, 1, 2, 3, 4, 5, 255], np.uint8), P),
                pod_ctr_begin=cb, ctr_val=val, ctr_scale=scale,
                node_cap_val=rng.integers(-3, 17, N).astype(np.int64),
                node_cap_scale=rng.choice(np.array([0, 0, -3, 1], np.int8), N))


@pytest.mark.parametrize("seed,P,N,max_ctr,big", [(1, 5_000, 40, 6, False), (2, 999, 5, 0, False),
                                                  (3, 2_000, 11, 700, False), (4, 50_000, 300, 9, True),
                                                  (5, 130, 2, 3000, True)])
def test_ragged_and_extreme(gpu, seed, P, N, max_ctr, big):
    """Pods with zero / hundreds / thousands of containers (multi-chunk path),
    negative and fractional quantities, every scale in [-18, 18], wrapping sums."""
    rng = np.random.default_rng(seed)
    host = gpu.HostSnapshot.from_arrays(_ragged(rng, P, N, max_ctr, big))
    _check(gpu, host)


def test_pod_on_unknown_node_ids(gpu):
    """Node ids >= N that are not ARENA_NONE are ignored like unknown node names."""
    rng = np.random.default_rng(9)
    cols = _ragged(rng, 4_000, 10, 4)
    cols["pod_node"][::7] = 12345        # also breaks the grouping
    host = gpu.HostSnapshot.from_arrays(cols)
    _check(gpu, host)


def test_one_shot_host_entry(gpu):
    """arena_top_node_host: HOST columns in, HOST table out (the e2e call)."""
    host = gpu.HostSnapshot.synth(200_000, 2_000, seed=21)
    res = gpu.top_node(host, details=True)
    tot, alloc, cl, pg = cref.top_node_columnar(host.cols, True)
    assert np.array_equal(res.node_total, tot) and np.array_equal(res.node_alloc, alloc)
    assert np.array_equal(res.cluster, cl) and np.array_equal(res.pod_gpu, pg)


def test_device_entry_with_torch_tensors(gpu):
    """arena_snapshot_wrap_device + arena_top_node_device on torch's stream."""
    import torch
    host = gpu.HostSnapshot.synth(500_000, 10_000, seed=31)
    t = {k: torch.from_numpy(v.copy()).cuda() for k, v in host.cols.items() if v.size and not k.startswith("pod_job")
         and k not in ("pod_role", "pod_created")}
    dev = gpu.DeviceSnapshot.wrap_tensors(t)
    d_total = torch.empty(host.n_nodes, dtype=torch.int64, device="cuda")
    d_alloc = torch.empty_like(d_total)
    d_cluster = torch.empty(2, dtype=torch.int64, device="cuda")
    for _ in range(3):
        gpu.top_node_device(dev, d_total, d_alloc, d_cluster, None, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    tot, alloc, cl, _ = cref.top_node_columnar(host.cols)
    assert np.array_equal(d_total.cpu().numpy(), tot)
    assert np.array_equal(d_alloc.cpu().numpy(), alloc)
    assert np.array_equal(d_cluster.cpu().numpy(), cl)


def test_full_size_properties(gpu):
    """BASELINE config 4 size (1e8 pods / 1e6 nodes, one GPU): exact against
    the C oracle's single pass, plus size-independent properties."""
    host = gpu.HostSnapshot.synth(100_000_000, 1_000_000, seed=1)
    dev = gpu.DeviceSnapshot.upload(host)
    res = gpu.top_node(dev)
    # checksum of checksums: cluster totals are the sums of the table
    assert int(res.node_total.sum()) == int(res.cluster[0])
    assert int(res.node_alloc.sum()) == int(res.cluster[1])
    # idempotence
    res2 = gpu.top_node(dev)
    assert np.array_equal(res.node_alloc, res2.node_alloc) and np.array_equal(res.cluster, res2.cluster)
    tot, alloc, cl = cref.top_node_columnar_threads(host.cols, 16)
    assert np.array_equal(res.node_total, tot)
    assert np.array_equal(res.node_alloc, alloc)
    assert np.array_equal(res.cluster, cl)
    # linearity: shards over disjoint node ranges reproduce the table
    parts = []
    for i in range(4):
        sh = host.shard(i, 4)
        r = gpu.top_node(gpu.DeviceSnapshot.upload(sh))
        parts.append(r)
    assert np.array_equal(np.concatenate([p.node_alloc for p in parts]), alloc)
    assert sum(int(p.cluster[1]) for p in parts) == int(cl[1])
    dev.close()


def test_full_size_skewed(gpu):
    """BASELINE configs[4] at full size: 1e8 pods / 1e6 nodes, power-law pods per node (a few nodes
    at the 1e6-pod cap, a median of 8) and GPU-sharing nodes, seed 5.  The wide kernel, the packed
    kernel (giant nodes go through its pieces path) and four node-range shards, one after the
    other, each against the C oracle's all-threads columnar pass."""
    host = gpu.HostSnapshot.synth(100_000_000, 1_000_000, seed=5, skewed=True)
    tot, alloc, cl = cref.top_node_columnar_threads(host.cols, 16)
    pods_per_node = np.bincount(host.cols["pod_node"][host.cols["pod_node"] != NONE], minlength=host.n_nodes)
    assert pods_per_node.max() >= 1_000_000 and np.median(pods_per_node) < 20, "not the skewed workload"
    assert (alloc > tot).any(), "no over-subscribed GPU-sharing node"
    del pods_per_node
    dev = gpu.DeviceSnapshot.upload(host)
    res = gpu.top_node(dev)
    assert np.array_equal(res.node_total, tot) and np.array_equal(res.node_alloc, alloc) and np.array_equal(res.cluster, cl)
    dev.close()
    packed = host.pack(pinned=False)
    assert int(packed.view.n_pieces) > 0, "no node above the packed kernel's piece size"
    res = gpu.top_node_packed(packed)
    assert np.array_equal(res.node_total, tot) and np.array_equal(res.node_alloc, alloc) and np.array_equal(res.cluster, cl)
    packed.close()
    key_cut, pod_cut = host.shard_plan(4)
    got_total, got_alloc, got_cl = [], [], np.zeros(2, np.int64)
    for i in range(4):
        sh = host.shard(i, 4)
        d = gpu.DeviceSnapshot.upload(sh)
        r = gpu.top_node(d)
        d.close()
        # this shard's slice of the table against the oracle's rows of the same node range
        lo, hi = int(key_cut[i]), int(key_cut[i + 1])
        assert np.array_equal(r.node_alloc, alloc[lo:hi]) and np.array_equal(r.node_total, tot[lo:hi]), f"shard {i}"
        got_total.append(r.node_total); got_alloc.append(r.node_alloc); got_cl += r.cluster
    assert np.array_equal(np.concatenate(got_alloc), alloc) and np.array_equal(np.concatenate(got_total), tot)
    assert got_cl.tolist() == cl.tolist()
    assert int(np.diff(pod_cut.astype(np.int64)).sum()) == host.n_pods


def test_quantity_values_kernel(gpu):
    rng = np.random.default_rng(3)
    val = np.concatenate([rng.integers(-2**63, 2**63 - 1, 20000, dtype=np.int64),
                          rng.integers(-3000, 3000, 20000).astype(np.int64),
                          np.array([0, 1, -1, 2**63 - 1, -2**63, 500, -500, 1500, 10, 922337203685477580,
                                    922337203685477581], np.int64)])
    scale = rng.integers(-20, 21, len(val)).astype(np.int8)
    got = gpu.quantity_values(val, scale)
    want = np.array([cref.quantity_value(int(v), int(s)) for v, s in zip(val, scale)], np.int64)
    assert np.array_equal(got, want)


def test_corrupt_csr_offsets_do_not_fault(gpu):
    """Garbage in pod_ctr_begin (non-monotonic offsets) must not crash the process:
    the fast path hands such tiles to the range-checked generic path.  The sums are
    then meaningless; only memory safety and a clean return are checked."""
    rng = np.random.default_rng(12)
    host = gpu.HostSnapshot.synth(50_000, 500, seed=12)
    cols = {k: v.copy() for k, v in host.cols.items() if v.size}
    cb = cols["pod_ctr_begin"]
    idx = rng.integers(1, len(cb) - 1, 2_000)
    cb[idx] = rng.integers(0, cb[-1] + 1, len(idx)).astype(np.uint32)      # keep them inside [0, C]
    bad = gpu.HostSnapshot.from_arrays({k: v for k, v in cols.items() if not k.startswith("pod_job")
                                        and k not in ("pod_role", "pod_created")})
    dev = gpu.DeviceSnapshot.upload(bad)
    res = gpu.top_node(dev, details=True)
    assert res.node_alloc.shape == (500,)
    # the library is still healthy afterwards
    good = gpu.DeviceSnapshot.upload(host)
    r2 = gpu.top_node(good)
    tot, alloc, cl, _ = cref.top_node_columnar(host.cols)
    assert np.array_equal(r2.node_alloc, alloc)


def test_fused_totals_exchange_with_itself(gpu):
    """World of one: the kernel posts its totals to its own mailbox and reads them back
    (the multi-rank form needs more GPUs than the test box has; bench.py runs it)."""
    import torch
    from arena_b200 import multigpu
    assert multigpu.setup_peer_exchange()
    try:
        host = gpu.HostSnapshot.synth(300_000, 4_000, seed=21)
        dev = gpu.DeviceSnapshot.upload(host)
        tot, alloc, cl, _ = cref.top_node_columnar(host.cols)
        d_total = torch.empty(host.n_nodes, dtype=torch.int64, device="cuda")
        d_alloc = torch.empty(host.n_nodes, dtype=torch.int64, device="cuda")
        d_cluster = torch.zeros(2, dtype=torch.int64, device="cuda")
        for _ in range(5):                                  # epochs alternate the mailbox parity
            d_cluster.zero_()
            gpu.top_node_device_fused(dev, d_total, d_alloc, d_cluster, torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            assert d_cluster.cpu().numpy().tolist() == cl.tolist()
            assert np.array_equal(d_alloc.cpu().numpy(), alloc) and np.array_equal(d_total.cpu().numpy(), tot)
        assert gpu.lib().arena_peer_exchange_timed_out() == 0
        # deferred collection: every launch delivers the totals of the step before it, the collect
        # call those of the last one.  Two snapshots alternate so that a stale total would show.
        host2 = gpu.HostSnapshot.synth(200_000, 3_000, seed=22)
        dev2 = gpu.DeviceSnapshot.upload(host2)
        tot2, alloc2, cl2, _ = cref.top_node_columnar(host2.cols)
        d_total2 = torch.empty(host2.n_nodes, dtype=torch.int64, device="cuda")
        d_alloc2 = torch.empty(host2.n_nodes, dtype=torch.int64, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        prev = cl                                           # the fused steps above posted `cl` last
        for k in range(6):
            d_cluster.fill_(-1)
            if k % 2 == 0:
                gpu.top_node_device_posted(dev2, d_total2, d_alloc2, d_cluster, st)
                mine, table, want_alloc = cl2, d_alloc2, alloc2
            else:
                gpu.top_node_device_posted(dev, d_total, d_alloc, d_cluster, st)
                mine, table, want_alloc = cl, d_alloc, alloc
            torch.cuda.synchronize()
            assert d_cluster.cpu().numpy().tolist() == prev.tolist(), f"step {k}: not the previous step's totals"
            assert np.array_equal(table.cpu().numpy(), want_alloc)
            prev = mine
        gpu.peer_exchange_collect(d_cluster, st)
        torch.cuda.synchronize()
        assert d_cluster.cpu().numpy().tolist() == prev.tolist()
        assert gpu.lib().arena_peer_exchange_timed_out() == 0
    finally:
        gpu.lib().arena_peer_exchange_destroy()


def test_unaligned_and_dirty_output_buffers(gpu):
    """node_alloc is accumulated with atomics, so the call must clear it first whatever it held, and
    only d_pod_gpu has an alignment requirement."""
    import torch
    host = gpu.HostSnapshot.synth(700_000, 9_001, seed=23)
    dev = gpu.DeviceSnapshot.upload(host)
    tot, alloc, cl, _ = cref.top_node_columnar(host.cols)
    N = host.n_nodes
    d_total = torch.empty(N, dtype=torch.int64, device="cuda")
    buf = torch.full((N + 1,), -3, dtype=torch.int64, device="cuda")       # garbage in both
    d_cluster = torch.empty(2, dtype=torch.int64, device="cuda")
    for view in (buf[:N], buf[1:]):                                          # aligned, then 8 bytes off
        view.fill_(-3)
        gpu.top_node_device(dev, d_total, view, d_cluster, None, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(view.cpu().numpy(), alloc) and d_cluster.cpu().numpy().tolist() == cl.tolist()


def test_pageable_columns_take_the_staged_upload(gpu):
    """Columns in ordinary (pageable) memory above 16 MiB go to the device through page-locked
    slots filled by several host threads; the results must not notice."""
    host = gpu.HostSnapshot.synth(6_000_000, 50_000, seed=13, pinned=False)      # pod_node alone is 24 MB
    tot, alloc, cl, _ = cref.top_node_columnar(host.cols)
    for _ in range(2):
        res = gpu.top_node(host)                                                 # arena_top_node_host
        assert np.array_equal(res.node_alloc, alloc) and np.array_equal(res.node_total, tot) and np.array_equal(res.cluster, cl)
    dev = gpu.DeviceSnapshot.upload(host)
    res = gpu.top_node(dev)
    assert np.array_equal(res.node_alloc, alloc) and np.array_equal(res.cluster, cl)
    res = gpu.top_node_packed(host.pack(pinned=False))
    assert np.array_equal(res.node_alloc, alloc) and np.array_equal(res.cluster, cl)


def test_bulk_copy_staged_variant_is_exact(gpu):
    """ARENA_STAGED=1 selects top_node_pods_staged_kernel (pod columns staged one iteration ahead by
    cp.async.bulk into an mbarrier ring; slower than the shipped kernel, kept with its evidence in
    profiles/r02_iterations.md).  The switch is read at arena_acct_init, so the check runs in a child."""
    import os
    import subprocess
    import sys
    code = (
        "import numpy as np, arena_b200 as ab\n"
        "from oracle import cref\n"
        "ab.init(0)\n"
        "for P, N, seed, kw in [(1_000_000, 10_000, 1, {}), (129, 3, 2, {}), (2_000_000, 20_000, 5, dict(skewed=True)), (100_003, 977, 7, {})]:\n"
        "    h = ab.HostSnapshot.synth(P, N, seed=seed, **kw)\n"
        "    d = ab.DeviceSnapshot.upload(h)\n"
        "    tot, alloc, cl, _ = cref.top_node_columnar(h.cols)\n"
        "    for _ in range(2):\n"
        "        r = ab.top_node(d)\n"
        "        assert np.array_equal(r.node_total, tot) and np.array_equal(r.node_alloc, alloc) and np.array_equal(r.cluster, cl), (P, N)\n"
        "print('staged ok')\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=root,
                       env=dict(os.environ, ARENA_STAGED="1", PYTHONPATH=root))
    assert p.returncode == 0 and "staged ok" in p.stdout, (p.stdout[-1000:], p.stderr[-2000:])
