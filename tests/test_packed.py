"""Packed host snapshots (arena_host_snapshot_pack / arena_top_node_host_packed): one byte per
pod, one per container, one row range per node.  CPU part: the packing itself against numpy;
GPU part: results bit-identical to the wide path and to the C oracle."""

import numpy as np
import pytest

from oracle import cref


def unpack(p):
    """PackedSnapshot -> the wide columns it stands for (numpy), for comparison."""
    import ctypes as C
    P, Cn, N = p.n_pods, p.n_containers, p.n_nodes

    def arr(ptr, n, dt):
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(n * np.dtype(dt).itemsize,)).view(dt).copy()

    begin = arr(p.view.node_pod_begin, N + 1, np.uint32)
    meta = arr(p.view.pod_meta, P, np.uint8) if P else np.zeros(0, np.uint8)
    if p.view.ctr_bits == 4:
        half = arr(p.view.ctr_gpu, (Cn + 1) // 2, np.uint8) if Cn else np.zeros(0, np.uint8)
        gpu = np.stack([half & 15, half >> 4], axis=1).reshape(-1)[:Cn]
    else:
        gpu = arr(p.view.ctr_gpu, Cn, np.uint8) if Cn else np.zeros(0, np.uint8)
    return begin, meta, gpu


@pytest.mark.parametrize("skewed,P,N", [(False, 30_000, 700), (True, 30_000, 700), (False, 1_500_000, 12_345),
                                        (True, 1_500_000, 12_345)])
def test_pack_is_the_wide_snapshot_in_fewer_bytes(built, skewed, P, N):
    """(From 2^20 rows + containers up the packing runs on every host thread, one row range each.)"""
    from arena_b200 import HostSnapshot
    h = HostSnapshot.synth(P, N, seed=11, skewed=skewed)
    p = h.pack()
    begin, meta, gpu = unpack(p)
    pn = h.cols["pod_node"].astype(np.int64)
    pn[pn == 0xFFFFFFFF] = h.n_nodes
    assert np.array_equal(begin, np.searchsorted(pn, np.arange(h.n_nodes + 1), side="left").astype(np.uint32))
    ph = h.cols["pod_phase"].copy()
    ph[ph == 255] = 15
    assert np.array_equal(meta & 15, ph)
    assert np.array_equal(meta >> 4, np.diff(h.cols["pod_ctr_begin"]).astype(np.uint8))
    want = np.array([cref.quantity_value(int(v), int(s)) for v, s in zip(h.cols["ctr_val"][:2000], h.cols["ctr_scale"][:2000])])
    assert np.array_equal(gpu[:2000].astype(np.int64), want)
    # every container, vectorised: the generator only writes scale 0 and scale -3 ("500m": rounded up)
    v, sc = h.cols["ctr_val"], h.cols["ctr_scale"]
    assert set(np.unique(sc).tolist()) <= {0, -3}
    assert np.array_equal(gpu.astype(np.int64), np.where(sc == 0, v, -((-v) // 1000)))
    assert p.nbytes < h.algorithmic_bytes_top_node() / 5


def test_snapshots_without_a_packed_form_say_why(built):
    from arena_b200 import ArenaError, HostSnapshot

    def snap(pod_node, phases, counts, vals, scales, n_nodes=3):
        cb = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
        return HostSnapshot.from_arrays(dict(
            pod_node=np.array(pod_node, np.uint32), pod_phase=np.array(phases, np.uint8), pod_ctr_begin=cb,
            ctr_val=np.array(vals, np.int64), ctr_scale=np.array(scales, np.int8),
            node_cap_val=np.full(n_nodes, 8, np.int64), node_cap_scale=np.zeros(n_nodes, np.int8)))

    ok = snap([0, 0, 2, 0xFFFFFFFF], [2, 3, 255, 0], [1, 0, 2, 1], [4, 1500, 255, 1], [0, -3, 0, 0]).pack()
    begin, meta, gpu = unpack(ok)
    assert begin.tolist() == [0, 2, 2, 3] and gpu.tolist() == [4, 2, 255, 1]       # 1500m rounds up to 2
    assert ok.view.ctr_bits == 8
    small = snap([0, 1, 1], [2, 2, 2], [2, 0, 3], [15, 0, 7, 8, 1], [0, 0, 0, 0, 0]).pack()
    assert small.view.ctr_bits == 4 and unpack(small)[2].tolist() == [15, 0, 7, 8, 1] and small.nbytes == 3 + 3 + 32 + 27
    assert (meta & 15).tolist() == [2, 3, 15, 0] and (meta >> 4).tolist() == [1, 0, 2, 1]
    for bad, why in ((snap([1, 0], [2, 2], [0, 0], [], []), "not grouped"),
                     (snap([0xFFFFFFFF, 0], [2, 2], [0, 0], [], []), "not grouped"),
                     (snap([0], [2], [16], [1] * 16, [0] * 16), "16 containers"),
                     (snap([0], [2], [1], [256], [0]), "256 GPUs"),
                     (snap([0], [2], [1], [-1], [0]), "-1 GPUs"),
                     (snap([0], [2], [1], [3], [2]), "300 GPUs"),
                     (snap([0], [9], [0], [], []), "phase code 9"),
                     (snap([3], [2], [0], [], []), "names node 3")):
        with pytest.raises(ArenaError) as e:
            bad.pack()
        assert e.value.code == -1 and why in str(e.value), str(e.value)
    empty = snap([], [], [], [], [], n_nodes=2).pack()
    assert unpack(empty)[0].tolist() == [0, 0, 0]


@pytest.mark.gpu
@pytest.mark.parametrize("P,N,skewed", [(0, 4, False), (1, 1, False), (3, 1, False), (1000, 7, False), (4097, 33, True),
                                        (200_000, 3000, False), (1_000_003, 50_000, True)])
def test_packed_path_is_bit_identical_to_the_wide_path(gpu, P, N, skewed):
    h = gpu.HostSnapshot.synth(P, N, seed=P + 7, skewed=skewed)
    packed = h.pack(pinned=True)
    got = gpu.top_node_packed(packed, details=True)
    tot, alloc, cl, pg = cref.top_node_columnar(h.cols, True)
    assert np.array_equal(got.node_total, tot) and np.array_equal(got.node_alloc, alloc)
    assert np.array_equal(got.cluster, cl) and np.array_equal(got.pod_gpu, pg)
    wide = gpu.top_node(h, details=True)                      # arena_top_node_host on the wide columns
    assert np.array_equal(wide.node_alloc, got.node_alloc) and np.array_equal(wide.pod_gpu, got.pod_gpu)
    for _ in range(2):                                        # summary: straight from the packed columns
        again = gpu.top_node_packed(packed)
        assert np.array_equal(again.node_alloc, alloc) and np.array_equal(again.cluster, cl)
        assert np.array_equal(again.node_total, tot)


@pytest.mark.gpu
def test_packed_summary_with_giant_nodes_and_byte_values(gpu):
    """Nodes above 2048 pods are cut into pieces; values above 15 keep a byte each."""
    rng = np.random.default_rng(3)
    sizes = [5000, 0, 1, 2049, 2048, 70000, 3, 0, 40]
    pod_node = np.concatenate([np.full(k, n, np.uint32) for n, k in enumerate(sizes)] + [np.full(77, 0xFFFFFFFF, np.uint32)])
    P = len(pod_node)
    counts = rng.integers(0, 5, P)
    cb = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
    for top in (15, 255):
        vals = rng.integers(0, top + 1, int(cb[-1])).astype(np.int64)
        h = gpu.HostSnapshot.from_arrays(dict(
            pod_node=pod_node, pod_phase=rng.choice([0, 1, 2, 3, 4, 5, 255], P).astype(np.uint8), pod_ctr_begin=cb,
            ctr_val=vals, ctr_scale=np.zeros(len(vals), np.int8),
            node_cap_val=rng.integers(0, 9, len(sizes)).astype(np.int64), node_cap_scale=np.zeros(len(sizes), np.int8)))
        packed = h.pack()
        assert packed.view.ctr_bits == (4 if top == 15 else 8) and packed.view.n_pieces == 3 + 2 + 35
        tot, alloc, cl, _ = cref.top_node_columnar(h.cols)
        got = gpu.top_node_packed(packed)
        assert np.array_equal(got.node_alloc, alloc) and np.array_equal(got.node_total, tot) and np.array_equal(got.cluster, cl)


@pytest.mark.gpu
def test_packed_path_on_the_json_fixture(gpu):
    """configs[0]: quantities like "1500m" or "2k" decide whether a packed form exists."""
    import os
    D = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "config1")
    rd = lambda n: open(os.path.join(D, n), "rb").read()
    h = gpu.HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"))
    tot, alloc, cl, _ = cref.top_node_columnar(h.cols)
    try:
        packed = h.pack()
    except gpu.ArenaError as e:
        assert "no packed form" in str(e)
        return
    got = gpu.top_node_packed(packed)
    assert np.array_equal(got.node_total, tot) and np.array_equal(got.node_alloc, alloc) and np.array_equal(got.cluster, cl)


@pytest.mark.gpu
def test_packed_one_shot_call_is_pipelined(gpu, monkeypatch):
    """Large snapshots (8 Mi pods per chunk; here the threshold is lowered) are cut into up to eight node-range
    chunks whose uploads, kernels and downloads overlap; page-locked and ordinary buffers, uniform and skewed
    (giant nodes straddle nothing: a node is never cut, its pieces go with its chunk)."""
    monkeypatch.setenv("ARENA_PACKED_CHUNK_PODS", "700000")
    for skewed, seed in ((False, 17), (True, 5)):
        h = gpu.HostSnapshot.synth(6_000_000, 60_000, seed=seed, skewed=skewed)
        tot, alloc, cl, _ = cref.top_node_columnar(h.cols)
        for pinned in (True, False):
            packed = h.pack(pinned=pinned)
            for _ in range(2):
                got = gpu.top_node_packed(packed)
                assert np.array_equal(got.node_total, tot) and np.array_equal(got.node_alloc, alloc) and np.array_equal(got.cluster, cl)
            packed.close()


@pytest.mark.gpu
def test_packed_snapshot_resident_on_the_device(gpu):
    """arena_snapshot_upload_packed + arena_top_node_packed_device: the summary kernel on packed columns that
    stay in HBM, against the C oracle -- generator snapshots, and pods with up to 15 containers each."""
    import torch
    rng = np.random.default_rng(8)
    P, N = 50_000, 300
    counts = rng.integers(0, 16, P)
    cb = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
    pn = np.sort(rng.integers(0, N + 1, P)).astype(np.uint32)
    pn[pn == N] = 0xFFFFFFFF                                   # the rows that name no node come last
    many = gpu.HostSnapshot.from_arrays(dict(
        pod_node=pn, pod_phase=rng.choice([0, 1, 2, 3, 4, 5, 255], P).astype(np.uint8), pod_ctr_begin=cb,
        ctr_val=rng.integers(0, 16, int(cb[-1])).astype(np.int64), ctr_scale=np.zeros(int(cb[-1]), np.int8),
        node_cap_val=rng.integers(0, 9, N).astype(np.int64), node_cap_scale=np.zeros(N, np.int8)))
    cases = [gpu.HostSnapshot.synth(300_000, 3_000, seed=31), gpu.HostSnapshot.synth(1_000_000, 9_000, seed=5, skewed=True), many,
             gpu.HostSnapshot.synth(5, 2, seed=1)]
    st = torch.cuda.current_stream().cuda_stream
    for h in cases:
        tot, alloc, cl, _ = cref.top_node_columnar(h.cols)
        dev = gpu.PackedDeviceSnapshot(h.pack())
        assert 0 < dev.launch_bytes < h.algorithmic_bytes_top_node()
        d_total = torch.full((h.n_nodes,), -1, dtype=torch.int64, device="cuda")
        d_alloc = torch.full((h.n_nodes,), -1, dtype=torch.int64, device="cuda")
        d_cluster = torch.full((2,), -1, dtype=torch.int64, device="cuda")
        for _ in range(3):
            gpu.top_node_packed_device(dev, d_total, d_alloc, d_cluster, st)
        torch.cuda.synchronize()
        assert np.array_equal(d_total.cpu().numpy(), tot) and np.array_equal(d_alloc.cpu().numpy(), alloc)
        assert d_cluster.cpu().numpy().tolist() == cl.tolist()
        dev.close()


# ---- the arithmetic of the packed kernel's one-pass lane sum -------------------------------------------
# kernels.cuh packed_lane_sum24 adds up a lane's eight pods in one pass: each pod sets a TOGGLE bit at its first
# value when its activity differs from the pod before it, a 256-entry table turns eight toggle bits into eight
# mask nibbles (running parity), and the masked words are summed with byte arithmetic.  This is that arithmetic
# restated in Python (32-bit wrap-around included) against a plain per-pod loop -- a check of the METHOD on the
# CPU; the kernel itself is checked against the oracle by the GPU tests above.
M32 = 0xFFFFFFFF


def _shl_clamp(x, n):
    return 0 if n >= 32 else (x << n) & M32


def _active_bytes(m):
    x = ((m & 0x0F0F0F0F) + 0x05050505) & M32
    y = (x & 0x1E1E1E1E) ^ 0x08080808
    return (((y + 0x7F7F7F7F) & M32) >> 7) & 0x01010101


def _parity_nibbles(b):
    m = par = 0
    for i in range(8):
        par ^= (b >> i) & 1
        if par:
            m |= 0xF << (4 * i)
    return m


_LUT = [_parity_nibbles(b) for b in range(256)]


def _lane_sum24(words, cl, ch, al, ah, start, total):
    funnel_r = lambda lo, hi, sh: ((hi << 32 | lo) >> sh) & M32
    pre_lo = (cl * 0x01010101) & M32
    pre_hi = (ch * 0x01010101 + (pre_lo >> 24) * 0x01010101) & M32
    ex_lo, ex_hi = (pre_lo - cl) & M32, (pre_hi - ch) & M32
    t_lo = al ^ ((al << 8) & M32)
    t_hi = ah ^ ((((ah << 32 | al) << 8) >> 32) & M32)
    sh_lo = ex_lo | (((t_lo ^ 0x01010101) << 5) & M32)
    sh_hi = ex_hi | (((t_hi ^ 0x01010101) << 5) & M32)
    d = _shl_clamp(ah >> 24, pre_hi >> 24)
    for w in (sh_lo, sh_hi):
        for j in range(4):
            d ^= _shl_clamp(1, (w >> (8 * j)) & 0xFF)
    par = lambda x: bin(x).count("1") & 1
    m = [_LUT[d & 0xFF], _LUT[(d >> 8) & 0xFF] ^ (M32 if par(d & 0xFF) else 0), _LUT[(d >> 16) & 0xFF] ^ (M32 if par(d & 0xFFFF) else 0)]
    w, sh = start >> 3, (start & 7) * 4
    y = 0
    for k in range(3 if total > 16 else 2):
        v = funnel_r(words[w + k], words[w + k + 1], sh) & m[k]
        y += (v & 0x0F0F0F0F) + ((v >> 4) & 0x0F0F0F0F)
    y = (y & 0x00FF00FF) + ((y >> 8) & 0x00FF00FF)
    return ((y * 0x00010001) & M32) >> 16


def test_one_pass_lane_sum_arithmetic():
    import random
    rng = random.Random(2)
    for _ in range(20000):
        n = rng.randint(0, 8)
        cnts, tot = [], 0
        for j in range(8):
            c = rng.choice([0, 1, 1, 2, 2, 3, 4, 7, 15]) if j < n else 0
            if tot + c > 24:
                c = 0
            cnts.append(c)
            tot += c
        phases = [rng.choice([0, 1, 2, 3, 4, 5, 15, 2, 2, 3]) if j < n else 0 for j in range(8)]
        start = rng.randint(0, 40)
        nib = [rng.randint(0, 15) for _ in range(start + tot + 48)]
        words = [sum(nib[8 * k + i] << (4 * i) for i in range(8)) for k in range(len(nib) // 8)]
        meta = [(cnts[j] << 4) | phases[j] for j in range(8)]
        m_lo = sum(meta[j] << (8 * j) for j in range(4))
        m_hi = sum(meta[4 + j] << (8 * j) for j in range(4))
        got = _lane_sum24(words, (m_lo >> 4) & 0x0F0F0F0F, (m_hi >> 4) & 0x0F0F0F0F, _active_bytes(m_lo), _active_bytes(m_hi), start, tot)
        want, pos = 0, start
        for j in range(8):
            v = sum(nib[pos:pos + cnts[j]])
            pos += cnts[j]
            if phases[j] not in (3, 4):
                want += v
        assert got == want, (cnts, phases, start)
