"""The C ABI: the library loads and exports every symbol the two headers
declare (no compute calls: no GPU needed), and fails loudly without CUDA."""

import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions(path):
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(arena_[a-z0-9_]+)\s*\(", text))


def test_every_declared_symbol_is_exported(built):
    from arena_b200 import _lib
    declared = header_functions(os.path.join(ROOT, "include", "arena_gpuacct.h")) | \
        header_functions(os.path.join(ROOT, "include", "arena_top.h"))
    assert len(declared) >= 30
    L = C.CDLL(_lib.LIB_PATH)
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, f"not exported: {missing}"
    # and the Python binding table covers the headers exactly
    assert set(_lib.DECLARED_SYMBOLS) == declared


def test_no_torch_types_and_no_oracle_in_the_product():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "arena_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, f"{f} mentions the oracle: the product must not route through it"
    for h in ("arena_gpuacct.h", "arena_top.h"):
        src = open(os.path.join(ROOT, "include", h)).read()
        assert "torch" not in src.replace("torchrun", "").replace("torch tensors", "")


def test_calls_fail_loudly_without_init_or_gpu(built):
    """No CPU fallback: without arena_acct_init (or without a GPU) compute entry
    points return an error instead of computing on the host."""
    import torch
    from arena_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the -m gpu tests")
    L = _lib.lib()
    rc = L.arena_acct_init(0)
    assert rc != 0
    assert L.arena_last_error()
    cluster = (C.c_int64 * 2)()
    assert L.arena_top_node(None, None, None, cluster, None) == -3      # ARENA_E_NOT_INIT
    v = _lib.SnapshotView()
    assert L.arena_top_node_host(C.byref(v), None, None, cluster, None) == -3


def test_newer_entry_points_also_fail_loudly_without_a_gpu(built):
    """The GPU loader, the packed one-shot call, pinned buffers and the peer exchange have no
    host stand-in either: without an initialised device they return ARENA_E_NOT_INIT."""
    import torch
    from arena_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the -m gpu tests")
    L = _lib.lib()
    cluster = (C.c_int64 * 2)()
    nodes, dev, buf = C.c_void_p(), C.c_void_p(), C.c_void_p()
    doc = b'{"items": []}'
    assert L.arena_load_json_device(doc, len(doc), doc, len(doc), C.byref(nodes), C.byref(dev), None) == -3
    assert nodes.value is None and dev.value is None
    assert L.arena_pinned_alloc(64, C.byref(buf)) == -3 and buf.value is None
    pv = _lib.PackedView()
    assert L.arena_top_node_host_packed(C.byref(pv), None, None, cluster, None) == -3
    handle = (C.c_ubyte * 64)()
    assert L.arena_peer_exchange_create(0, 1, handle) == -3
    assert L.arena_peer_exchange_connect(handle) == -3
    assert L.arena_top_node_device_fused(None, None, None, cluster, None) == -3
    assert L.arena_peer_exchange_timed_out() == 0
    L.arena_peer_exchange_destroy()                      # harmless without a mailbox
    assert b"arena_acct_init" in L.arena_last_error()


def test_round2_entry_points_fail_loudly_without_a_gpu(built):
    """Multi-device, packed-resident, deferred-exchange entry points: no host stand-in either.  The
    host-only ones (regroup, age ordering, the list printer) work without a device."""
    import numpy as np
    import torch
    from arena_b200 import HostSnapshot, _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the -m gpu tests")
    L = _lib.lib()
    cluster = (C.c_int64 * 2)()
    ids = (C.c_int * 2)(0, 1)
    assert L.arena_acct_init_devices(2, ids) != 0 and L.arena_acct_device_count() == 0
    assert L.arena_acct_init_devices(2, (C.c_int * 2)(0, 0)) == -1           # a device listed twice
    h = HostSnapshot.synth(1000, 10, seed=1, n_jobs=5)
    ms = C.c_void_p()
    assert L.arena_multi_snapshot_upload(C.byref(h.view), 0, C.byref(ms)) == -3 and ms.value is None
    assert L.arena_top_node_multi(None, None, None, cluster, None) == -1
    assert L.arena_top_job_multi(None, None, None, None, None, cluster) == -1
    pd = C.c_void_p()
    p = h.pack()
    assert L.arena_snapshot_upload_packed(C.byref(p.view), C.byref(pd)) == -3 and pd.value is None
    assert L.arena_top_node_packed_device(None, None, None, cluster, None) == -3
    assert L.arena_top_node_device_posted(None, None, None, cluster, None) == -3
    assert L.arena_peer_exchange_collect(cluster, None) == -3
    # host-only helpers
    byjob = h.regroup(by_job=True)
    pj = byjob.cols["pod_job"].astype(np.int64)
    pj[pj == 0xFFFFFFFF] = 5
    assert np.all(pj[:-1] <= pj[1:]) and byjob.n_pods == 1000
