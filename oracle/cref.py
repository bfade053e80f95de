"""ctypes access to oracle/_build/liboracle.so (the plain-C restatement in
oracle/acct_ref.c).  ORACLE = test infrastructure: only tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs may import this."""

from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "acct_ref.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.oracle_quantity_value.restype = C.c_int64
        _lib.oracle_quantity_value.argtypes = [C.c_int64, C.c_int32]
        _lib.oracle_aos_from_columnar.restype = C.c_void_p
        _lib.oracle_threads_top_node_columnar.restype = None
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def quantity_value(value: int, scale: int) -> int:
    return lib().oracle_quantity_value(C.c_int64(value), C.c_int32(scale))


def top_node_columnar(snap, want_pod_gpu: bool = False):
    """snap: dict of numpy columns (pod_node u32, pod_phase u8, pod_ctr_begin
    u32[P+1], ctr_val i64, ctr_scale i8, node_cap_val i64, node_cap_scale i8)."""
    P = len(snap["pod_node"])
    N = len(snap["node_cap_val"])
    total = np.zeros(N, np.int64)
    alloc = np.zeros(N, np.int64)
    cluster = np.zeros(2, np.int64)
    pod_gpu = np.zeros(P, np.int64) if want_pod_gpu else None
    lib().oracle_top_node_columnar(
        C.c_uint64(P), C.c_uint64(N),
        _p(snap["pod_node"], C.c_uint32), _p(snap["pod_phase"], C.c_uint8),
        _p(snap["pod_ctr_begin"], C.c_uint32), _p(snap["ctr_val"], C.c_int64),
        _p(snap["ctr_scale"], C.c_int8), _p(snap["node_cap_val"], C.c_int64),
        _p(snap["node_cap_scale"], C.c_int8), _p(total, C.c_int64), _p(alloc, C.c_int64),
        _p(cluster, C.c_int64), _p(pod_gpu, C.c_int64) if want_pod_gpu else None)
    return total, alloc, cluster, pod_gpu


def top_node_columnar_threads(snap, threads: int):
    """Same result as top_node_columnar, nodes split into `threads` contiguous
    ranges summed by one pthread each (requires GROUPED pods).  Used as the
    all-cores CPU baseline."""
    P = len(snap["pod_node"])
    N = len(snap["node_cap_val"])
    total = np.zeros(N, np.int64)
    alloc = np.zeros(N, np.int64)
    cluster = np.zeros(2, np.int64)
    lib().oracle_threads_top_node_columnar(
        C.c_int(threads), C.c_uint64(P), C.c_uint64(N),
        _p(snap["pod_node"], C.c_uint32), _p(snap["pod_phase"], C.c_uint8),
        _p(snap["pod_ctr_begin"], C.c_uint32), _p(snap["ctr_val"], C.c_int64),
        _p(snap["ctr_scale"], C.c_int8), _p(snap["node_cap_val"], C.c_int64),
        _p(snap["node_cap_scale"], C.c_int8), _p(total, C.c_int64), _p(alloc, C.c_int64),
        _p(cluster, C.c_int64))
    return total, alloc, cluster


def top_job_columnar(snap, J: int):
    P = len(snap["pod_job"])
    req = np.zeros(J, np.int64)
    alloc = np.zeros(J, np.int64)
    latest = np.zeros(J, np.uint32)
    totals = np.zeros(2, np.int64)
    lib().oracle_top_job_columnar(
        C.c_uint64(P), C.c_uint64(J),
        _p(snap["pod_job"], C.c_uint32), _p(snap["pod_role"], C.c_uint8),
        _p(snap["pod_created"], C.c_int64), _p(snap["pod_phase"], C.c_uint8),
        _p(snap["pod_ctr_begin"], C.c_uint32), _p(snap["ctr_val"], C.c_int64),
        _p(snap["ctr_scale"], C.c_int8), _p(req, C.c_int64), _p(alloc, C.c_int64),
        _p(latest, C.c_uint32), _p(totals, C.c_int64))
    return req, alloc, latest, totals


class AoS:
    """Array-of-structs form of a columnar snapshot (string node names, keyed
    limits): what the reference walks.  nested=True is the faithful O(N*P)
    loop of top_node.go:77-97; nested=False joins through a string hash."""

    def __init__(self, snap):
        self.P = len(snap["pod_node"])
        self.N = len(snap["node_cap_val"])
        self._keep = snap
        self._h = C.c_void_p(lib().oracle_aos_from_columnar(
            C.c_uint64(self.P), C.c_uint64(self.N),
            _p(snap["pod_node"], C.c_uint32), _p(snap["pod_phase"], C.c_uint8),
            _p(snap["pod_ctr_begin"], C.c_uint32), _p(snap["ctr_val"], C.c_int64),
            _p(snap["ctr_scale"], C.c_int8), _p(snap["node_cap_val"], C.c_int64),
            _p(snap["node_cap_scale"], C.c_int8)))

    def run(self, nested: bool, threads: int = 1):
        total = np.zeros(self.N, np.int64)
        alloc = np.zeros(self.N, np.int64)
        cluster = np.zeros(2, np.int64)
        if threads > 1 and not nested:
            lib().oracle_aos_run_threads(self._h, C.c_int(threads), _p(total, C.c_int64),
                                         _p(alloc, C.c_int64), _p(cluster, C.c_int64))
        else:
            lib().oracle_aos_run(self._h, C.c_int(1 if nested else 0), _p(total, C.c_int64),
                                 _p(alloc, C.c_int64), _p(cluster, C.c_int64))
        return total, alloc, cluster

    def close(self):
        if self._h:
            lib().oracle_aos_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
