"""Host / device snapshot handles over the C ABI.

`HostSnapshot` is what the reference gets from its two List calls
(cmd/arena/commands/pod_helper.go:15-43, top_node.go:71), in columnar form;
`DeviceSnapshot` is the same data resident in HBM.
"""

from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np

from . import _lib
from ._lib import ArenaError, LoadDeviceStats, PackedView, SnapshotView, SynthParams, check, lib

_COLS = [
    # name, dtype, length as a function of the view
    ("pod_node", np.uint32, lambda v: v.n_pods),
    ("pod_phase", np.uint8, lambda v: v.n_pods),
    ("pod_ctr_begin", np.uint32, lambda v: v.n_pods + 1),
    ("ctr_val", np.int64, lambda v: v.n_containers),
    ("ctr_scale", np.int8, lambda v: v.n_containers),
    ("node_cap_val", np.int64, lambda v: v.n_nodes),
    ("node_cap_scale", np.int8, lambda v: v.n_nodes),
    ("pod_job", np.uint32, lambda v: v.n_pods),
    ("pod_role", np.uint8, lambda v: v.n_pods),
    ("pod_created", np.int64, lambda v: v.n_pods),
]


def _np_view(addr: int, dtype, n: int) -> np.ndarray:
    if n == 0 or not addr:
        return np.zeros(0, dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(addr)
    return np.frombuffer(buf, dtype=dtype, count=n)


class HostSnapshot:
    """Owns an `arena_host_snapshot*` (or keeps numpy arrays alive for a view
    built from arrays).  `.cols` maps column name -> numpy array (zero-copy)."""

    def __init__(self, handle: Optional[int], view: SnapshotView, keep=None):
        self._h = handle
        self.view = view
        self._keep = keep
        self.cols: Dict[str, np.ndarray] = {}
        for name, dt, n in _COLS:
            self.cols[name] = _np_view(getattr(view, name), dt, int(n(view)))

    # -- constructors -------------------------------------------------------
    @classmethod
    def from_handle(cls, handle: int, keep=None) -> "HostSnapshot":
        v = SnapshotView()
        check(lib().arena_host_snapshot_view(handle, C.byref(v)))
        return cls(handle, v, keep)

    @classmethod
    def from_arrays(cls, cols: Dict[str, np.ndarray], n_jobs: int = 0) -> "HostSnapshot":
        """Build a view over caller-owned numpy columns (kept alive here)."""
        keep = {}
        v = SnapshotView()
        P = len(cols["pod_node"])
        for name, dt, _ in _COLS:
            a = cols.get(name)
            if a is None:
                setattr(v, name, None)
                continue
            a = np.ascontiguousarray(a, dtype=dt)
            keep[name] = a
            setattr(v, name, a.ctypes.data if a.size else None)
        if "pod_ctr_begin" not in keep or len(keep["pod_ctr_begin"]) != P + 1:
            raise ValueError("pod_ctr_begin must have P+1 entries")
        v.n_pods = P
        v.n_containers = int(keep["pod_ctr_begin"][P])
        v.n_nodes = len(cols["node_cap_val"])
        v.n_jobs = n_jobs
        if v.n_containers != len(cols["ctr_val"]):
            raise ValueError("pod_ctr_begin[P] != len(ctr_val)")
        # pod_ctr_begin may be size 1 with P == 0: keep its address
        v.pod_ctr_begin = keep["pod_ctr_begin"].ctypes.data
        hs = cls(None, v, keep)
        return hs

    @classmethod
    def synth(cls, n_pods: int, n_nodes: int, seed: int = 1, n_jobs: int = 0, skewed: bool = False,
              grouped: bool = True, pinned: bool = False, threads: int = 0) -> "HostSnapshot":
        """Seeded synthetic snapshot (arena_synth_snapshot)."""
        sp = SynthParams(n_pods, n_nodes, n_jobs, seed, int(skewed), int(grouped), int(pinned), threads)
        h = C.c_void_p()
        check(lib().arena_synth_snapshot(C.byref(sp), C.byref(h)))
        return cls.from_handle(h.value)

    @classmethod
    def from_json(cls, podlist: bytes, nodelist: bytes, joblist: Optional[bytes] = None,
                  tfjoblist: Optional[bytes] = None, helm_list: Optional[bytes] = None,
                  group_by_job: bool = False) -> "HostSnapshot":
        """arena_load_json: PodList/NodeList(/JobList/TFJobList/`helm list`) -> columns."""
        h = C.c_void_p()
        n = lambda b: len(b) if b is not None else 0
        check(lib().arena_load_json(podlist, len(podlist), nodelist, len(nodelist), joblist, n(joblist),
                                    tfjoblist, n(tfjoblist), helm_list, n(helm_list), int(group_by_job),
                                    C.byref(h)))
        return cls.from_handle(h.value)

    @classmethod
    def from_json_gpu(cls, podlist, nodelist: bytes, joblist: Optional[bytes] = None,
                      tfjoblist: Optional[bytes] = None, helm_list: Optional[bytes] = None,
                      group_by_job: bool = False):
        """arena_load_json_gpu: the same snapshot as from_json, the PodList (bytes or a PinnedBuffer)
        decoded on the GPU.  Returns (snapshot, stats dict)."""
        h = C.c_void_p()
        st = LoadDeviceStats()
        n = lambda b: len(b) if b is not None else 0
        if isinstance(podlist, PinnedBuffer):
            ptr, plen = C.c_void_p(podlist.ptr), podlist.nbytes
        else:
            ptr, plen = podlist, len(podlist)
        check(lib().arena_load_json_gpu(ptr, plen, nodelist, len(nodelist), joblist, n(joblist), tfjoblist, n(tfjoblist),
                                        helm_list, n(helm_list), int(group_by_job), C.byref(h), C.byref(st)))
        return cls.from_handle(h.value), {k: getattr(st, k) for k, _ in LoadDeviceStats._fields_}

    @classmethod
    def from_json_hostscan(cls, podlist: bytes, nodelist: bytes, joblist: Optional[bytes] = None,
                           tfjoblist: Optional[bytes] = None, helm_list: Optional[bytes] = None,
                           group_by_job: bool = False):
        """arena_load_json_hostscan (diagnostic): the device scanner's source run on the host.
        Returns (snapshot, items handed to the host's item code)."""
        h = C.c_void_p()
        handed = C.c_uint64(0)
        n = lambda b: len(b) if b is not None else 0
        check(lib().arena_load_json_hostscan(podlist, len(podlist), nodelist, len(nodelist), joblist, n(joblist),
                                             tfjoblist, n(tfjoblist), helm_list, n(helm_list), int(group_by_job),
                                             C.byref(h), C.byref(handed)))
        return cls.from_handle(h.value), int(handed.value)

    # -- properties -----------------------------------------------------------
    @property
    def handle(self) -> Optional[int]:
        return self._h

    @property
    def n_pods(self) -> int:
        return int(self.view.n_pods)

    @property
    def n_containers(self) -> int:
        return int(self.view.n_containers)

    @property
    def n_nodes(self) -> int:
        return int(self.view.n_nodes)

    @property
    def n_jobs(self) -> int:
        return int(self.view.n_jobs)

    def algorithmic_bytes_top_node(self, pod_gpu: bool = False) -> int:
        """B = 9P + 9C + 25N + 20, +8P when pod_gpu is written."""
        return 9 * self.n_pods + 9 * self.n_containers + 25 * self.n_nodes + 20 + (8 * self.n_pods if pod_gpu else 0)

    def pod_order(self) -> Optional[np.ndarray]:
        if not self._h:
            return None
        p = lib().arena_host_pod_order(self._h)
        return _np_view(p, np.uint32, self.n_pods).copy() if p else None

    def job_name(self, j: int) -> str:
        return lib().arena_host_job_name(self._h, j).decode()

    def job_trainer(self, j: int) -> str:
        return lib().arena_host_job_trainer(self._h, j).decode()

    def node_name(self, n: int) -> str:
        return lib().arena_host_node_name(self._h, n).decode()

    def pod_name(self, row: int) -> str:
        return lib().arena_host_pod_name(self._h, row).decode()

    def pack(self, pinned: bool = False) -> "PackedSnapshot":
        """arena_host_snapshot_pack (raises ArenaError -1 when there is no packed form)."""
        return PackedSnapshot(self, pinned)

    def regroup(self, by_job: bool = False) -> "HostSnapshot":
        """arena_host_snapshot_regroup: rows grouped (stably) by node id or by job id."""
        h = C.c_void_p()
        check(lib().arena_host_snapshot_regroup(C.byref(self.view), int(by_job), C.byref(h)))
        return HostSnapshot.from_handle(h.value)

    # -- sharding ---------------------------------------------------------------
    def shard_plan(self, n_shards: int, by_job: bool = False):
        key_cut = np.zeros(n_shards + 1, np.uint64)
        pod_cut = np.zeros(n_shards + 1, np.uint64)
        check(lib().arena_shard_plan(C.byref(self.view), n_shards, int(by_job), key_cut.ctypes.data,
                                     pod_cut.ctypes.data))
        return key_cut, pod_cut

    def shard(self, i: int, n_shards: int, by_job: bool = False) -> "HostSnapshot":
        key_cut, pod_cut = self.shard_plan(n_shards, by_job)
        h = C.c_void_p()
        check(lib().arena_shard_extract(C.byref(self.view), int(by_job), int(key_cut[i]), int(key_cut[i + 1]),
                                        int(pod_cut[i]), int(pod_cut[i + 1]), C.byref(h)))
        s = HostSnapshot.from_handle(h.value, keep=self)     # aliases the parent's columns
        s.key_lo, s.key_hi = int(key_cut[i]), int(key_cut[i + 1])
        s.pod_lo, s.pod_hi = int(pod_cut[i]), int(pod_cut[i + 1])
        return s

    def scan_podlist(self, podlist: bytes, capacity: int):
        """arena_scan_podlist_host: the DEVICE loader's per-item scanner run on the
        host against this (NodeList) snapshot.  Returns numpy arrays."""
        n = C.c_uint64()
        node = np.zeros(capacity, np.uint32)
        phase = np.zeros(capacity, np.uint8)
        nctr = np.zeros(capacity, np.uint32)
        vals = np.zeros((capacity, 8), np.int64)
        scales = np.zeros((capacity, 8), np.int8)
        flag = np.zeros(capacity, np.uint8)
        check(lib().arena_scan_podlist_host(podlist, len(podlist), self._h, capacity, C.byref(n), node.ctypes.data,
                                            phase.ctypes.data, nctr.ctypes.data, vals.ctypes.data, scales.ctypes.data,
                                            flag.ctypes.data))
        k = int(n.value)
        return {"n": k, "node": node[:k], "phase": phase[:k], "n_ctrs": nctr[:k], "ctr_val": vals[:k], "ctr_scale": scales[:k],
                "flag": flag[:k]}

    def close(self) -> None:
        if self._h:
            lib().arena_host_snapshot_free(self._h)
            self._h = None
        self.cols = {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PackedSnapshot:
    """`arena_packed_snapshot*`: one byte per pod, one per container, one row range per node
    (the host side of arena_top_node_host_packed)."""

    def __init__(self, host: "HostSnapshot", pinned: bool = False):
        h = C.c_void_p()
        check(lib().arena_host_snapshot_pack(C.byref(host.view), int(pinned), C.byref(h)))
        self._h = h.value
        self.view = PackedView()
        check(lib().arena_packed_snapshot_view(self._h, C.byref(self.view)))
        self.n_pods, self.n_containers, self.n_nodes = (int(self.view.n_pods), int(self.view.n_containers),
                                                        int(self.view.n_nodes))

    @property
    def nbytes(self) -> int:
        """bytes the one-shot call copies to the device"""
        ctr = (self.n_containers + 1) // 2 if self.view.ctr_bits == 4 else self.n_containers
        return self.n_pods + ctr + 8 * (self.n_nodes + 1) + 9 * self.n_nodes + 16 * int(self.view.n_pieces)

    def close(self) -> None:
        if self._h:
            lib().arena_packed_snapshot_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PackedDeviceSnapshot:
    """`arena_packed_device*`: the packed columns resident in HBM (arena_snapshot_upload_packed)."""

    def __init__(self, packed: "PackedSnapshot"):
        h = C.c_void_p()
        check(lib().arena_snapshot_upload_packed(C.byref(packed.view), C.byref(h)))
        self._h = h.value
        self.n_pods, self.n_containers, self.n_nodes = packed.n_pods, packed.n_containers, packed.n_nodes
        b = C.c_uint64()
        check(lib().arena_packed_device_bytes(self._h, C.byref(b)))
        self.launch_bytes = int(b.value)

    @property
    def handle(self) -> int:
        return self._h

    def close(self) -> None:
        if self._h:
            lib().arena_packed_device_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PinnedBuffer:
    """arena_pinned_alloc: page-locked host bytes (read a PodList into one of these and the
    copy to the device runs at PCIe speed)."""

    def __init__(self, nbytes: int):
        p = C.c_void_p()
        check(lib().arena_pinned_alloc(int(nbytes), C.byref(p)))
        self.ptr, self.capacity, self.nbytes = p.value, int(nbytes), 0

    def fill(self, data: bytes) -> "PinnedBuffer":
        assert len(data) <= self.capacity
        C.memmove(self.ptr, data, len(data))
        self.nbytes = len(data)
        return self

    def close(self) -> None:
        if self.ptr:
            lib().arena_pinned_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MultiSnapshot:
    """`arena_multi_snapshot*`: ONE host snapshot cut into key ranges, one per device of this process
    (init_devices first).  by_job=False: node ranges for top_node_multi; by_job=True: job-id ranges
    for top_job_multi (rows grouped by job: HostSnapshot.regroup(by_job=True))."""

    def __init__(self, host: HostSnapshot, by_job: bool = False):
        h = C.c_void_p()
        check(lib().arena_multi_snapshot_upload(C.byref(host.view), int(by_job), C.byref(h)))
        self._h = h.value
        self.host, self.by_job = host, by_job
        self.n_pods, self.n_nodes, self.n_jobs = host.n_pods, host.n_nodes, host.n_jobs
        n = C.c_int()
        check(lib().arena_multi_snapshot_shards(self._h, C.byref(n), None, None))
        self.n_shards = n.value
        self.key_cut = np.zeros(self.n_shards + 1, np.uint64)
        self.pod_cut = np.zeros(self.n_shards + 1, np.uint64)
        check(lib().arena_multi_snapshot_shards(self._h, None, self.key_cut.ctypes.data, self.pod_cut.ctypes.data))

    @property
    def handle(self) -> int:
        return self._h

    def close(self) -> None:
        if self._h:
            lib().arena_multi_snapshot_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceSnapshot:
    """`arena_snapshot*`: columns resident in HBM."""

    def __init__(self, handle: int, host: Optional[HostSnapshot], keep=None):
        self._h = handle
        self.host = host
        self._keep = keep
        src = host.view if host is not None else keep["view"]
        self.n_pods, self.n_containers = int(src.n_pods), int(src.n_containers)
        self.n_nodes, self.n_jobs = int(src.n_nodes), int(src.n_jobs)

    @classmethod
    def upload(cls, host: HostSnapshot) -> "DeviceSnapshot":
        h = C.c_void_p()
        check(lib().arena_snapshot_upload(C.byref(host.view), C.byref(h)))
        return cls(h.value, host)

    @classmethod
    def wrap_tensors(cls, tensors: dict, n_jobs: int = 0) -> "DeviceSnapshot":
        """Wrap CUDA torch tensors (same names / dtypes as the host columns)
        without copying; the tensors are kept alive by this object."""
        v = SnapshotView()
        P = tensors["pod_node"].numel()
        v.n_pods = P
        v.n_containers = tensors["ctr_val"].numel()
        v.n_nodes = tensors["node_cap_val"].numel()
        v.n_jobs = n_jobs
        for name, _, _ in _COLS:
            t = tensors.get(name)
            setattr(v, name, t.data_ptr() if t is not None and t.numel() else None)
        v.pod_ctr_begin = tensors["pod_ctr_begin"].data_ptr()
        h = C.c_void_p()
        check(lib().arena_snapshot_wrap_device(C.byref(v), C.byref(h)))
        return cls(h.value, None, keep={"tensors": tensors, "view": v})

    @classmethod
    def from_json_device(cls, podlist, nodelist: bytes):
        """arena_load_json_device: the PodList (bytes or a PinnedBuffer) is parsed on the
        GPU.  Returns (DeviceSnapshot, stats dict); .host is the NodeList's host snapshot
        (zero pods unless the host loader had to take over)."""
        nodes, dev, st = C.c_void_p(), C.c_void_p(), LoadDeviceStats()
        if isinstance(podlist, PinnedBuffer):
            ptr, n = C.c_void_p(podlist.ptr), podlist.nbytes
        else:
            ptr, n = podlist, len(podlist)
        check(lib().arena_load_json_device(ptr, n, nodelist, len(nodelist), C.byref(nodes), C.byref(dev), C.byref(st)))
        host = HostSnapshot.from_handle(nodes.value)
        counts = (C.c_uint64 * 4)()
        check(lib().arena_snapshot_counts(dev, counts))
        v = SnapshotView()
        v.n_pods, v.n_containers, v.n_nodes, v.n_jobs = counts[0], counts[1], counts[2], counts[3]
        out = cls(dev.value, None, keep={"view": v})
        out.host = host
        stats = {k: getattr(st, k) for k, _ in LoadDeviceStats._fields_}
        return out, stats

    def download(self) -> dict:
        """Device columns copied back to numpy arrays (tests / inspection)."""
        P, Cn, N = self.n_pods, self.n_containers, self.n_nodes
        cols = {"pod_node": np.zeros(P, np.uint32), "pod_phase": np.zeros(P, np.uint8),
                "pod_ctr_begin": np.zeros(P + 1, np.uint32), "ctr_val": np.zeros(Cn, np.int64),
                "ctr_scale": np.zeros(Cn, np.int8), "node_cap_val": np.zeros(N, np.int64),
                "node_cap_scale": np.zeros(N, np.int8)}
        v = SnapshotView()
        for k, a in cols.items():
            setattr(v, k, a.ctypes.data)
        check(lib().arena_snapshot_download(self._h, C.byref(v)))
        return cols

    @property
    def handle(self) -> int:
        return self._h

    def close(self) -> None:
        if self._h:
            lib().arena_snapshot_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
