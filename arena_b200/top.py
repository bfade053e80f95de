"""`arena top node` / `arena top job` over the C ABI.

Host-side mirror of the reference's seams (paths relative to /root/reference):
  top_node()          NodeDescriber.getAllNodeInfos + caculateNodeGPU + cluster
                      totals, cmd/arena/commands/top_node.go:69-97, 232-242, 109-118
  top_job()           RequestedGPU / AllocatedGPU per TrainingJob, job_info.go:51-70,
                      ordered like makeTrainingJobOrderdByGPUCount, trainer.go:68-75
  format_top_node()   displayTopNodeSummary / Details, top_node.go:107-229
  format_top_job()    displayTrainingJobList(jobs, true), list.go:82-133
The arithmetic runs in the CUDA kernels; nothing here computes a sum.
"""

from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Union

import numpy as np

from ._lib import check, lib
from .snapshot import DeviceSnapshot, HostSnapshot


@dataclass
class TopNodeResult:
    node_total: np.ndarray          # int64[N]  gpuInNode
    node_alloc: np.ndarray          # int64[N]  sum of gpuInPod over the node's active pods
    cluster: np.ndarray             # int64[2]  {total, allocated}
    pod_gpu: Optional[np.ndarray]   # int64[P]  gpuInPod (details view) or None

    def usage_percent(self) -> int:
        """top_node.go:146-153: int64(float64(alloc)/float64(total)*100), 0 when total <= 0."""
        total, alloc = int(self.cluster[0]), int(self.cluster[1])
        if total > 0:
            return _go_int64(float(alloc) / float(total) * 100)
        return 0


def _go_int64(x: float) -> int:
    # Go float64 -> int64 conversion truncates toward zero
    return int(x)


@dataclass
class TopJobResult:
    job_req: np.ndarray       # int64[J]
    job_alloc: np.ndarray     # int64[J]
    job_order: np.ndarray     # uint32[J]  requested desc, id asc
    job_latest: np.ndarray    # uint32[J]  winning ROLE_LATEST pod row or 0xFFFFFFFF
    totals: np.ndarray        # int64[2]   {requested, allocated}


def _ptr(a: Optional[np.ndarray]):
    return a.ctypes.data if a is not None and a.size else None


def _outputs(N: int, P: int, details: bool, out):
    """Result arrays: fresh ones, or the caller's (e.g. page-locked, reused every call)."""
    if out is not None:
        assert len(out.node_total) == N and len(out.node_alloc) == N and (not details or len(out.pod_gpu) == P)
        return out.node_total, out.node_alloc, out.cluster, out.pod_gpu if details else None
    return (np.zeros(N, np.int64), np.zeros(N, np.int64), np.zeros(2, np.int64), np.zeros(P, np.int64) if details else None)


def top_node(snap: Union[DeviceSnapshot, HostSnapshot], details: bool = False, out: Optional[TopNodeResult] = None) -> TopNodeResult:
    """Per-node table + cluster totals.  A HostSnapshot goes through the
    one-shot arena_top_node_host (upload + kernels + download)."""
    N, P = snap.n_nodes, snap.n_pods
    total, alloc, cluster, pod_gpu = _outputs(N, P, details, out)
    if isinstance(snap, HostSnapshot):
        check(lib().arena_top_node_host(C.byref(snap.view), _ptr(total), _ptr(alloc), cluster.ctypes.data,
                                        pod_gpu.ctypes.data if details and P else None))
    else:
        check(lib().arena_top_node(snap.handle, _ptr(total), _ptr(alloc), cluster.ctypes.data,
                                   pod_gpu.ctypes.data if details and P else None))
    return TopNodeResult(total, alloc, cluster, pod_gpu)


def top_node_device_fused(snap: DeviceSnapshot, d_node_total, d_node_alloc, d_cluster, stream: Optional[int] = None) -> None:
    """arena_top_node_device_fused: the kernels plus the in-kernel totals exchange with the
    peer GPUs (multigpu.setup_peer_exchange first); d_cluster receives the GLOBAL totals."""
    check(lib().arena_top_node_device_fused(
        snap.handle,
        d_node_total.data_ptr() if d_node_total is not None and d_node_total.numel() else None,
        d_node_alloc.data_ptr() if d_node_alloc is not None and d_node_alloc.numel() else None,
        d_cluster.data_ptr(), _stream_arg(stream)))


def top_node_device_posted(snap: DeviceSnapshot, d_node_total, d_node_alloc, d_cluster_prev, stream: Optional[int] = None) -> None:
    """arena_top_node_device_posted: as top_node_device_fused, but the launch collects the PREVIOUS
    step's global totals (into d_cluster_prev) and only posts its own; peer_exchange_collect()
    delivers the last step's."""
    check(lib().arena_top_node_device_posted(
        snap.handle,
        d_node_total.data_ptr() if d_node_total is not None and d_node_total.numel() else None,
        d_node_alloc.data_ptr() if d_node_alloc is not None and d_node_alloc.numel() else None,
        d_cluster_prev.data_ptr(), _stream_arg(stream)))


def peer_exchange_collect(d_cluster, stream: Optional[int] = None) -> None:
    """arena_peer_exchange_collect: the global totals of the most recent posted step."""
    check(lib().arena_peer_exchange_collect(d_cluster.data_ptr(), _stream_arg(stream)))


def top_node_packed_device(snap, d_node_total, d_node_alloc, d_cluster, stream: Optional[int] = None) -> None:
    """arena_top_node_packed_device: the summary kernel on a device-resident packed snapshot; no sync."""
    check(lib().arena_top_node_packed_device(
        snap.handle,
        d_node_total.data_ptr() if d_node_total is not None and d_node_total.numel() else None,
        d_node_alloc.data_ptr() if d_node_alloc is not None and d_node_alloc.numel() else None,
        d_cluster.data_ptr(), _stream_arg(stream)))


def top_node_packed(packed, details: bool = False, out: Optional[TopNodeResult] = None) -> TopNodeResult:
    """arena_top_node_host_packed: one-shot upload of a PackedSnapshot + kernels + download."""
    N, P = packed.n_nodes, packed.n_pods
    total, alloc, cluster, pod_gpu = _outputs(N, P, details, out)
    check(lib().arena_top_node_host_packed(C.byref(packed.view), _ptr(total), _ptr(alloc), cluster.ctypes.data,
                                           pod_gpu.ctypes.data if details and P else None))
    return TopNodeResult(total, alloc, cluster, pod_gpu)


def _stream_arg(stream: Optional[int]):
    """None -> the library's own stream; 0 (torch's default stream handle) ->
    cudaStreamLegacy, because a NULL stream means "library stream" in the ABI."""
    if stream is None:
        return None
    return C.c_void_p(stream if stream else 1)


def top_node_device(snap: DeviceSnapshot, d_node_total, d_node_alloc, d_cluster, d_pod_gpu=None,
                    stream: Optional[int] = None) -> None:
    """Enqueue the kernels with torch CUDA tensors as outputs; no sync."""
    check(lib().arena_top_node_device(
        snap.handle,
        d_node_total.data_ptr() if d_node_total is not None and d_node_total.numel() else None,
        d_node_alloc.data_ptr() if d_node_alloc is not None and d_node_alloc.numel() else None,
        d_cluster.data_ptr(), d_pod_gpu.data_ptr() if d_pod_gpu is not None else None,
        _stream_arg(stream)))


def top_node_multi(snap, details: bool = False, out: Optional[TopNodeResult] = None) -> TopNodeResult:
    """arena_top_node_multi: every device of the process computes its node range; one table comes back
    (`out`: page-locked arrays make the per-device downloads asynchronous and concurrent)."""
    N, P = snap.n_nodes, snap.n_pods
    total, alloc, cluster, pod_gpu = _outputs(N, P, details, out)
    check(lib().arena_top_node_multi(snap.handle, _ptr(total), _ptr(alloc), cluster.ctypes.data,
                                     pod_gpu.ctypes.data if details and P else None))
    return TopNodeResult(total, alloc, cluster, pod_gpu)


def top_job_multi(snap) -> TopJobResult:
    """arena_top_job_multi: job-id ranges, one per device; merged ordering."""
    J = snap.n_jobs
    req, alloc = np.zeros(J, np.int64), np.zeros(J, np.int64)
    order, latest = np.zeros(J, np.uint32), np.zeros(J, np.uint32)
    totals = np.zeros(2, np.int64)
    check(lib().arena_top_job_multi(snap.handle, _ptr(req), _ptr(alloc), _ptr(order), _ptr(latest), totals.ctypes.data))
    return TopJobResult(req, alloc, order, latest, totals)


def top_job(snap: DeviceSnapshot) -> TopJobResult:
    J = snap.n_jobs
    req = np.zeros(J, np.int64)
    alloc = np.zeros(J, np.int64)
    order = np.zeros(J, np.uint32)
    latest = np.zeros(J, np.uint32)
    totals = np.zeros(2, np.int64)
    check(lib().arena_top_job(snap.handle, _ptr(req), _ptr(alloc), _ptr(order), _ptr(latest), totals.ctypes.data))
    return TopJobResult(req, alloc, order, latest, totals)


def top_job_device(snap: DeviceSnapshot, d_req, d_alloc, d_order, d_latest, d_totals,
                   stream: Optional[int] = None) -> None:
    check(lib().arena_top_job_device(
        snap.handle, d_req.data_ptr() if d_req.numel() else None, d_alloc.data_ptr() if d_alloc.numel() else None,
        d_order.data_ptr() if d_order.numel() else None,
        d_latest.data_ptr() if d_latest is not None and d_latest.numel() else None, d_totals.data_ptr(),
        _stream_arg(stream)))


def quantity_values(val: np.ndarray, scale: np.ndarray) -> np.ndarray:
    """Quantity.Value() of int64Amount pairs, on the GPU."""
    val = np.ascontiguousarray(val, np.int64)
    scale = np.ascontiguousarray(scale, np.int8)
    out = np.zeros(len(val), np.int64)
    check(lib().arena_quantity_values(_ptr(val), _ptr(scale), len(val), _ptr(out)))
    return out


def _take_text(p: C.c_void_p, n: C.c_size_t) -> str:
    try:
        return C.string_at(p.value, n.value).decode("utf-8", "replace")
    finally:
        lib().arena_free_text(p)


def format_top_node(host: HostSnapshot, res: TopNodeResult, details: bool = False) -> str:
    p, n = C.c_void_p(), C.c_size_t()
    check(lib().arena_format_top_node(host.handle, _ptr(res.node_total), _ptr(res.node_alloc),
                                      _ptr(res.pod_gpu) if details else None, int(details), C.byref(p), C.byref(n)))
    return _take_text(p, n)


def format_top_job(host: HostSnapshot, res: TopJobResult, now_unix: int) -> str:
    p, n = C.c_void_p(), C.c_size_t()
    check(lib().arena_format_top_job(host.handle, _ptr(res.job_req), _ptr(res.job_alloc), _ptr(res.job_order),
                                     _ptr(res.job_latest), now_unix, C.byref(p), C.byref(n)))
    return _take_text(p, n)


def format_job_list(host: HostSnapshot, res: TopJobResult, now_unix: int) -> str:
    """`arena list` (list.go:16-80): the jobs of `res` ordered by age (arena_job_order_by_age) and
    printed without the GPU columns (arena_format_job_list, display_gpu = 0)."""
    order = np.zeros(host.n_jobs, np.uint32)
    check(lib().arena_job_order_by_age(host.handle, _ptr(order)))
    p, n = C.c_void_p(), C.c_size_t()
    check(lib().arena_format_job_list(host.handle, None, None, _ptr(order), _ptr(res.job_latest), now_unix, 0,
                                      C.byref(p), C.byref(n)))
    return _take_text(p, n)


def top_job_device_noalloc(tensors: dict, n_jobs: int, d_req, d_alloc, d_order, d_latest, d_totals,
                           stream: Optional[int] = None, n_pods: Optional[int] = None) -> None:
    """arena_top_job_device_noalloc: `top job` over CUDA tensors holding the
    columns, for callers that must not allocate device memory on the query
    path (the library keeps the work arrays in a fixed device array instead of
    cudaMalloc'ing them).  At most ARENA_NOALLOC_MAX_JOBS (2048) jobs; calls
    share that one array, so do not overlap them on the GPU; make the first
    call on a device before starting any stream capture.  Enqueues only."""
    from ._lib import SnapshotView
    from .snapshot import _COLS
    v = SnapshotView()
    v.n_pods = tensors["pod_job"].numel() if n_pods is None else n_pods      # pod_node is optional here
    v.n_containers = tensors["ctr_val"].numel()
    v.n_nodes = tensors["node_cap_val"].numel() if "node_cap_val" in tensors else 0
    v.n_jobs = n_jobs
    for name, _, _ in _COLS:
        t = tensors.get(name)
        setattr(v, name, t.data_ptr() if t is not None and t.numel() else None)
    v.pod_ctr_begin = tensors["pod_ctr_begin"].data_ptr()
    check(lib().arena_top_job_device_noalloc(
        C.byref(v), d_req.data_ptr() if d_req.numel() else None, d_alloc.data_ptr() if d_alloc.numel() else None,
        d_order.data_ptr() if d_order.numel() else None, d_latest.data_ptr() if d_latest.numel() else None,
        d_totals.data_ptr(), _stream_arg(stream)))
