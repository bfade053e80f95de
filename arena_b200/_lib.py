"""ctypes binding of libarena_gpuacct.so (include/arena_gpuacct.h, arena_top.h).

There is no fallback: if the library is missing, or CUDA is, the calls raise.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# ARENA_GPUACCT_LIB: A/B-test another build of the same library (kernel tuning only)
LIB_PATH = os.environ.get("ARENA_GPUACCT_LIB") or os.path.join(_HERE, "libarena_gpuacct.so")

ARENA_NONE = 0xFFFFFFFF


class ArenaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"arena_gpuacct error {code}: {msg}")
        self.code = code


class SnapshotView(C.Structure):
    """struct arena_snapshot_view"""
    _fields_ = [
        ("n_pods", C.c_uint64), ("n_containers", C.c_uint64),
        ("n_nodes", C.c_uint64), ("n_jobs", C.c_uint64),
        ("pod_node", C.c_void_p), ("pod_phase", C.c_void_p), ("pod_ctr_begin", C.c_void_p),
        ("ctr_val", C.c_void_p), ("ctr_scale", C.c_void_p),
        ("node_cap_val", C.c_void_p), ("node_cap_scale", C.c_void_p),
        ("pod_job", C.c_void_p), ("pod_role", C.c_void_p), ("pod_created", C.c_void_p),
    ]


class PackedView(C.Structure):
    """struct arena_packed_view"""
    _fields_ = [
        ("n_pods", C.c_uint64), ("n_containers", C.c_uint64), ("n_nodes", C.c_uint64),
        ("node_pod_begin", C.c_void_p), ("pod_meta", C.c_void_p), ("ctr_gpu", C.c_void_p),
        ("node_cap_val", C.c_void_p), ("node_cap_scale", C.c_void_p), ("ctr_bits", C.c_uint32),
        ("node_ctr_begin", C.c_void_p), ("pieces", C.c_void_p), ("n_pieces", C.c_uint64),
    ]


class LoadDeviceStats(C.Structure):
    """struct arena_load_device_stats"""
    _fields_ = [
        ("n_items", C.c_uint64), ("flagged_item", C.c_uint64), ("flag_reason", C.c_uint32), ("fell_back", C.c_uint32),
        ("ms_h2d", C.c_double), ("ms_structure", C.c_double), ("ms_items", C.c_double), ("ms_device", C.c_double),
        ("n_handed_back", C.c_uint64),
    ]


class SynthParams(C.Structure):
    """struct arena_synth_params"""
    _fields_ = [
        ("n_pods", C.c_uint64), ("n_nodes", C.c_uint64), ("n_jobs", C.c_uint64), ("seed", C.c_uint64),
        ("skewed", C.c_int), ("grouped", C.c_int), ("pinned", C.c_int), ("threads", C.c_int),
    ]


_lib = None

# every symbol the two headers declare: (name, restype, argtypes)
_VP = C.c_void_p
_SIGS = [
    ("arena_acct_init", C.c_int, [C.c_int]),
    ("arena_acct_shutdown", None, []),
    ("arena_acct_device", C.c_int, []),
    ("arena_acct_sm_count", C.c_int, []),
    ("arena_last_error", C.c_char_p, []),
    ("arena_snapshot_upload", C.c_int, [C.POINTER(SnapshotView), C.POINTER(_VP)]),
    ("arena_snapshot_wrap_device", C.c_int, [C.POINTER(SnapshotView), C.POINTER(_VP)]),
    ("arena_snapshot_free", None, [_VP]),
    ("arena_snapshot_counts", C.c_int, [_VP, C.POINTER(C.c_uint64)]),
    ("arena_snapshot_download", C.c_int, [_VP, C.POINTER(SnapshotView)]),
    ("arena_top_node", C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    ("arena_top_node_device", C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    ("arena_top_node_host", C.c_int, [C.POINTER(SnapshotView), _VP, _VP, _VP, _VP]),
    ("arena_peer_exchange_create", C.c_int, [C.c_int, C.c_int, _VP]),
    ("arena_peer_exchange_connect", C.c_int, [_VP]),
    ("arena_peer_exchange_destroy", None, []),
    ("arena_peer_exchange_timed_out", C.c_int, []),
    ("arena_top_node_device_fused", C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    ("arena_top_node_device_posted", C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    ("arena_peer_exchange_collect", C.c_int, [_VP, _VP]),
    ("arena_top_node_host_packed", C.c_int, [C.POINTER(PackedView), _VP, _VP, _VP, _VP]),
    ("arena_snapshot_upload_packed", C.c_int, [C.POINTER(PackedView), C.POINTER(_VP)]),
    ("arena_packed_device_free", None, [_VP]),
    ("arena_packed_device_bytes", C.c_int, [_VP, C.POINTER(C.c_uint64)]),
    ("arena_top_node_packed_device", C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    ("arena_acct_init_devices", C.c_int, [C.c_int, C.POINTER(C.c_int)]),
    ("arena_acct_device_count", C.c_int, []),
    ("arena_multi_snapshot_upload", C.c_int, [C.POINTER(SnapshotView), C.c_int, C.POINTER(_VP)]),
    ("arena_multi_snapshot_free", None, [_VP]),
    ("arena_multi_snapshot_shards", C.c_int, [_VP, C.POINTER(C.c_int), _VP, _VP]),
    ("arena_top_node_multi", C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    ("arena_top_job_multi", C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    ("arena_top_job", C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    ("arena_top_job_device", C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    ("arena_top_job_device_noalloc", C.c_int, [C.POINTER(SnapshotView), _VP, _VP, _VP, _VP, _VP, _VP]),
    ("arena_quantity_values", C.c_int, [_VP, _VP, C.c_uint64, _VP]),
    ("arena_kernel_launches", C.c_uint64, []),
    ("arena_calibrate_stream_read", C.c_int, [_VP, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    ("arena_calibrate_stream_pattern", C.c_int, [_VP, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    ("arena_profile_begin", C.c_int, [C.c_int]),
    ("arena_profile_begin_sampled", C.c_int, [C.c_int, C.c_int]),
    ("arena_profile_end", C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    # arena_top.h
    ("arena_host_snapshot_free", None, [_VP]),
    ("arena_host_snapshot_view", C.c_int, [_VP, C.POINTER(SnapshotView)]),
    ("arena_load_json", C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t,
                                  C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_int, C.POINTER(_VP)]),
    ("arena_pinned_alloc", C.c_int, [C.c_size_t, C.POINTER(_VP)]),
    ("arena_pinned_free", None, [_VP]),
    ("arena_load_json_device", C.c_int, [_VP, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(_VP), C.POINTER(_VP),
                                         C.POINTER(LoadDeviceStats)]),
    ("arena_load_json_gpu", C.c_int, [_VP, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t,
                                      C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_int, C.POINTER(_VP),
                                      C.POINTER(LoadDeviceStats)]),
    ("arena_load_json_hostscan", C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t,
                                           C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_int, C.POINTER(_VP),
                                           C.POINTER(C.c_uint64)]),
    ("arena_scan_podlist_host", C.c_int, [C.c_char_p, C.c_size_t, _VP, C.c_uint64, C.POINTER(C.c_uint64), _VP, _VP, _VP,
                                          _VP, _VP, _VP]),
    ("arena_host_snapshot_pack", C.c_int, [C.POINTER(SnapshotView), C.c_int, C.POINTER(_VP)]),
    ("arena_packed_snapshot_view", C.c_int, [_VP, C.POINTER(PackedView)]),
    ("arena_packed_snapshot_free", None, [_VP]),
    ("arena_host_pod_order", _VP, [_VP]),
    ("arena_host_job_name", C.c_char_p, [_VP, C.c_uint32]),
    ("arena_host_job_trainer", C.c_char_p, [_VP, C.c_uint32]),
    ("arena_host_node_name", C.c_char_p, [_VP, C.c_uint32]),
    ("arena_host_pod_name", C.c_char_p, [_VP, C.c_uint32]),
    ("arena_quantity_parse", C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                       C.POINTER(C.c_int32), C.POINTER(C.c_int)]),
    ("arena_format_top_node", C.c_int, [_VP, _VP, _VP, _VP, C.c_int, C.POINTER(_VP), C.POINTER(C.c_size_t)]),
    ("arena_format_top_job", C.c_int, [_VP, _VP, _VP, _VP, _VP, C.c_int64, C.POINTER(_VP), C.POINTER(C.c_size_t)]),
    ("arena_job_order_by_age", C.c_int, [_VP, _VP]),
    ("arena_format_job_list", C.c_int, [_VP, _VP, _VP, _VP, _VP, C.c_int64, C.c_int, C.POINTER(_VP), C.POINTER(C.c_size_t)]),
    ("arena_free_text", None, [_VP]),
    ("arena_synth_snapshot", C.c_int, [C.POINTER(SynthParams), C.POINTER(_VP)]),
    ("arena_shard_plan", C.c_int, [C.POINTER(SnapshotView), C.c_int, C.c_int, _VP, _VP]),
    ("arena_host_snapshot_regroup", C.c_int, [C.POINTER(SnapshotView), C.c_int, C.POINTER(_VP)]),
    ("arena_shard_extract", C.c_int, [C.POINTER(SnapshotView), C.c_int, C.c_uint64, C.c_uint64, C.c_uint64,
                                      C.c_uint64, C.POINTER(_VP)]),
]

DECLARED_SYMBOLS = [s[0] for s in _SIGS]


def lib():
    """Load the shared library (built in-tree by __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ArenaError(-2, f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                 "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, res, args in _SIGS:
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise ArenaError(rc, lib().arena_last_error().decode("utf-8", "replace"))


_initialised_device = None


def init_devices(devices) -> None:
    """arena_acct_init_devices: one process, several GPUs (devices[0] is also the device of the
    single-device calls)."""
    global _initialised_device
    ids = (C.c_int * len(devices))(*[int(d) for d in devices])
    check(lib().arena_acct_init_devices(len(devices), ids))
    _initialised_device = int(devices[0])


def init(device: int = 0) -> None:
    """arena_acct_init: bind this process to one GPU (idempotent)."""
    global _initialised_device
    check(lib().arena_acct_init(int(device)))
    _initialised_device = int(device)
