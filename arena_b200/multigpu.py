"""Multi-GPU driver: one process per GPU, node-range shards, one all-reduce.

Nodes are independent and a pod only touches its own node, so a snapshot cut
into contiguous node ranges (arena_shard_plan: about equal pod counts, cuts at
node boundaries only) needs no data-path exchange: every rank produces a
disjoint slice of the per-node table.  The only exchange is the pair of
cluster totals {sum node_total, sum node_alloc} (top_node.go:109-118): one
int64[2] all-reduce (NCCL over NVLink on GPUs, gloo in the CPU tests).
Integer sums are order-independent, so any rank count gives identical bits.
"""

from __future__ import annotations

import os
from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist


def env_rank() -> Tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_process_group(backend: Optional[str] = None) -> Tuple[int, int, int]:
    """Initialise torch.distributed from the torchrun environment (idempotent)."""
    rank, world, local = env_rank()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, world, local


def all_reduce_totals(totals: torch.Tensor) -> torch.Tensor:
    """Sum an int64[2] (or [4]) totals tensor over all ranks, in place.  With
    one rank this is a no-op (no collective is issued)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(totals, op=dist.ReduceOp.SUM)
    return totals


def gather_tables(local_total: np.ndarray, local_alloc: np.ndarray, dst: int = 0):
    """Concatenate the per-rank slices of the node table on rank `dst` (the
    ranks hold disjoint, ordered node ranges).  Returns (total, alloc) on dst,
    (None, None) elsewhere."""
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return local_total, local_alloc
    world, rank = dist.get_world_size(), dist.get_rank()
    out: Optional[List] = [None] * world if rank == dst else None
    dist.gather_object((local_total, local_alloc), out, dst=dst)
    if rank != dst:
        return None, None
    return np.concatenate([o[0] for o in out]), np.concatenate([o[1] for o in out])


def bind_to_gpu_numa_node(local_rank: int):
    """Run this process on the CPUs of the NUMA node its GPU hangs off, so that the page-locked
    staging buffers it allocates afterwards (first touch) sit next to the GPU's PCIe root: with
    one process per GPU all uploading at once, remote-socket memory is what limits them.
    Returns the node number, or None when the topology cannot be read (nothing is changed)."""
    import os
    import subprocess
    try:
        bus = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(local_rank)],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]                                   # sysfs uses a 4-digit PCI domain
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return None
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def setup_peer_exchange() -> bool:
    """Map every rank's totals mailbox into every process (CUDA IPC over NVLink) so that
    top_node_device_fused can trade the cluster totals inside the kernel.  Collective: every
    rank calls it.  Returns False (and leaves NCCL as the only way) when IPC is unavailable."""
    import ctypes as C
    from ._lib import lib
    rank, world, _ = env_rank()
    if world > 16:
        return False
    mine = (C.c_ubyte * 64)()
    ok = lib().arena_peer_exchange_create(rank, world, mine) == 0
    handles = [None] * world
    if dist.is_initialized() and world > 1:
        dist.all_gather_object(handles, (ok, bytes(mine)))
    else:
        handles = [(ok, bytes(mine))]
    if all(h[0] for h in handles):
        blob = b"".join(h[1] for h in handles)
        ok = lib().arena_peer_exchange_connect(blob) == 0
    else:
        ok = False
    flags = [None] * world
    if dist.is_initialized() and world > 1:
        dist.all_gather_object(flags, ok)
    else:
        flags = [ok]
    if not all(flags):
        lib().arena_peer_exchange_destroy()
        return False
    return True


def top_node_sharded(host_shard, details: bool = False, out=None):
    """End-to-end sharded `top node` for this rank's shard (a HostSnapshot, or a
    PackedSnapshot of it): upload + kernels + download through the C ABI, then the totals
    all-reduce.  `out`: a TopNodeResult to write into (page-locked arrays make the download a
    straight DMA).  Returns (TopNodeResult of the shard, global cluster totals int64[2])."""
    from .snapshot import PackedSnapshot
    from .top import top_node, top_node_packed
    if isinstance(host_shard, PackedSnapshot):
        res = top_node_packed(host_shard, details=details, out=out)
    else:
        res = top_node(host_shard, details=details, out=out)
    t = torch.from_numpy(res.cluster.copy())
    if dist.is_initialized() and dist.get_world_size() > 1 and dist.get_backend() == "nccl":
        t = t.cuda()
    all_reduce_totals(t)
    return res, t.cpu().numpy()


def merge_job_orders(reqs, orders, key_cut):
    """The global `top job` ordering from per-shard results.  Shard i holds the jobs
    [key_cut[i], key_cut[i+1]) and has them ordered by (requested desc, id asc) already
    (trainer.go:25-37 + the documented tie-break), so the global order is a MERGE of runs:
    a stable sort by requested of the runs laid end to end in shard order, because equal
    requested values then stay in ascending-id order.  reqs[i]: int64 per-job requested of shard i
    (shard-local ids), orders[i]: uint32 shard-local order.  Returns global job ids, uint32."""
    runs_id, runs_req = [], []
    for i, (req, order) in enumerate(zip(reqs, orders)):
        order = np.asarray(order, np.int64)
        runs_id.append(order + int(key_cut[i]))
        runs_req.append(np.asarray(req, np.int64)[order])
    ids = np.concatenate(runs_id) if runs_id else np.zeros(0, np.int64)
    req = np.concatenate(runs_req) if runs_req else np.zeros(0, np.int64)
    # descending by requested, stable: flip the sign bit trick is not needed with int64 keys
    # unless requested can be INT64_MIN; negate through uint64 to stay exact
    key = (~req.view(np.uint64)) ^ np.uint64(0x8000000000000000)      # ascending key == requested descending
    return ids[np.argsort(key, kind="stable")].astype(np.uint32)


def top_job_sharded(host_by_job, out_device=None):
    """Multi-GPU `top job`: jobs span nodes, so the snapshot -- rows grouped by job,
    HostSnapshot.regroup(by_job=True) -- is cut into job-id ranges of about equal pod count, one per
    rank; each rank runs the ordinary `top job` kernels on its range; the per-job tables are disjoint
    slices, the two totals (list.go:102-103) are all-reduced, and the per-rank orderings are merged.
    Every rank passes the same full snapshot.  Returns (TopJobResult of this rank's shard with
    job_latest as GLOBAL row indices, key range (lo, hi), global totals int64[2], and on rank 0 the
    global job order (else None))."""
    from .snapshot import DeviceSnapshot
    from .top import top_job
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
    key_cut, pod_cut = host_by_job.shard_plan(world, by_job=True)
    shard = host_by_job.shard(rank, world, by_job=True)
    dev = DeviceSnapshot.upload(shard)
    res = top_job(dev)
    dev.close()
    latest = res.job_latest.astype(np.int64)
    latest[res.job_latest != 0xFFFFFFFF] += int(pod_cut[rank])
    res.job_latest = latest.astype(np.uint32)
    t = torch.from_numpy(res.totals.copy())
    if dist.is_initialized() and world > 1 and dist.get_backend() == "nccl":
        t = t.cuda()
    all_reduce_totals(t)
    order = None
    if world > 1:
        gathered = [None] * world if rank == 0 else None
        dist.gather_object((res.job_req, res.job_order), gathered, dst=0)
        if rank == 0:
            order = merge_job_orders([g[0] for g in gathered], [g[1] for g in gathered], key_cut)
    else:
        order = res.job_order
    return res, (int(key_cut[rank]), int(key_cut[rank + 1])), t.cpu().numpy(), order
