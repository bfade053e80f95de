"""Multi-GPU parity under torchrun (one rank per GPU): every rank's slice of the `top node` table
against the C oracle on its shard -- uniform and skewed snapshots, the in-kernel totals exchange
collected by the same launch and one step late, and NCCL -- and the job-range `top job`.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/multigpu_check.py [pods]
Exit status 0 only when every rank agrees with the oracle everywhere."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import arena_b200 as ab
from arena_b200 import multigpu
from oracle import cref

P = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
rank, world, local = multigpu.env_rank()
torch.cuda.set_device(local)
multigpu.init_process_group("nccl")
ab.init(local)
fails = 0


def check(name, ok):
    global fails
    bad = torch.tensor([0 if ok else 1], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(bad)
    if rank == 0:
        print(f"{name}: {'OK' if int(bad.item()) == 0 else 'FAIL on %d rank(s)' % int(bad.item())}", flush=True)
    fails += int(bad.item())


assert multigpu.setup_peer_exchange() or world == 1
st = torch.cuda.current_stream().cuda_stream
for skewed, seed in ((False, 1), (True, 5)):
    full = ab.HostSnapshot.synth(P, P // 100, seed=seed, skewed=skewed)
    ftot, falloc, fcl, _ = cref.top_node_columnar(full.cols)
    key_cut, pod_cut = full.shard_plan(world)
    sh = full.shard(rank, world)
    dev = ab.DeviceSnapshot.upload(sh)
    lo, hi = int(key_cut[rank]), int(key_cut[rank + 1])
    d_total = torch.empty(hi - lo, dtype=torch.int64, device="cuda")
    d_alloc = torch.empty_like(d_total)
    d_cl = torch.zeros(2, dtype=torch.int64, device="cuda")
    tag = "skewed" if skewed else "uniform"
    for mode in ("fused", "posted", "nccl"):
        d_alloc.fill_(-5)
        if mode == "fused":
            ab.top_node_device_fused(dev, d_total, d_alloc, d_cl, st)
        elif mode == "posted":
            for _ in range(3):
                ab.top_node_device_posted(dev, d_total, d_alloc, d_cl, st)
            ab.peer_exchange_collect(d_cl, st)
        else:
            ab.top_node_device(dev, d_total, d_alloc, d_cl, None, st)
            multigpu.all_reduce_totals(d_cl)
        torch.cuda.synchronize()
        ok = (np.array_equal(d_alloc.cpu().numpy(), falloc[lo:hi]) and np.array_equal(d_total.cpu().numpy(), ftot[lo:hi])
              and d_cl.cpu().numpy().tolist() == fcl.tolist())
        check(f"top node {tag} {mode} x{world}", ok)
    # the one-shot call with host buffers (wide and packed) + all-reduce
    for snap, nm in ((sh, "host wide"), (sh.pack(), "host packed")):
        res, totals = multigpu.top_node_sharded(snap)
        check(f"top node {tag} {nm} x{world}", np.array_equal(res.node_alloc, falloc[lo:hi]) and totals.tolist() == fcl.tolist())
    dev.close()
assert ab.lib().arena_peer_exchange_timed_out() == 0

J = max(world * 3, P // 40)
full = ab.HostSnapshot.synth(P, P // 100, seed=7, n_jobs=J)
req, alloc, latest, totals = cref.top_job_columnar(full.cols, J)
byjob = full.regroup(by_job=True)
src = byjob.pod_order()
res, (lo, hi), gtot, order = multigpu.top_job_sharded(byjob)
has = res.job_latest != 0xFFFFFFFF
ok = (np.array_equal(res.job_req, req[lo:hi]) and np.array_equal(res.job_alloc, alloc[lo:hi]) and gtot.tolist() == totals.tolist()
      and np.array_equal(src[res.job_latest[has]], latest[lo:hi][has]) and np.array_equal(has, latest[lo:hi] != 0xFFFFFFFF))
check(f"top job job-range shards x{world}", ok)
if rank == 0:
    want = np.lexsort((np.arange(J), -req)).astype(np.uint32)
    ok_order = np.array_equal(order, want)
else:
    ok_order = True
check(f"top job merged order x{world}", ok_order)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
sys.exit(1 if fails else 0)
