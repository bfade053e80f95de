"""Under torchrun, one rank per GPU: host -> device copy rate from page-locked memory, every rank at once, and
the same for rank 0 alone.  Names the limiter of the multi-GPU end-to-end numbers (all uploads share the host's
memory and PCIe complex).  python -m torch.distributed.run --nproc-per-node N tools/pcie_aggregate.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from arena_b200 import multigpu

rank, world, local = multigpu.env_rank()
torch.cuda.set_device(local)
multigpu.init_process_group("nccl")
MB = 256
src = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
dst = torch.empty(MB << 20, dtype=torch.uint8, device="cuda")


def rate(active):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if active:
        for _ in range(8):
            dst.copy_(src, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return 8 * MB / 1024 / (float(t.item()) / 1e3)      # GB/s per active rank, by the slowest rank's clock


rate(True)
alone = rate(rank == 0)
together = rate(True)
if rank == 0:
    print(f"H2D from pinned memory, {MB} MiB x 8: rank 0 alone {alone:.1f} GB/s; {world} ranks at once {together:.1f} GB/s each = "
          f"{together * world:.1f} GB/s aggregate", flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
