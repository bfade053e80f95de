"""GPU box: what a `top node` STEP costs beyond its kernel at a small shard (1.25e7 pods, the 8-GPU share of
BASELINE configs[3]): per-call event profiling on / off, and the step replayed from a CUDA graph.  Since the
library replays the step from a graph of its own, run with ARENA_NO_GRAPH=1 to see plain launches in the first line."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import arena_b200 as ab

P = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 500
ab.init(0)
host = ab.HostSnapshot.synth(P, P // 100, seed=1)
dev = ab.DeviceSnapshot.upload(host)
N = host.n_nodes
d_total = torch.empty(N, dtype=torch.int64, device="cuda")
d_alloc = torch.empty(N, dtype=torch.int64, device="cuda")
d_cluster = torch.empty(2, dtype=torch.int64, device="cuda")
st = torch.cuda.Stream()


def loop(fn, label):
    with torch.cuda.stream(st):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(K):
            fn()
        e1.record(st)
        torch.cuda.synchronize()
    print(f"{label}: {e0.elapsed_time(e1) / K * 1e3:.1f} us/step", flush=True)


step = lambda: ab.top_node_device(dev, d_total, d_alloc, d_cluster, None, st.cuda_stream)
loop(step, "library calls, no profiling (graph replay inside the library unless ARENA_NO_GRAPH=1)")
ab._lib.check(ab.lib().arena_profile_begin(K + 10))
loop(step, "library calls, per-call events (arena_profile_begin: plain launches)")
a, b, c = C.c_double(), C.c_double(), C.c_uint64()
ab._lib.check(ab.lib().arena_profile_end(C.byref(a), C.byref(b), C.byref(c)))
print(f"   events: memset {a.value / c.value * 1e3:.1f} us, kernel {b.value / c.value * 1e3:.1f} us", flush=True)
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(st):
    step()
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=st):
        step()
loop(g.replay, "CUDA graph of the step (memset + kernel)")
g4 = torch.cuda.CUDAGraph()
with torch.cuda.stream(st):
    with torch.cuda.graph(g4, stream=st):
        for _ in range(4):
            step()
K //= 4
loop(g4.replay, "CUDA graph of 4 steps (x4 per replay)")
