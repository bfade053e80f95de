"""GPU box: the packed summary kernel on device-resident packed columns (parity against the C oracle, then
CUDA-event timing), and the pipelined one-shot call with host buffers.  python tools/kbench_packed.py [pods] [nodes]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import arena_b200 as ab
from oracle import cref

P = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
N = int(sys.argv[2]) if len(sys.argv) > 2 else P // 100
ab.init(0)
for skewed, seed in (((False, 1),) if os.environ.get("KBENCH_UNIFORM_ONLY") else ((False, 1), (True, 5))):
    h = ab.HostSnapshot.synth(P, N, seed=seed, skewed=skewed)
    tot, alloc, cl = cref.top_node_columnar_threads(h.cols, 16)
    t0 = time.perf_counter()
    p = h.pack(pinned=True)
    pack_ms = (time.perf_counter() - t0) * 1e3
    dev = ab.PackedDeviceSnapshot(p)
    d_total = torch.empty(N, dtype=torch.int64, device="cuda")
    d_alloc = torch.empty(N, dtype=torch.int64, device="cuda")
    d_cl = torch.empty(2, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream()
    for _ in range(5):
        ab.top_node_packed_device(dev, d_total, d_alloc, d_cl, st.cuda_stream)
    torch.cuda.synchronize()
    ok = np.array_equal(d_alloc.cpu().numpy(), alloc) and np.array_equal(d_total.cpu().numpy(), tot) and d_cl.cpu().numpy().tolist() == cl.tolist()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 50
    e0.record(st)
    for _ in range(K):
        ab.top_node_packed_device(dev, d_total, d_alloc, d_cl, st.cuda_stream)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    print(f"{'skewed' if skewed else 'uniform'} P={P}: packed kernel {ms*1e3:.1f} us  {dev.launch_bytes/ms/1e6:.0f} GB/s of its own {dev.launch_bytes} B  "
          f"{P/ms*1e3:.3g} pods/s  parity={'OK' if ok else 'FAIL'}  pack {pack_ms:.0f} ms  pieces {int(p.view.n_pieces)}", flush=True)
    pin = lambda n: torch.zeros(max(n, 1), dtype=torch.int64).pin_memory().numpy()[:n]
    out = ab.TopNodeResult(pin(N), pin(N), pin(2), None)
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        r = ab.top_node_packed(p, out=out)
        best = min(best, time.perf_counter() - t0)
    ok2 = np.array_equal(r.node_alloc, alloc) and r.cluster.tolist() == cl.tolist()
    print(f"   one-shot call (pinned in/out): {best*1e3:.3f} ms for {p.nbytes} B up  parity={'OK' if ok2 else 'FAIL'}", flush=True)
    dev.close()
