"""GPU box: time `top job` (device entry point) at a few sizes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import arena_b200 as ab
from oracle import cref
ab.init(0)
for P, N, J in [(10_000, 1_000, 1_000), (1_000_000, 100_000, 50_000), (20_000_000, 200_000, 1_000_000)]:
    host = ab.HostSnapshot.synth(P, N, seed=3, n_jobs=J)
    dev = ab.DeviceSnapshot.upload(host)
    d_req = torch.empty(J, dtype=torch.int64, device="cuda"); d_alloc = torch.empty_like(d_req)
    d_order = torch.empty(J, dtype=torch.int32, device="cuda"); d_latest = torch.empty_like(d_order)
    d_tot = torch.empty(2, dtype=torch.int64, device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            ab.top_job_device(dev, d_req, d_alloc, d_order, d_latest, d_tot, st.cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        K = 20
        for _ in range(K):
            ab.top_job_device(dev, d_req, d_alloc, d_order, d_latest, d_tot, st.cuda_stream)
        e1.record(st)
        torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
    ok = np.array_equal(d_req.cpu().numpy(), req) and np.array_equal(d_tot.cpu().numpy(), totals)
    byts = 22 * P + 9 * host.n_containers
    print(f"top job P={P} J={J}: {ms*1e3:.1f} us/query  {P/(ms/1e3):.3g} pods/s  ~{byts/ms/1e6:.0f} GB/s  parity={'OK' if ok else 'FAIL'}", flush=True)
