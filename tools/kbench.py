"""Kernel iteration helper (GPU box): quick parity check + pods-kernel timing.
usage: python tools/kbench.py [pods] [nodes] [steps]"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import arena_b200 as ab
from oracle import cref

P = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
N = int(sys.argv[2]) if len(sys.argv) > 2 else P // 100
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
ab.init(0)

# parity on a few shapes
for (p, n, seed, kw) in [] if os.environ.get("KBENCH_NOPARITY") else [(1_000_000, 10_000, 1, {}), (129, 3, 2, {}), (300_000, 3_000, 3, dict(grouped=False)),
                         (2_000_000, 20_000, 5, dict(skewed=True))]:
    h = ab.HostSnapshot.synth(p, n, seed=seed, **kw)
    d = ab.DeviceSnapshot.upload(h)
    r = ab.top_node(d, details=True)
    tot, alloc, cl, pg = cref.top_node_columnar(h.cols, True)
    ok = (np.array_equal(r.node_total, tot) and np.array_equal(r.node_alloc, alloc)
          and np.array_equal(r.cluster, cl) and np.array_equal(r.pod_gpu, pg))
    print("parity", p, n, kw, "OK" if ok else "MISMATCH", flush=True)
    if not ok:
        sys.exit(1)
    d.close()

host = ab.HostSnapshot.synth(P, N, seed=1, skewed=bool(os.environ.get("KBENCH_SKEWED")))
dev = ab.DeviceSnapshot.upload(host)
d_total = torch.empty(N, dtype=torch.int64, device="cuda")
d_alloc = torch.empty(N, dtype=torch.int64, device="cuda")
d_cluster = torch.empty(2, dtype=torch.int64, device="cuda")
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    for _ in range(5):
        ab.top_node_device(dev, d_total, d_alloc, d_cluster, None, st.cuda_stream)
    torch.cuda.synchronize()
    ab._lib.check(ab.lib().arena_profile_begin(steps))
    for _ in range(steps):
        ab.top_node_device(dev, d_total, d_alloc, d_cluster, None, st.cuda_stream)
    torch.cuda.synchronize()
a, b, c = C.c_double(), C.c_double(), C.c_uint64()
ab._lib.check(ab.lib().arena_profile_end(C.byref(a), C.byref(b), C.byref(c)))
pods_ms = b.value / c.value
byts = 9 * host.n_pods + 9 * host.n_containers + 17 * host.n_nodes
print(f"P={P} pods_kernel {pods_ms*1e3:.1f} us  {byts/pods_ms/1e6:.0f} GB/s  frac_of_6538.9={byts/pods_ms/1e6/6538.9:.3f}  "
      f"memset {a.value/c.value*1e3:.1f} us", flush=True)

g, m = C.c_double(), C.c_double()
ab._lib.check(ab.lib().arena_calibrate_stream_read(dev.handle, 20, C.byref(g), C.byref(m)))
print(f"stream-read calibration: {m.value*1e3:.1f} us  {g.value:.0f} GB/s  frac_of_6538.9={g.value/6538.9:.3f}", flush=True)
ab._lib.check(ab.lib().arena_calibrate_stream_pattern(dev.handle, 20, C.byref(g), C.byref(m)))
print(f"tile-pattern calibration: {m.value*1e3:.1f} us  {g.value:.0f} GB/s requested", flush=True)
