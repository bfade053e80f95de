"""GPU box: top node on pods in arbitrary list order (not grouped by node)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import arena_b200 as ab
from oracle import cref
ab.init(0)
for P, N in [(20_000_000, 200_000), (100_000_000, 1_000_000)]:
    host = ab.HostSnapshot.synth(P, N, seed=1, grouped=False)
    dev = ab.DeviceSnapshot.upload(host)
    d_total = torch.empty(N, dtype=torch.int64, device="cuda"); d_alloc = torch.empty_like(d_total)
    d_cl = torch.empty(2, dtype=torch.int64, device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            ab.top_node_device(dev, d_total, d_alloc, d_cl, None, st.cuda_stream)
        torch.cuda.synchronize()
        ab._lib.check(ab.lib().arena_profile_begin(20))
        for _ in range(20):
            ab.top_node_device(dev, d_total, d_alloc, d_cl, None, st.cuda_stream)
        torch.cuda.synchronize()
    a, b, c = C.c_double(), C.c_double(), C.c_uint64()
    ab._lib.check(ab.lib().arena_profile_end(C.byref(a), C.byref(b), C.byref(c)))
    tot, alloc, cl = cref.top_node_columnar_threads(host.cols, 16)
    ok = np.array_equal(d_alloc.cpu().numpy(), alloc) and np.array_equal(d_cl.cpu().numpy(), cl)
    ms = b.value / c.value
    print(f"ungrouped P={P}: kernel {ms*1e3:.1f} us  {P/(ms/1e3):.3g} pods/s  parity={'OK' if ok else 'FAIL'}", flush=True)
