"""GPU box, debug build (tools/variants.sh timeline "-DARENA_TIMELINE"; ARENA_GPUACCT_LIB=build/lib_timeline.so):
where the time of ONE small top-node launch goes, from per-CTA global-timer marks.
usage: python tools/timeline.py [pods]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import arena_b200 as ab

P = int(sys.argv[1]) if len(sys.argv) > 1 else 909_312
ab.init(0)
host = ab.HostSnapshot.synth(P, max(1, P // 100), seed=1)
dev = ab.DeviceSnapshot.upload(host)
N = host.n_nodes
d_total = torch.empty(N, dtype=torch.int64, device="cuda")
d_alloc = torch.empty(N, dtype=torch.int64, device="cuda")
d_cluster = torch.empty(2, dtype=torch.int64, device="cuda")
st = torch.cuda.Stream()
os.environ["ARENA_NO_GRAPH"] = "1"
with torch.cuda.stream(st):
    for _ in range(20):
        ab.top_node_device(dev, d_total, d_alloc, d_cluster, None, st.cuda_stream)
    torch.cuda.synchronize()
lib = C.CDLL(os.environ["ARENA_GPUACCT_LIB"])
n_cta = int(os.environ.get("N_CTA", "888"))
buf = (C.c_ulonglong * (n_cta * 6))()
assert lib.arena_debug_timeline(buf, n_cta * 6) == 0
raw = np.array(buf, dtype=np.uint64).reshape(n_cta, 6)
smid = (raw[:, 0] >> np.uint64(48)).astype(np.int64)
t = raw.astype(np.int64)
t[:, 0] = (raw[:, 0] & np.uint64(0x0000FFFFFFFFFFFF)).astype(np.int64)
t[:, 1:] = (raw[:, 1:] & np.uint64(0x0000FFFFFFFFFFFF)).astype(np.int64)
t0 = t[:, 0].min()
names = ["entry", "last tile done", "loop left", "capacities done", "before totals", "exit"]
for k, name in enumerate(names):
    col = t[:, k] - t0
    print(f"{name:18s} min {col.min() / 1e3:7.2f}  median {np.median(col) / 1e3:7.2f}  max {col.max() / 1e3:7.2f} us")

# which CTAs are late: by SM
done = (t[:, 2] - t0) / 1e3
per_sm = {}
for b in range(n_cta):
    per_sm.setdefault(int(smid[b]), []).append(done[b])
means = sorted((float(np.mean(v)), sm, len(v), float(np.max(v) - np.min(v))) for sm, v in per_sm.items())
print("SMs:", len(per_sm), " CTAs per SM:", sorted(set(len(v) for v in per_sm.values())))
print("loop-left time per SM (mean over its CTAs): fastest", [(round(m, 1), sm) for m, sm, _, _ in means[:8]])
print("                                            slowest", [(round(m, 1), sm) for m, sm, _, _ in means[-8:]])
print("spread inside an SM (max - min over its CTAs): median %.2f us, max %.2f us" % (np.median([x[3] for x in means]), max(x[3] for x in means)))
print("spread of the per-SM means: %.2f .. %.2f us" % (means[0][0], means[-1][0]))
order = np.argsort(smid)
half = [float(np.mean([m for m, sm, _, _ in means if sm % 2 == p])) for p in (0, 1)]
print("mean by SM id parity:", [round(x, 2) for x in half], " by SM id < 74 / >= 74:",
      round(float(np.mean([m for m, sm, _, _ in means if sm < 74])), 2), round(float(np.mean([m for m, sm, _, _ in means if sm >= 74])), 2))
print("per-SM means in SM id order:", " ".join(f"{m:.0f}" for m, sm in sorted(((m, sm) for m, sm, _, _ in means), key=lambda x: x[1])))
