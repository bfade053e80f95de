"""GPU box: ONE process driving every GPU through the C ABI (arena_acct_init_devices + arena_*_multi) on
BASELINE configs[3] / configs[4] at full size, against the C oracle, with wall-clock timings of the calls.
  python tools/multi_device_check.py [n_gpus] [pods]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import arena_b200 as ab
from oracle import cref

G = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
P = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
ab.init_devices(list(range(G)))
ok_all = True
for skewed, seed in ((False, 1), (True, 5)):
    host = ab.HostSnapshot.synth(P, P // 100, seed=seed, skewed=skewed)
    tot, alloc, cl = cref.top_node_columnar_threads(host.cols, 16)
    t0 = time.perf_counter()
    ms = ab.MultiSnapshot(host)
    t_up = time.perf_counter() - t0
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        r = ab.top_node_multi(ms)
        best = min(best, time.perf_counter() - t0)
    ok = np.array_equal(r.node_total, tot) and np.array_equal(r.node_alloc, alloc) and np.array_equal(r.cluster, cl)
    pin = lambda n: torch.zeros(max(n, 1), dtype=torch.int64).pin_memory().numpy()[:n]
    out = ab.TopNodeResult(pin(host.n_nodes), pin(host.n_nodes), pin(2), None)
    best_pinned = 1e9
    for _ in range(5):
        t0 = time.perf_counter()
        r2 = ab.top_node_multi(ms, out=out)
        best_pinned = min(best_pinned, time.perf_counter() - t0)
    ok = ok and np.array_equal(r2.node_alloc, alloc) and np.array_equal(r2.cluster, cl)
    ok_all &= ok
    sizes = np.diff(ms.pod_cut.astype(np.int64))
    print(f"top node x{G} one process, {'skewed' if skewed else 'uniform'} P={P}: upload {t_up*1e3:.0f} ms, query (kernels + download of the "
          f"{host.n_nodes}-row table) {best*1e3:.2f} ms = {P/best:.3g} pods/s into pageable arrays, {best_pinned*1e3:.2f} ms = {P/best_pinned:.3g} pods/s into "
          f"page-locked ones, largest shard {sizes.max()*G/P:.4f} of the mean, "
          f"parity={'OK' if ok else 'FAIL'}", flush=True)
    ms.close()
J = P // 20
host = ab.HostSnapshot.synth(P // 5, P // 500, seed=3, n_jobs=J)
req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
byjob = host.regroup(by_job=True)
ms = ab.MultiSnapshot(byjob, by_job=True)
best = 1e9
for _ in range(5):
    t0 = time.perf_counter()
    r = ab.top_job_multi(ms)
    best = min(best, time.perf_counter() - t0)
ok = (np.array_equal(r.job_req, req) and np.array_equal(r.job_alloc, alloc) and np.array_equal(r.totals, totals)
      and np.array_equal(r.job_order, np.lexsort((np.arange(J), -req)).astype(np.uint32)))
ok_all &= ok
print(f"top job x{G} one process, {P // 5} pods / {J} jobs: query (kernels + download + merge) {best*1e3:.2f} ms, parity={'OK' if ok else 'FAIL'}", flush=True)
sys.exit(0 if ok_all else 1)
