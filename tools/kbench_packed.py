"""Launch list of the packed one-shot call (run under ncu): python tools/kbench_packed.py [pods] [nodes]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import arena_b200 as ab
P = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
ab.init(0)
h = ab.HostSnapshot.synth(P, N, seed=1, pinned=False)
p = h.pack(pinned=True)
for _ in range(3):
    t0 = time.perf_counter(); r = ab.top_node_packed(p); dt = time.perf_counter() - t0
print("packed call %.3f ms, bytes %d, cluster %s" % (dt * 1e3, p.nbytes, r.cluster.tolist()))
