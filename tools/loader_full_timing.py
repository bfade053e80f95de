"""GPU box: where the time of arena_load_json_gpu goes (10k-pod API-server-shaped PodList).
ARENA_LOADER_TRACE=1 prints the host loader's phases to stderr."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import arena_b200 as ab
import loader_timing

n_pods = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000
pods, nodes = loader_timing.make_docs(n_pods, max(1, n_pods // 10), seed=1)
ab.init(0)
pinned = ab.PinnedBuffer(len(pods)).fill(pods)
for name, fn in (("host arena_load_json", lambda: (ab.HostSnapshot.from_json(pods, nodes), None)),
                 ("gpu  arena_load_json_gpu (pinned)", lambda: ab.HostSnapshot.from_json_gpu(pinned, nodes)),
                 ("gpu  arena_load_json_device (pinned, top-node columns only)", lambda: ab.DeviceSnapshot.from_json_device(pinned, nodes))):
    best, st = 1e9, None
    for _ in range(6):
        t0 = time.perf_counter()
        _, st = fn()
        best = min(best, (time.perf_counter() - t0) * 1e3)
    print(f"{name}: {best:.2f} ms", {k: round(v, 2) if isinstance(v, float) else v for k, v in (st or {}).items()}, flush=True)
os.environ["ARENA_LOADER_TRACE"] = "1"
ab.HostSnapshot.from_json_gpu(pinned, nodes)
