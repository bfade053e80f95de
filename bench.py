#!/usr/bin/env python
"""bench.py -- pod-records/s and HBM GB/s of `arena top node` GPU accounting.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path (memset of node_alloc + the top-node
launch, plus the exchange of the int64[2] cluster totals when N > 1) over one
synthetic columnar snapshot that is already resident in HBM.  Workload at N = 1:
BASELINE.json configs[3]'s snapshot (1e8 pods / 1e6 nodes, seed 1; 2.3 GB, larger
than the 126 MB L2, so no flush is needed between steps) on ONE GPU -- the
configuration the >= 70 % of HBM bandwidth target is quoted on.  At N > 1 the SAME
snapshot is cut into N node ranges of about equal pod count, one per rank (strong
scaling: BASELINE configs[3] as written); `--scaling weak` instead gives every rank
a snapshot of the full size (N x 1e8 pods, seed 1 + rank).

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement".
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "pod-records/sec & HBM GB/s for top-node GPU accounting at 1/2/4/8 B200"
UNIT = "pod-records/s"
PODS_FULL = 100_000_000
NODES_FULL = 1_000_000


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while running."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def stop(self, t0: float, t1: float):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, ln in self.lines:
            if ts < t0 - 0.05 or ts > t1 + 0.25:
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples in timed region"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------
# CPU legs (the ONLY place bench.py touches oracle/): bounded samples.
# --------------------------------------------------------------------------
def oracle_parity(host, d_total, d_alloc):
    """This rank's slice of the table against the C oracle's columnar pass over the same shard
    (oracle/ used as the checker, after the timed region).  Returns (ok, oracle cluster totals)."""
    from oracle import cref
    tot, alloc, cl = cref.top_node_columnar_threads(host.cols, min(16, os.cpu_count() or 1))
    import numpy as np
    ok = bool(np.array_equal(d_total.cpu().numpy(), tot) and np.array_equal(d_alloc.cpu().numpy(), alloc))
    return ok, cl


def cpu_reference_run(steps: int, warmup: int, sample_pods: int, threads: int, min_seconds: float = 0.0):
    """The reference's CPU algorithm (C restatement, oracle/acct_ref.c) on a
    bounded sample of the workload: array-of-structs pods with string node
    names and keyed limits, string-hash join, all host threads."""
    import numpy as np  # noqa: F401
    from arena_b200 import HostSnapshot
    from oracle import cref
    sample_nodes = max(1, sample_pods // 100)
    host = HostSnapshot.synth(sample_pods, sample_nodes, seed=1)
    aos = cref.AoS(host.cols)
    for _ in range(warmup):
        aos.run(False, threads)
    t0 = time.perf_counter()
    done = 0
    while done < steps or (time.perf_counter() - t0) < min_seconds:
        total, alloc, cluster = aos.run(False, threads)
        done += 1
    steps = done
    dt = time.perf_counter() - t0
    # single-thread columnar pass (B3) and faithful nested loop (B1) for context
    t1 = time.perf_counter()
    cref.top_node_columnar(host.cols)
    b3 = host.n_pods / (time.perf_counter() - t1)
    # the same columnar pass on all host threads: the comparator for a caller that already holds columns
    cref.top_node_columnar_threads(host.cols, threads)
    t1 = time.perf_counter()
    reps = 0
    while reps < 3 or time.perf_counter() - t1 < 1.0:
        cref.top_node_columnar_threads(host.cols, threads)
        reps += 1
    b3_all = host.n_pods * reps / (time.perf_counter() - t1)
    small = HostSnapshot.synth(20_000, 200, seed=1)
    a2 = cref.AoS(small.cols)
    t2 = time.perf_counter()
    a2.run(True)
    b1 = small.n_pods / (time.perf_counter() - t2)
    a2.close()
    aos.close()
    return {
        "value": host.n_pods * steps / dt, "ms_per_step": dt / steps * 1e3, "sample_pods": host.n_pods, "passes": steps,
        "seconds": dt,
        "sample_nodes": sample_nodes, "b3_columnar_1thread": b3, "b3_columnar_all_threads": b3_all,
        "b1_nested_loop_1thread": b1,
        "b1_sample": "20000 pods x 200 nodes",
    }


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    r = cpu_reference_run(args.steps, max(1, min(args.warmup, 2)), args.sample_pods, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": args.scaling if args.gpus > 1 else "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": dict(workload_config(args.gpus, args.scaling, collective=args.collective),
                       pods_per_gpu=PODS_FULL // (args.gpus if args.scaling == "strong" else 1),
                       nodes_per_gpu=NODES_FULL // (args.gpus if args.scaling == "strong" else 1),
                       total_pods=PODS_FULL * (args.gpus if args.scaling == "weak" else 1)),
        "cpu_baseline": {
            "value": r["value"], "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"each step = first {r['sample_pods']} pods / {r['sample_nodes']} nodes of the workload's "
                      "generator (seed 1) as array-of-structs with string node names and keyed limits, string-hash "
                      "join (C restatement of gpu.go + top_node.go; the Go reference cannot be built: no toolchain)",
            "columnar_single_pass_1thread": r["b3_columnar_1thread"],
            "columnar_all_threads": r["b3_columnar_all_threads"],
            "faithful_nested_loop_1thread": r["b1_nested_loop_1thread"],
            "faithful_nested_loop_sample": r["b1_sample"],
        },
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


def wire_format_leg(ab, n_pods=10_000, n_nodes=1000, reps=5):
    """JSON PodList/NodeList -> columns -> `top node` table, wall clock around the C-ABI calls."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import loader_timing
    log(f"[rank 0] wire_format leg ({n_pods} pods)...")
    pods, nodes = loader_timing.make_docs(n_pods, n_nodes, seed=1)
    pinned = ab.PinnedBuffer(len(pods)).fill(pods)

    def best(fn):
        t_best, out = float("inf"), None
        for _ in range(reps):
            t0 = time.perf_counter()
            out = fn()
            t_best = min(t_best, time.perf_counter() - t0)
        return t_best * 1e3, out

    def host_path():
        h = ab.HostSnapshot.from_json(pods, nodes)
        return ab.top_node(ab.DeviceSnapshot.upload(h)).cluster.tolist()

    def device_path(src):
        dev, st = ab.DeviceSnapshot.from_json_device(src, nodes)
        return ab.top_node(dev).cluster.tolist(), st

    def full_path(src):
        # the whole host snapshot (names, labels: what -d / top job / list need), PodList decoded on the GPU
        h, st = ab.HostSnapshot.from_json_gpu(src, nodes)
        return ab.top_node(ab.DeviceSnapshot.upload(h)).cluster.tolist(), st

    host_ms, host_cluster = best(host_path)
    dev_ms, (dev_cluster, st) = best(lambda: device_path(pods))
    pin_ms, (pin_cluster, st_pin) = best(lambda: device_path(pinned))
    full_ms, (full_cluster, st_full) = best(lambda: full_path(pinned))
    assert host_cluster == dev_cluster == pin_cluster == full_cluster, "loaders disagree"
    return {
        "workload": f"API-server-shaped PodList, {n_pods} pods / {n_nodes} nodes (BASELINE configs[1] scale), "
                    f"{len(pods)} bytes of JSON, then one `top node`",
        "host_loader_ms": host_ms, "host_threads": os.cpu_count(),
        "gpu_loader_ms": dev_ms, "gpu_loader_pinned_input_ms": pin_ms,
        "gpu_loader_h2d_ms": st_pin["ms_h2d"], "fell_back": int(st["fell_back"] or st_pin["fell_back"] or st_full["fell_back"]),
        "gpu_loader_full_snapshot_ms": full_ms,
        "pods_per_s_host": n_pods / host_ms * 1e3, "pods_per_s_gpu_pinned": n_pods / pin_ms * 1e3,
        "json_gb_per_s_gpu_pinned": len(pods) / pin_ms / 1e6,
    }


EXCHANGES = {
    "posted": "cluster totals traded inside the top-node kernel over NVLink peer memory (CUDA IPC mailboxes), "
              "collected one step late (no collective launch)",
    "fused": "cluster totals traded inside the top-node kernel over NVLink peer memory (CUDA IPC mailboxes), "
             "collected by the same launch (no collective launch)",
    "nccl": "NCCL all-reduce of the int64[2] cluster totals after the kernel",
}


def workload_config(n_gpus: int, scaling: str, skewed: bool = False, collective: str = "posted"):
    ex = EXCHANGES[collective]
    if skewed:
        wl = ("BASELINE configs[4]: skewed snapshot of 1e8 pods / 1e6 nodes (power-law pods per node, GPU-sharing nodes, "
              f"seed 5) cut into {n_gpus} node range(s) of equal pod count" + (f"; {ex}" if n_gpus > 1 else ""))
    elif scaling == "strong" and n_gpus > 1:
        wl = (f"BASELINE configs[3]: synthetic snapshot of 1e8 pods / 1e6 nodes (seed 1) cut into {n_gpus} node ranges; {ex}")
    elif n_gpus > 1:
        wl = (f"BASELINE configs[3] per GPU: {n_gpus} node-range shards of 1e8 pods / 1e6 nodes each "
              f"(cluster of {n_gpus}e8 pods / {n_gpus}e6 nodes, seeds 1..{n_gpus}); {ex}")
    else:
        wl = "BASELINE configs[3] on one GPU: synthetic snapshot of 1e8 pods / 1e6 nodes, mixed phases, multi-container pods, seed 1"
    return {"workload": wl, "command": "arena top node (summary)", "pods_per_gpu": None, "nodes_per_gpu": None,
            "l2": "inputs (2.3 GB per GPU) exceed the 126 MB L2; no flush between steps",
            "parallelism": f"node-range shards x{n_gpus}" if n_gpus > 1 else "single GPU"}


def run_job_arm(args):
    """`--command job`: multi-GPU `top job`.  ONE snapshot (default 2e7 pods / 1e6 jobs,
    seed 3), rows regrouped by job and cut into N job-id ranges of about equal pod count; a step is
    the `top job` query of every rank on its resident shard (per-job sums, latest-pod rule, ordering)
    plus the all-reduce of the two totals.  Device-timed, max over ranks; per-rank tables and the
    global totals are checked against the C oracle after the timed region."""
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import numpy as np
    import torch
    import torch.distributed as dist
    import arena_b200 as ab
    from arena_b200 import multigpu
    from oracle import cref
    rank, world, local = multigpu.env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    multigpu.init_process_group("nccl")
    ab.init(local)
    P, J = args.job_pods, args.jobs
    full = ab.HostSnapshot.synth(P, max(1, P // 100), seed=3, n_jobs=J).regroup(by_job=True)
    key_cut, pod_cut = full.shard_plan(world, by_job=True)
    shard = full.shard(rank, world, by_job=True)
    Js = shard.n_jobs
    dev = ab.DeviceSnapshot.upload(shard)
    d_req = torch.empty(Js, dtype=torch.int64, device="cuda")
    d_alloc = torch.empty_like(d_req)
    d_order = torch.empty(Js, dtype=torch.int32, device="cuda")
    d_latest = torch.empty_like(d_order)
    d_tot = torch.empty(2, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream()

    def step():
        ab.top_job_device(dev, d_req, d_alloc, d_order, d_latest, d_tot, stream.cuda_stream)
        multigpu.all_reduce_totals(d_tot)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    L0 = ab.lib().arena_kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    barrier()
    t1 = time.time()
    launches = int(ab.lib().arena_kernel_launches() - L0)
    tmax = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_per_step = float(tmax.item()) / args.steps
    # parity: this rank's job range against the oracle on the shard, totals against the oracle on the whole
    req, alloc, latest, totals = cref.top_job_columnar(shard.cols, Js)
    ok = (np.array_equal(d_req.cpu().numpy(), req) and np.array_equal(d_alloc.cpu().numpy(), alloc)
          and np.array_equal(d_latest.cpu().numpy().view(np.uint32), latest))
    want_order = np.lexsort((np.arange(Js), -req)).astype(np.uint32) if Js else np.zeros(0, np.uint32)
    ok = ok and np.array_equal(d_order.cpu().numpy().view(np.uint32), want_order)
    _, _, _, full_totals = cref.top_job_columnar(full.cols, J)
    ok = ok and d_tot.cpu().numpy().tolist() == full_totals.tolist()
    bad = torch.tensor([0 if ok else 1], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(bad)
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    if rank == 0:
        line = {
            "metric": "pod-records/sec for top-job GPU accounting", "value": P / (ms_per_step / 1e3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "parity_ok": int(bad.item()) == 0,
            "config": {"workload": f"synthetic snapshot of {P} pods / {J} jobs (seed 3) regrouped by job and cut into "
                                   f"{world} job-id range(s); per-rank `top job` query + NCCL all-reduce of the int64[2] totals",
                       "command": "arena top job", "pods_per_gpu": shard.n_pods, "jobs_per_gpu": Js,
                       "parallelism": f"job-range shards x{world}" if world > 1 else "single GPU"},
            "clocks": clocks, "gpu_launches": launches,
        }
        os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if int(bad.item()):
        raise SystemExit("bench.py --command job: parity failure")
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default=None, choices=["weak", "strong"],
                    help="N > 1: strong (default) = ONE snapshot of --pods cut into N node ranges (BASELINE configs[3]); "
                         "weak = --pods per GPU")
    ap.add_argument("--pods", type=int, default=PODS_FULL, help="pods per GPU (weak) or in total (strong)")
    ap.add_argument("--nodes", type=int, default=NODES_FULL)
    ap.add_argument("--e2e-steps", type=int, default=0, help="0 = min(steps, 10)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--skewed", action="store_true",
                    help="BASELINE configs[4]: power-law pods per node + GPU-sharing nodes (seed 5); with --gpus > 1 one "
                         "snapshot is cut into node ranges (implies --scaling strong) and the pod-count imbalance is reported")
    ap.add_argument("--collective", default="posted", choices=["posted", "fused", "nccl"],
                    help="multi-GPU totals: traded inside the kernel over NVLink peer memory and collected one step late "
                         "(posted) or by the same launch (fused), or a NCCL all-reduce after the kernel (nccl)")
    ap.add_argument("--no-wire", action="store_true", help="skip the wire_format leg (JSON PodList -> columns)")
    ap.add_argument("--sample-pods", type=int, default=10_000_000, help="pods per step of the CPU legs (bounded sample)")
    ap.add_argument("--graph", action="store_true", help="also time the step replayed from a CUDA graph (reported as graph_ms_per_step)")
    ap.add_argument("--command", default="node", choices=["node", "job"],
                    help="node: the headline `top node` path; job: the multi-GPU `top job` arm (job-id range shards)")
    ap.add_argument("--job-pods", type=int, default=20_000_000)
    ap.add_argument("--jobs", type=int, default=1_000_000)
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3                         # timing rule: W >= 3
    if args.scaling is None:
        args.scaling = "strong" if args.gpus > 1 else "weak"

    if args.impl == "reference":
        return run_reference_arm(args)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # launched directly: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"),
               os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)

    if args.command == "job":
        return run_job_arm(args)

    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's
    # version banner) are sent to stderr for the duration of the run
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    import numpy as np
    import torch
    import torch.distributed as dist

    import arena_b200 as ab
    from arena_b200 import multigpu

    rank, world, local = multigpu.env_rank()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    numa = multigpu.bind_to_gpu_numa_node(local) if world > 1 else None
    multigpu.init_process_group("nccl")
    ab.init(local)
    n_gpus = world

    # ---- workload ------------------------------------------------------------
    t_gen = time.time()
    if args.skewed:
        args.scaling = "strong"
    seed0 = 5 if args.skewed else 1
    if args.scaling == "strong" and world > 1:
        full = ab.HostSnapshot.synth(args.pods, args.nodes, seed=seed0, pinned=False, skewed=args.skewed)
        host = full.shard(rank, world)
        # pinned staging of the shard for the e2e leg
        host = ab.HostSnapshot.from_arrays({k: v.copy() for k, v in host.cols.items() if v.size or k == "pod_ctr_begin"})
        del full
    else:
        host = ab.HostSnapshot.synth(args.pods, args.nodes, seed=seed0 + rank, pinned=True, skewed=args.skewed)
    P, C, N = host.n_pods, host.n_containers, host.n_nodes
    log(f"[rank {rank}] snapshot: P={P} C={C} N={N} generated in {time.time() - t_gen:.1f}s (NUMA node {numa})")
    total_pods = torch.tensor([P], dtype=torch.int64, device="cuda")
    max_pods = torch.tensor([P], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(total_pods)
        dist.all_reduce(max_pods, op=dist.ReduceOp.MAX)
    total_pods = int(total_pods.item())
    imbalance = int(max_pods.item()) * world / max(1, total_pods)      # 1.0 = every rank holds the same number of pods

    dev = ab.DeviceSnapshot.upload(host)
    d_total = torch.empty(N, dtype=torch.int64, device="cuda")
    d_alloc = torch.empty(N, dtype=torch.int64, device="cuda")
    d_cluster = torch.empty(2, dtype=torch.int64, device="cuda")
    # a stream of its own: the library replays the step from a CUDA graph, which the legacy default stream cannot do
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)

    fused = False
    if world > 1 and args.collective in ("fused", "posted"):
        fused = multigpu.setup_peer_exchange()
        log(f"[rank {rank}] in-kernel totals exchange over peer memory: {'on' if fused else 'unavailable, using NCCL'}")
        if not fused:
            args.collective = "nccl"
    posted = fused and args.collective == "posted"

    def step():
        if posted:
            ab.top_node_device_posted(dev, d_total, d_alloc, d_cluster, stream.cuda_stream)
        elif fused:
            ab.top_node_device_fused(dev, d_total, d_alloc, d_cluster, stream.cuda_stream)
        else:
            ab.top_node_device(dev, d_total, d_alloc, d_cluster, None, stream.cuda_stream)
            multigpu.all_reduce_totals(d_cluster)

    def finish_steps():
        # deferred collection: the last step's totals are fetched by one more (one-warp) launch
        if posted:
            ab.peer_exchange_collect(d_cluster, stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    finish_steps()
    barrier()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)

    # ---- timed region: exactly K steps, device-timed, max over ranks -----------
    L0 = ab.lib().arena_kernel_launches()
    # the kernel's own duration is sampled on every 8th step of the timed region (recording the events on
    # every step costs ~10 us of stream time per step, tools/kbench_step.py)
    ab._lib.check(ab.lib().arena_profile_begin_sampled(args.steps, 8))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    finish_steps()
    ev1.record(stream)
    barrier()
    t_wall1 = time.time()
    ms = ev0.elapsed_time(ev1)
    launches = int(ab.lib().arena_kernel_launches() - L0)
    import ctypes as C_
    nodes_ms, pods_ms, calls = C_.c_double(), C_.c_double(), C_.c_uint64()
    ab._lib.check(ab.lib().arena_profile_end(C_.byref(nodes_ms), C_.byref(pods_ms), C_.byref(calls)))
    tmax = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_total = float(tmax.item())
    ms_per_step = ms_total / args.steps
    value = total_pods * args.steps / (ms_total / 1e3)

    # the same step replayed from a CUDA graph (captured through the C ABI on this stream), reported beside the
    # headline, which is timed on plain launches
    graph_ms = None
    if args.graph and (world == 1 or fused):
        try:
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                def gstep():
                    if posted:
                        ab.top_node_device_posted(dev, d_total, d_alloc, d_cluster, side.cuda_stream)
                    elif fused:
                        ab.top_node_device_fused(dev, d_total, d_alloc, d_cluster, side.cuda_stream)
                    else:
                        ab.top_node_device(dev, d_total, d_alloc, d_cluster, None, side.cuda_stream)
                with torch.cuda.graph(g, stream=side):
                    gstep()
                for _ in range(args.warmup):
                    g.replay()
                barrier()
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                g0.record(side)
                for _ in range(args.steps):
                    g.replay()
                if posted:
                    ab.peer_exchange_collect(d_cluster, side.cuda_stream)
                g1.record(side)
                barrier()
                gms = torch.tensor([g0.elapsed_time(g1)], dtype=torch.float64, device="cuda")
            stream.wait_stream(side)
            if world > 1:
                dist.all_reduce(gms, op=dist.ReduceOp.MAX)
            graph_ms = float(gms.item()) / args.steps
        except Exception as e:                      # a side measurement: never lose the headline over it
            log(f"[rank {rank}] graph replay leg failed: {e!r}")
            graph_ms = None

    if fused and ab.lib().arena_peer_exchange_timed_out():
        raise SystemExit("bench.py: a peer did not post its totals in time (in-kernel exchange)")
    # sanity: table sums equal the totals this rank contributed (cheap; parity proper is tests/)
    cl = d_cluster.cpu().numpy()
    local_alloc_sum = int(d_alloc.sum().item())
    local_total_sum = int(d_total.sum().item())
    sums = torch.tensor([local_total_sum, local_alloc_sum], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(sums)
    if sums.cpu().numpy().tolist() != cl.tolist():
        raise SystemExit(f"bench.py: cluster totals {cl.tolist()} != table sums {sums.tolist()}")
    # parity proper, outside the timed region: THIS rank's slice of the table against the C oracle on the
    # same shard, and the exchanged global totals against the sum of the per-rank oracle totals
    ok_local, oracle_cl = oracle_parity(host, d_total, d_alloc)
    oc = torch.tensor(oracle_cl.tolist() + [0 if ok_local else 1], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(oc)
    oc = oc.cpu().numpy().tolist()
    parity_ok = oc[2] == 0 and oc[:2] == cl.tolist()
    if not parity_ok:
        raise SystemExit(f"bench.py: parity failure: {oc[2]} rank(s) differ from the oracle; totals {cl.tolist()} vs oracle {oc[:2]}")

    # ---- e2e: the public one-shot call with HOST buffers, copies in the timed region.  The snapshot goes over
    # PCIe in its packed form (1 byte per pod + 1 per container) when it has one; the wide form (9 + 9) is timed
    # beside it.
    e2e_steps = args.e2e_steps or min(args.steps, 10)

    # the per-node table lands in page-locked host arrays that the caller keeps between calls
    pin = lambda n: torch.zeros(max(n, 1), dtype=torch.int64).pin_memory().numpy()[:n]
    e2e_out = ab.TopNodeResult(pin(N), pin(N), pin(2), None)

    def e2e_leg(snapshot, steps):
        for _ in range(2):
            multigpu.top_node_sharded(snapshot, out=e2e_out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            _, totals = multigpu.top_node_sharded(snapshot, out=e2e_out)
        barrier()
        secs = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(secs, op=dist.ReduceOp.MAX)
        if totals.tolist() != cl.tolist():
            raise SystemExit(f"bench.py: e2e totals {totals.tolist()} != device-path totals {cl.tolist()}")
        return float(secs.item())

    wide_s = e2e_leg(host, max(2, e2e_steps // 2))
    wide_steps = max(2, e2e_steps // 2)
    h2d_wide = 9 * P + 4 + 9 * C + 9 * N     # columns uploaded per step (top node needs no job columns)
    d2h = 16 * N + 16
    pack_s = None
    try:
        t_pack = time.perf_counter()
        packed = host.pack(pinned=True)
        pack_s = time.perf_counter() - t_pack              # host time of arena_host_snapshot_pack, once per snapshot
    except ab.ArenaError as e:
        log(f"[rank {rank}] no packed form: {e}")
        packed = None
    flags = torch.tensor([1 if packed is not None else 0], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if int(flags.item()):
        e2e_s = e2e_leg(packed, e2e_steps)
        h2d, e2e_call = packed.nbytes, "arena_top_node_host_packed (packed upload, one kernel straight on the packed columns, download)"
    else:
        e2e_s, e2e_steps, h2d = wide_s, wide_steps, h2d_wide
        e2e_call = "arena_top_node_host (upload + kernels + download)"
    e2e_call += " + totals all-reduce"
    # ---- the packed columns RESIDENT on the device: the same summary from 0.2 GB instead of 2.3 GB.  Its own
    # roofline entry: the kernel reads other columns than the wide one, so its bytes are recomputed
    packed_roof = None
    if packed is not None and world == 1:
        pdev = ab.PackedDeviceSnapshot(packed)
        p_total, p_alloc, p_cluster = torch.empty_like(d_total), torch.empty_like(d_alloc), torch.empty_like(d_cluster)
        for _ in range(args.warmup):
            ab.top_node_packed_device(pdev, p_total, p_alloc, p_cluster, stream.cuda_stream)
        pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        pe0.record(stream)
        for _ in range(args.steps):
            ab.top_node_packed_device(pdev, p_total, p_alloc, p_cluster, stream.cuda_stream)
        pe1.record(stream)
        torch.cuda.synchronize()
        p_ms = pe0.elapsed_time(pe1) / args.steps
        if not (torch.equal(p_alloc, d_alloc) and torch.equal(p_total, d_total) and p_cluster.cpu().numpy().tolist() == cl.tolist()):
            raise SystemExit("bench.py: the packed kernel's table differs from the wide kernel's")
        peak_p, _ = measured_peak_hbm()
        packed_roof = {"bound": "hbm", "kernel": "arena::top_node_packed_kernel", "kernel_ms": p_ms,
                       "algorithmic_bytes": pdev.launch_bytes, "achieved": pdev.launch_bytes / (p_ms / 1e3) / 1e9,
                       "peak": peak_p, "unit": "GB/s", "frac": pdev.launch_bytes / (p_ms / 1e3) / 1e9 / peak_p,
                       "pod_records_per_s": P / (p_ms / 1e3),
                       "layout": "packed, device-resident (arena_snapshot_upload_packed): 1 B per pod, a nibble or byte per "
                                 "container, 8 B of offsets per node; Quantity.Value() folded at pack time"}
        pdev.close()
    t_wall2 = time.time()

    clocks = sampler.stop(t_wall0, t_wall2) if rank == 0 else None

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (pods kernel) ----------------------------
    peak, peak_src = measured_peak_hbm()
    # pod_node+phase+ctr_begin, ctr_val+ctr_scale, node capacities in, node_total out (SURVEY 8d per-unit
    # sizes); the 8N bytes of node_alloc are zeroed by the memset before the kernel and not counted here
    pods_bytes = 9 * P + 9 * C + 4 + 17 * N
    pods_ms_avg = pods_ms.value / max(1, calls.value)
    achieved = pods_bytes / (pods_ms_avg / 1e3) / 1e9
    step_bytes = host.algorithmic_bytes_top_node()
    roofline = {
        "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": None, "kernel": "arena::top_node_pods_kernel", "kernel_ms": pods_ms_avg,
        "algorithmic_bytes": pods_bytes, "peak_source": peak_src,
        "node_alloc_memset_ms": nodes_ms.value / max(1, calls.value),
        "step_algorithmic_bytes": step_bytes,
        "step_gbs": step_bytes / (ms_per_step / 1e3) / 1e9 * (1 if world == 1 else 1),
        "frac_of_nominal_8000": achieved / 8000.0,
    }
    # calibration: the same five columns read with plain coalesced loads, no arithmetic
    cal_gbs, cal_ms = C_.c_double(), C_.c_double()
    ab._lib.check(ab.lib().arena_calibrate_stream_read(dev.handle, 20, C_.byref(cal_gbs), C_.byref(cal_ms)))
    roofline["read_only_stream_gbs"] = cal_gbs.value
    roofline["frac_of_read_only_stream"] = achieved / cal_gbs.value
    pat_gbs, pat_ms = C_.c_double(), C_.c_double()
    ab._lib.check(ab.lib().arena_calibrate_stream_pattern(dev.handle, 20, C_.byref(pat_gbs), C_.byref(pat_ms)))
    roofline["tile_pattern_only_ms"] = pat_ms.value       # same access pattern, no arithmetic, no dependent loads
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            with open(tr) as f:
                roofline["traffic"] = json.load(f).get("pods_kernel_dram_bytes_per_launch")
        except Exception:
            pass

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling if world > 1 else "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "parity_ok": parity_ok,
        "config": dict(workload_config(n_gpus, args.scaling, args.skewed, args.collective),
                       l2=f"inputs ({step_bytes / 1e9:.2f} GB on this rank) exceed the 126 MB L2; no flush between steps",
                       pods_per_gpu=P, nodes_per_gpu=N,
                       containers_per_gpu=C, total_pods=total_pods, shard_imbalance_max_over_mean=imbalance),
        "clocks": clocks,
        "e2e": {"value": total_pods * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": e2e_s / e2e_steps * 1e3,
                "call": e2e_call,
                "pack_ms": pack_s * 1e3 if pack_s is not None and packed is not None else None,
                "pack_note": "host time of arena_host_snapshot_pack (wide columns -> packed form), once per snapshot, "
                             "outside the timed e2e steps; e2e_incl_pack adds it to EVERY step (a caller that repacks "
                             "per query)",
                "e2e_incl_pack": (total_pods / (e2e_s / e2e_steps + pack_s) if pack_s is not None and packed is not None
                                  and int(flags.item()) else None),
                "wide": {"value": total_pods * wide_steps / wide_s, "h2d_bytes_per_step": h2d_wide,
                         "ms_per_step": wide_s / wide_steps * 1e3,
                         "call": "arena_top_node_host (wide columns: 9 B per pod + 9 B per container)"}},
        "gpu_launches": launches,
        "graph_ms_per_step": graph_ms,
        "collective": EXCHANGES[args.collective] if world > 1 else None,
        "roofline": roofline,
        "roofline_packed": packed_roof,
        "hbm_gbs_step": step_bytes / (ms_per_step / 1e3) / 1e9,
    }

    if not args.no_cpu and world == 1:
        threads = os.cpu_count() or 1
        log("[rank 0] cpu_baseline leg (bounded sample)...")
        r = cpu_reference_run(steps=5, warmup=1, sample_pods=args.sample_pods, threads=threads, min_seconds=10.0)
        line["cpu_baseline"] = {
            "value": r["value"], "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"{r['passes']} passes ({r['seconds']:.1f} s) over the first {r['sample_pods']} pods / "
                      f"{r['sample_nodes']} nodes of the generator (seed 1), array-of-structs + string-hash join, all "
                      "host threads (C restatement of the Go path; the Go reference cannot be built here)",
            "columnar_single_pass_1thread": r["b3_columnar_1thread"],
            "columnar_all_threads": r["b3_columnar_all_threads"],
            "faithful_nested_loop_1thread": r["b1_nested_loop_1thread"],
            "faithful_nested_loop_sample": r["b1_sample"],
        }
    if not args.no_wire and world == 1:
        # the step in FRONT of the hot path, reported beside it: an API-server-shaped PodList of
        # BASELINE configs[1]'s size parsed by the host loader and by the GPU loader
        try:
            line["wire_format"] = wire_format_leg(ab)
        except Exception as e:                      # never lose the headline over the side measurement
            line["wire_format"] = {"error": repr(e)}
    sys.stdout.flush()
    os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
