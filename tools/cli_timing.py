"""Write config-2-sized JSON fixtures to /tmp and time the arena-top CLI on them (GPU box)."""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_fixtures as MF
for pods, nodes in [(10_000, 1_000), (200_000, 10_000)]:
    podlist, nodelist, joblist, tfjoblist, helm = MF.config2(n_nodes=nodes, n_pods_target=pods)
    d = f"/tmp/cfg_{pods}"
    os.makedirs(d, exist_ok=True)
    for name, obj in [("pods", podlist), ("nodes", nodelist), ("jobs", joblist), ("tfjobs", tfjoblist)]:
        json.dump(obj, open(f"{d}/{name}.json", "w"))
    open(f"{d}/helm.txt", "w").write(helm)
    sz = os.path.getsize(f"{d}/pods.json") / 1e6
    cli = os.path.join(ROOT, "arena_b200", "arena-top")
    for args in (["node"], ["job", "--jobs", f"{d}/jobs.json", "--tfjobs", f"{d}/tfjobs.json", "--releases", f"{d}/helm.txt", "--now", "1532304000"]):
        t = time.time()
        p = subprocess.run([cli, args[0], "--pods", f"{d}/pods.json", "--nodes", f"{d}/nodes.json", "--time"] + args[1:], capture_output=True, text=True)
        print(f"pods={len(podlist['items'])} ({sz:.1f} MB JSON) top {args[0]}: wall {time.time()-t:.2f}s rc={p.returncode} :: {p.stderr.strip()}  [{len(p.stdout.splitlines())} lines]")
