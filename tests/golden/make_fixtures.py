#!/usr/bin/env python
"""Writes the JSON fixtures of BASELINE configs[0] (16 nodes x 8 GPUs, 200 pods)
and configs[1] (1e4 TFJob/Horovod pods on 1000 nodes) plus the expected outputs.

The reference (Go) cannot be run here, so the expected tables / texts are
produced by the Python oracle (oracle/top_ref.py) and, for every edge case of
SURVEY.md section 8c, ALSO written down by hand below (HAND_NODE_TABLE,
HAND_JOB_TABLE): tests/test_fixtures.py checks oracle == hand values first and
only then product == oracle.

    python tests/golden/make_fixtures.py        # rewrites tests/golden/*.json|txt
"""

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

G = "nvidia.com/gpu"
D = "alpha.kubernetes.io/nvidia-gpu"
NOW = 1532304000 + 3 * 3600          # 2018-07-23T03:00:00Z


def node(name, capacity=None, allocatable=None, addresses=None, labels=None):
    st = {}
    if capacity is not None:
        st["capacity"] = dict({"cpu": "32", "memory": "131072Ki"}, **capacity)
    if allocatable is not None:
        st["allocatable"] = allocatable
    if addresses is not None:
        st["addresses"] = [{"type": t, "address": a} for t, a in addresses]
    md = {"name": name}
    if labels is not None:
        md["labels"] = labels
    return {"metadata": md, "status": st}


def ctr(limits=None, requests=None, name="c"):
    res = {}
    if limits is not None:
        res["limits"] = limits
    if requests is not None:
        res["requests"] = requests
    return {"name": name, "image": "img", "resources": res}


def pod(name, node_name, phase, containers, ns="default", labels=None, owner=None, created="2018-07-23T00:00:00Z",
        init=None, host_ip=None):
    md = {"name": name, "namespace": ns, "creationTimestamp": created}
    if labels:
        md["labels"] = labels
    if owner:
        md["ownerReferences"] = [{"kind": owner, "name": "o"}]
    spec = {"containers": containers}
    if node_name is not None:
        spec["nodeName"] = node_name
    if init:
        spec["initContainers"] = init
    st = {}
    if phase is not None:
        st["phase"] = phase
    if host_ip:
        st["hostIP"] = host_ip
    return {"metadata": md, "spec": spec, "status": st}


def config1():
    ip = lambda i: f"10.0.0.{10 + i}"
    nodes = [
        node("node-00", {G: "8"}, None, [("InternalIP", ip(0)), ("Hostname", "node-00")],
             {"node-role.kubernetes.io/master": ""}),                                    # master, label value empty
        node("node-01", {G: "8"}, None, [("Hostname", "node-01"), ("InternalIP", ip(1))]),   # summary/details differ
        node("node-02", {}, {G: "8"}, [("InternalIP", ip(2))]),                           # only allocatable => 0 (F3)
        node("node-03", {D: "4"}, None, [("InternalIP", ip(3))]),                         # deprecated capacity key
        node("node-04", {G: "8"}, None, None),                                            # no address => "unknown"
        node("node-05", {G: "0", D: "8"}, None, [("InternalIP", ip(5))]),                 # primary present wins even "0"
        node("node-06", {G: None, D: "8"}, None, [("ExternalIP", "1.2.3.4")]),            # null: key present, value 0
        node("node-07", {G: "8"}, None, [("InternalIP", ip(7))]),                         # GPU sharing: over-subscribed
    ] + [node(f"node-{i:02d}", {G: "8"}, {G: "8"}, [("InternalIP", ip(i))]) for i in range(8, 16)]

    pods = []
    add = pods.append
    # --- node-08: container/key selection edge cases (all Running) -> 4 + 2 + 0 + 0 + 1 = 7
    add(pod("three-ctrs", "node-08", "Running", [ctr({G: "1"}), ctr({G: "2"}), ctr({G: "500m"})]))      # 1+2+1
    add(pod("deprecated-only", "node-08", "Running", [ctr({D: "2"})]))                                  # 2
    add(pod("both-keys", "node-08", "Running", [ctr({G: "0", D: "4"})]))                                # 0: primary wins
    add(pod("requests-only", "node-08", "Running", [ctr(None, {G: "4"})]))                              # 0: limits (F2)
    add(pod("init-gpu", "node-08", "Running", [ctr({G: "1"})], init=[ctr({G: "8"})]))                   # 1: init ignored
    # --- node-09: one 1-GPU pod per phase -> "", Pending, Running, Unknown, Weird are active = 5
    for ph in [None, "Pending", "Running", "Succeeded", "Failed", "Unknown", "Weird"]:
        add(pod(f"phase-{ph or 'empty'}".lower(), "node-09", ph, [ctr({G: "1"})]))
    # --- nowhere: unscheduled, unknown node
    add(pod("unscheduled", None, "Pending", [ctr({G: "2"})]))
    add(pod("empty-node-name", "", "Pending", [ctr({G: "2"})]))
    add(pod("ghost", "ghost-node", "Running", [ctr({G: "2"})]))
    # --- node-10: quantity forms: 2 + 1000 + 1000 + 1024 + 1 + 0 + 0 = 3027
    add(pod("unquoted", "node-10", "Running", [{"name": "c", "resources": {"limits": {G: 2}}}]))        # bare number
    add(pod("kilo", "node-10", "Running", [ctr({G: "1k"})]))
    add(pod("exp", "node-10", "Running", [ctr({G: "1e3"})]))
    add(pod("kibi", "node-10", "Running", [ctr({G: "1Ki"})]))
    add(pod("half", "node-10", "Running", [ctr({G: "0.5"})]))
    add(pod("null-limit", "node-10", "Running", [ctr({G: None, D: "3"})]))                              # null => 0, no fallback
    add(pod("overflow", "node-10", "Running", [ctr({G: "10E"})]))                                       # Value() overflow => 0
    # --- node-11: negative limits: 3 + (-1) + (-1) = 1
    add(pod("plus3", "node-11", "Running", [ctr({G: "3"})]))
    add(pod("minus1", "node-11", "Running", [ctr({G: "-1"})]))
    add(pod("minus-half", "node-11", "Running", [ctr({G: "-500m"})]))
    # --- node-07: 30 pods of "500m" (GPU sharing) -> 30 allocated on 8
    for i in range(30):
        add(pod(f"share-{i:02d}", "node-07", "Running", [ctr({G: "500m" if i % 2 else "250m"})]))
    # --- jobs (also used by `top job`) ------------------------------------------------
    tf = lambda typ, idx: {"app": "tfjob", "release": "tf-dist", "group_name": "kubeflow.org",
                           "tf-replica-type": typ, "tf-replica-index": str(idx)}
    add(pod("tf-dist-ps-0", "node-12", "Running", [ctr({})], labels=tf("ps", 0), host_ip=ip(12)))
    add(pod("tf-dist-worker-0", "node-12", "Running", [ctr({G: "1"})], labels=tf("worker", 0), host_ip=ip(12)))
    add(pod("tf-dist-worker-1", "node-13", "Succeeded", [ctr({G: "1"})], labels=tf("worker", 1), host_ip=ip(13)))
    add(pod("tf-dist-worker-2", None, "Pending", [ctr({G: "1"})], labels=tf("worker", 2)))
    add(pod("tf-dist-stray", "node-13", "Running", [ctr({G: "4"})],
            labels={"app": "tfjob", "release": "tf-dist"}))                                              # no group_name: not a job pod
    hv = lambda role: {"app": "tf-horovod", "release": "hvd", "role": role}
    add(pod("hvd-launcher-old", "node-14", "Failed", [ctr({G: "1"})], labels=hv("mpimaster"), owner="Job",
            created="2018-07-23T00:00:10Z", host_ip=ip(14)))
    add(pod("hvd-launcher-new", "node-14", "Running", [ctr({G: "2"})], labels=hv("mpimaster"), owner="Job",
            created="2018-07-23T00:05:00Z", host_ip=ip(14)))
    add(pod("hvd-0", "node-14", "Running", [ctr({G: "2"})], labels=hv("mpiworker"), owner="StatefulSet"))
    add(pod("hvd-1", "node-15", "Running", [ctr({G: "2"})], labels=hv("mpiworker"), owner="StatefulSet"))
    add(pod("hvd-jobmon", "node-15", "Running", [ctr({G: "8"})], ns="arena-system", labels=hv("jobmon"), owner="Job"))
    ht = {"app": "tf-horovod", "release": "hvd-tie"}
    add(pod("tie-a", "node-15", "Running", [ctr({G: "1"})], labels=ht, owner="Job", created="2018-07-23T01:00:00Z",
            host_ip=ip(15)))
    add(pod("tie-b", "node-12", "Running", [ctr({G: "3"})], labels=ht, owner="Job", created="2018-07-23T01:00:00Z",
            host_ip=ip(12)))                                                                             # equal time: tie-a wins
    so = {"app": "training", "release": "solo"}
    add(pod("solo-try1", "node-13", "Failed", [ctr({G: "4"})], labels=so, owner="Job", created="2018-07-22T23:00:00Z"))
    add(pod("solo-try2", "node-13", "Pending", [ctr({G: "4"})], labels=so, owner="Job", created="2018-07-22T23:30:00+00:00",
            host_ip=ip(13)))
    # --- filler: 1-GPU running pods on nodes 12..15 and 00..01, up to 200 pods
    i = 0
    while len(pods) < 200:
        n = [12, 13, 14, 15, 0, 1][i % 6]
        add(pod(f"filler-{i:03d}", f"node-{n:02d}", "Running" if i % 5 else "Succeeded", [ctr({G: "1"}), ctr({})]))
        i += 1
    podlist = {"kind": "PodList", "apiVersion": "v1", "items": pods}
    nodelist = {"kind": "NodeList", "apiVersion": "v1", "items": nodes}

    job = lambda name, ns, app, rel, status: {"metadata": {"name": name, "namespace": ns, "labels": {"app": app, "release": rel}},
                                              "status": status}
    joblist = {"kind": "JobList", "items": [
        job("hvd-jobmon", "arena-system", "tf-horovod", "hvd", {"active": 1, "startTime": "2018-07-23T00:00:00Z"}),
        job("hvd", "default", "tf-horovod", "hvd", {"active": 1, "startTime": "2018-07-23T00:00:05Z"}),
        job("hvd-tie", "default", "tf-horovod", "hvd-tie", {"active": 1, "startTime": "2018-07-23T01:00:00Z"}),
        job("solo", "default", "training", "solo", {"active": 1, "failed": 1, "startTime": "2018-07-22T23:00:00Z"}),
    ]}
    tfjoblist = {"kind": "TFJobList", "items": [
        {"metadata": {"name": "tf-dist", "namespace": "default", "labels": {"app": "tfjob", "release": "tf-dist"}},
         "status": {"startTime": "2018-07-23T02:59:43Z",
                    "conditions": [{"type": "Created", "status": "False"}, {"type": "Running", "status": "True"}]}},
    ]}
    helm = ("NAME       \tREVISION\tUPDATED                 \tSTATUS  \tCHART           \tNAMESPACE\n"
            "hvd        \t1       \tMon Jul 23 00:00:00 2018\tDEPLOYED\ttf-horovod-0.1.0\tdefault  \n"
            "hvd-tie    \t1       \tMon Jul 23 01:00:00 2018\tDEPLOYED\ttf-horovod-0.1.0\tdefault  \n"
            "solo       \t1       \tSun Jul 22 23:00:00 2018\tDEPLOYED\ttraining-0.1.0  \tdefault  \n"
            "tf-dist    \t1       \tMon Jul 23 02:59:00 2018\tDEPLOYED\ttfjob-0.1.0     \tdefault  \n"
            "other-chart\t1       \tMon Jul 23 02:59:00 2018\tDEPLOYED\tnginx-1.0.0     \tdefault  \n")
    return podlist, nodelist, joblist, tfjoblist, helm


# Hand-derived expectations for config 1 (see the comments in config1()).
# filler pods: i = 0..; node cycle [12,13,14,15,0,1]; Succeeded when i % 5 == 0.
HAND_NODE_TABLE_EDGE = {
    # name: (total, allocated)
    "node-02": (0, 0), "node-03": (4, 0), "node-05": (0, 0), "node-06": (0, 0),
    "node-07": (8, 30), "node-08": (8, 7), "node-09": (8, 5), "node-10": (8, 3027), "node-11": (8, 1),
}
HAND_JOB_TABLE = {
    # name: (trainer, requested, allocated)
    "tf-dist": ("tfjob", 3, 2),        # ps 0 + workers 1+1+1; worker-1 Succeeded => allocated 2
    "hvd": ("mpijob", 6, 6),           # workers 2+2 + newest launcher (2, Running); old launcher ignored
    "hvd-tie": ("mpijob", 1, 1),       # equal timestamps: first in list (tie-a, 1 GPU) wins
    "solo": ("standalonejob", 4, 4),   # only the newest pod (try2, Pending => active)
}


def config2(seed=20180723, n_nodes=1000, n_pods_target=10_000):
    """Generated: ~1000 jobs (70% tfjob: 1 PS + 2-16 workers; 30% horovod: 1 launcher
    + 1-15 workers; 5% of launchers retried once); worker limit in {1,2,4,8};
    phases Running 70 / Pending 10 / Succeeded 12 / Failed 6 / Unknown 2 %."""
    import random
    rnd = random.Random(seed)
    nodes = [node(f"n{i:04d}", {G: "8"}, {G: "8"}, [("InternalIP", f"10.{i // 250}.{i % 250}.1")]) for i in range(n_nodes)]
    phases = ["Running"] * 70 + ["Pending"] * 10 + ["Succeeded"] * 12 + ["Failed"] * 6 + ["Unknown"] * 2
    pods, jobs, tfjobs, helm_lines = [], [], [], ["NAME\tREVISION\tUPDATED\tSTATUS\tCHART\tNAMESPACE"]
    j = 0
    while len(pods) < n_pods_target:
        name = f"job-{j:04d}"
        ns = "default" if j % 3 else "team-a"
        gpu = str(rnd.choice([1, 2, 4, 8]))
        t0 = 1532304000 - rnd.randrange(0, 86400)
        ts = lambda dt=0: __import__("time").strftime("%Y-%m-%dT%H:%M:%SZ", __import__("time").gmtime(t0 + dt))
        pick_node = lambda: f"n{rnd.randrange(n_nodes):04d}"
        if rnd.random() < 0.7:
            lab = lambda typ, idx: {"app": "tfjob", "release": name, "group_name": "kubeflow.org",
                                    "tf-replica-type": typ, "tf-replica-index": str(idx)}
            pods.append(pod(f"{name}-ps-0", pick_node(), rnd.choice(phases), [ctr({})], ns=ns, labels=lab("ps", 0),
                            created=ts(), host_ip="10.9.9.9"))
            for w in range(rnd.randrange(2, 17)):
                pods.append(pod(f"{name}-worker-{w}", pick_node(), rnd.choice(phases), [ctr({G: gpu}), ctr({})], ns=ns,
                                labels=lab("worker", w), created=ts(), host_ip="10.9.9.9"))
            cond = rnd.choice([[("Running", "True")], [("Created", "True")], [("Succeeded", "True")],
                               [("Failed", "True")], [("Created", "False"), ("Running", "True")]])
            tfjobs.append({"metadata": {"name": name, "namespace": ns, "labels": {"app": "tfjob", "release": name}},
                           "status": {"startTime": ts(), "conditions": [{"type": t, "status": s} for t, s in cond]}})
            chart = "tfjob-0.1.0"
        else:
            lab = lambda role: {"app": "tf-horovod", "release": name, "role": role}
            pods.append(pod(f"{name}-launcher", pick_node(), rnd.choice(phases), [ctr({G: gpu})], ns=ns,
                            labels=lab("mpimaster"), owner="Job", created=ts(), host_ip="10.8.8.8"))
            if rnd.random() < 0.05:
                dt = rnd.choice([0, 0, 30])          # equal timestamps happen (second resolution)
                pods.append(pod(f"{name}-launcher-retry", pick_node(), rnd.choice(phases), [ctr({G: gpu}), ctr({G: "1"})],
                                ns=ns, labels=lab("mpimaster"), owner="Job", created=ts(dt), host_ip="10.8.8.7"))
            for w in range(rnd.randrange(1, 16)):
                pods.append(pod(f"{name}-{w}", pick_node(), rnd.choice(phases), [ctr({G: gpu})], ns=ns,
                                labels=lab("mpiworker"), owner="StatefulSet", created=ts()))
            jobs.append({"metadata": {"name": name, "namespace": ns, "labels": {"app": "tf-horovod", "release": name}},
                         "status": dict(rnd.choice([{"active": 1}, {"succeeded": 1}, {"failed": 1}]), startTime=ts())})
            chart = "tf-horovod-0.1.0"
        helm_lines.append(f"{name}\t1\tMon Jul 23 00:00:00 2018\tDEPLOYED\t{chart}\t{ns}")
        j += 1
    rnd.shuffle(pods)                                 # PodList order is arbitrary
    return ({"kind": "PodList", "items": pods}, {"kind": "NodeList", "items": nodes}, {"kind": "JobList", "items": jobs},
            {"kind": "TFJobList", "items": tfjobs}, "\n".join(helm_lines) + "\n")


def dump(obj, path):
    with open(path, "w") as f:
        json.dump(obj, f, indent=None, separators=(",", ":"))
        f.write("\n")


def main():
    from oracle import top_ref
    podlist, nodelist, joblist, tfjoblist, helm = config1()
    d = os.path.join(HERE, "config1")
    os.makedirs(d, exist_ok=True)
    dump(podlist, os.path.join(d, "pods.json"))
    dump(nodelist, os.path.join(d, "nodes.json"))
    dump(joblist, os.path.join(d, "jobs.json"))
    dump(tfjoblist, os.path.join(d, "tfjobs.json"))
    open(os.path.join(d, "helm_list.txt"), "w").write(helm)
    P, N = json.dumps(podlist), json.dumps(nodelist)
    text, table = top_ref.top_node(P, N, details=False)
    open(os.path.join(d, "expected_top_node.txt"), "w").write(text)
    dump([list(r) for r in table], os.path.join(d, "expected_node_table.json"))
    text_d, _ = top_ref.top_node(P, N, details=True)
    open(os.path.join(d, "expected_top_node_details.txt"), "w").write(text_d)
    jtext, rows = top_ref.top_job(P, json.dumps(joblist), json.dumps(tfjoblist), helm, NOW)
    open(os.path.join(d, "expected_top_job.txt"), "w").write(jtext)
    dump([list(r) for r in rows], os.path.join(d, "expected_job_table.json"))
    ltext, _ = top_ref.job_list(P, json.dumps(joblist), json.dumps(tfjoblist), helm, NOW)
    open(os.path.join(d, "expected_list.txt"), "w").write(ltext)
    print(text)
    print(jtext)
    return 0


if __name__ == "__main__":
    sys.exit(main())
