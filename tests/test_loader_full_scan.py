"""arena_load_json_gpu's host twin (arena_load_json_hostscan): the device scanner's source, compiled
for the host, reports per pod the accounting columns AND the places of the fields `top job`,
`top node -d` and `arena list` read; the loader's own record building and job rules then run on them.
The snapshot must be the one arena_load_json (loader.cc, pinned to the Python oracle by
tests/test_fixtures.py) builds from the same five documents: every column, row order, every name, the
job table, and the printed texts.  The GPU run of the same comparison is tests/test_gpu_json_device.py."""

import json
import os
import random
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
sys.path.insert(0, os.path.join(HERE, "golden"))
import loader_timing as LT  # noqa: E402
import make_fixtures as MF  # noqa: E402

D = os.path.join(HERE, "golden", "config1")


def rd(name):
    with open(os.path.join(D, name), "rb") as f:
        return f.read()


def same_snapshot(a, b):
    """Every observable of two HostSnapshots: columns, row order, names, jobs, printed texts."""
    from arena_b200 import TopJobResult, TopNodeResult, format_job_list, format_top_job, format_top_node
    from oracle import cref
    assert (a.n_pods, a.n_containers, a.n_nodes, a.n_jobs) == (b.n_pods, b.n_containers, b.n_nodes, b.n_jobs)
    for k in a.cols:
        assert np.array_equal(a.cols[k], b.cols[k]), k
    assert np.array_equal(a.pod_order(), b.pod_order())
    assert [a.pod_name(r) for r in range(a.n_pods)] == [b.pod_name(r) for r in range(b.n_pods)]
    assert [a.node_name(n) for n in range(a.n_nodes)] == [b.node_name(n) for n in range(b.n_nodes)]
    assert [(a.job_name(j), a.job_trainer(j)) for j in range(a.n_jobs)] == [(b.job_name(j), b.job_trainer(j)) for j in range(b.n_jobs)]
    texts = []
    for h in (a, b):
        tot, alloc, cl, pod_gpu = cref.top_node_columnar(h.cols, True)
        t = [format_top_node(h, TopNodeResult(tot, alloc, cl, pod_gpu), details=True)]          # namespaces, pod names
        if h.n_jobs:
            req, al, latest, totals = cref.top_job_columnar(h.cols, h.n_jobs)
            order = np.array(sorted(range(h.n_jobs), key=lambda j: (-int(req[j]), j)), np.uint32)
            res = TopJobResult(req, al, order, latest, totals)
            t.append(format_top_job(h, res, MF.NOW))                                                # status, age, node (host IP)
            t.append(format_job_list(h, res, MF.NOW))
        texts.append(t)
    assert texts[0] == texts[1]


def both(pods, nodes, jobs=None, tfjobs=None, helm=None, by_job=False):
    from arena_b200 import HostSnapshot
    want = HostSnapshot.from_json(pods, nodes, jobs, tfjobs, helm, group_by_job=by_job)
    got, handed = HostSnapshot.from_json_hostscan(pods, nodes, jobs, tfjobs, helm, group_by_job=by_job)
    same_snapshot(got, want)
    return handed, want


def test_config1_fixture_all_documents(built):
    for by_job in (False, True):
        handed, h = both(rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"), rd("helm_list.txt"), by_job)
        assert h.n_jobs > 0
        # the fixture carries deliberately exotic items (escaped strings, odd quantities): some go through the
        # host's item code, most must not
        assert handed < h.n_pods // 2, handed


def test_api_server_shaped_pods_with_jobs(built):
    """10 releases x 4 pods, Job-owned, the shape tools/loader_timing.py gives the API server's output."""
    rng = random.Random(7)
    n_pods, n_nodes = 200, 12
    pods = json.dumps({"kind": "PodList", "items": [LT.make_pod(i, rng, n_nodes) for i in range(n_pods)]}).encode()
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(n_nodes)]}).encode()
    jobs = json.dumps({"items": [
        {"metadata": {"name": f"job-{j:05d}", "namespace": "default", "labels": {"release": f"job-{j:05d}", "app": "tf-horovod"}},
         "status": {"active": 1, "startTime": "2024-03-01T00:00:00Z"}} for j in range(0, n_pods // 4, 2)]}).encode()
    helm = ("NAME\tREVISION\tUPDATED\tSTATUS\tCHART\tNAMESPACE\n" +
            "".join(f"job-{j:05d}\t1\tx\tDEPLOYED\ttf-horovod-0.1.0\tdefault\n" for j in range(n_pods // 4))).encode()
    handed, h = both(pods, nodes, jobs, None, helm)
    assert handed == 0 and h.n_jobs == n_pods // 8
    assert int((h.cols["pod_job"] != 0xFFFFFFFF).sum()) == n_pods // 2


LABELS = ["release", "app", "group_name", "tf-replica-type", "tf-replica-index", "other"]
LABEL_VALUES = ["r0", "r1", "tf-horovod", "training", "tfjob", "kubeflow.org", "worker", "0", "1", "", None, "x\"y", "é", 5]


def mutate_meta(pod, rng):
    """Spellings around the fields the job rules read."""
    m = pod["metadata"]
    r = rng.random()
    if r < 0.2:
        m["labels"] = {k: rng.choice(LABEL_VALUES) for k in rng.sample(LABELS, rng.randrange(0, 6))}
    elif r < 0.25:
        m["labels"] = rng.choice([None, {}, [], "x"])
    r = rng.random()
    if r < 0.15:
        m["ownerReferences"] = [{"kind": rng.choice(["Job", "ReplicaSet", "job", "", None, "Jo\\u0062"]), "name": "o"}
                                for _ in range(rng.randrange(0, 3))]
    elif r < 0.2:
        m["ownerReferences"] = rng.choice([None, [], [None], {}, [5]])
    r = rng.random()
    if r < 0.1:
        m["creationTimestamp"] = rng.choice([None, "2024-02-29T23:59:59Z", "2023-02-29T00:00:00Z", "2024-03-01T00:00:00+08:00",
                                             "2024-03-01T00:00:00.5Z", "", 5, "2024-03-01t00:00:00z"])
    r = rng.random()
    if r < 0.1:
        m["name"] = rng.choice([None, "", "p\"q", "näme", 7])
    if rng.random() < 0.1:
        m["namespace"] = rng.choice([None, "", "other", "default"])
    if rng.random() < 0.1:
        pod["status"]["hostIP"] = rng.choice([None, "", "10.0.0.1", 5, "1\\u0030.0.0.2"])
    if rng.random() < 0.05:
        pod["metadata"] = rng.choice([None, m])


def test_fuzz_job_fields_against_host_loader(built):
    from arena_b200 import ArenaError, HostSnapshot
    rng = random.Random(424242)
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(5)]}).encode()
    helm = b"NAME\tNAMESPACE\nr0\t1\tdefault\nr1\t1\tdefault\njob-00000\t1\tdefault\n"
    jobs = json.dumps({"items": [
        {"metadata": {"namespace": "default", "labels": {"release": "r0", "app": "tf-horovod"}}, "status": {"active": 2}},
        {"metadata": {"namespace": "default", "labels": {"release": "r1", "app": "training"}}, "status": {"succeeded": 1}},
        {"metadata": {"namespace": "default", "labels": {"release": "job-00000", "app": "tf-horovod"}}, "status": {}}]}).encode()
    tfjobs = json.dumps({"items": []}).encode()
    seen = {"same": 0, "host_err": 0, "handed": 0}
    for trial in range(600):
        items = []
        for i in range(rng.randrange(1, 7)):
            pod = LT.make_pod(rng.randrange(8), rng, 6)
            pod["metadata"]["managedFields"] = []
            mutate_meta(pod, rng)
            items.append(pod)
        text = json.dumps({"kind": "PodList", "items": items}, ensure_ascii=rng.random() < 0.5).replace("\\\\u00", "\\u00").encode()
        by_job = rng.random() < 0.3
        try:
            want = HostSnapshot.from_json(text, nodes, jobs, tfjobs, helm, group_by_job=by_job)
        except ArenaError:
            seen["host_err"] += 1
            with pytest.raises(ArenaError):            # the twin reports "not taken"; arena_load_json_gpu then runs the host loader
                HostSnapshot.from_json_hostscan(text, nodes, jobs, tfjobs, helm, group_by_job=by_job)
            continue
        got, handed = HostSnapshot.from_json_hostscan(text, nodes, jobs, tfjobs, helm, group_by_job=by_job)
        same_snapshot(got, want)
        seen["same"] += 1
        seen["handed"] += 1 if handed else 0
    assert seen["same"] > 200 and seen["host_err"] > 20 and seen["handed"] > 20, seen


def test_canonical_time_matches_the_loader(built):
    """creationTimestamp is converted on the host from the place the scanner reports: leap years, month ends."""
    from arena_b200 import HostSnapshot
    nodes = json.dumps({"items": [LT.make_node(0)]}).encode()
    rng = random.Random(3)
    stamps = ["0001-01-01T00:00:00Z", "1970-01-01T00:00:00Z", "2000-02-29T12:34:56Z", "2100-03-01T00:00:00Z", "9999-12-31T23:59:59Z"]
    stamps += [f"{rng.randrange(1, 9999):04d}-{rng.randrange(1, 13):02d}-{rng.randrange(1, 29):02d}T{rng.randrange(24):02d}:{rng.randrange(60):02d}:{rng.randrange(60):02d}Z"
               for _ in range(200)]
    items = []
    for i, ts in enumerate(stamps):
        pod = LT.make_pod(i, rng, 1)
        pod["metadata"]["managedFields"] = []
        pod["metadata"]["creationTimestamp"] = ts
        items.append(pod)
    text = json.dumps({"items": items}).encode()
    want = HostSnapshot.from_json(text, nodes)
    got, handed = HostSnapshot.from_json_hostscan(text, nodes)
    assert handed == 0
    assert np.array_equal(got.cols["pod_created"], want.cols["pod_created"])
