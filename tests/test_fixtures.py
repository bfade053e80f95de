"""BASELINE configs[0] fixture (16 nodes x 8 GPUs, 200 pods, every edge case
of SURVEY.md 8c) without a GPU: oracle vs hand-derived values, then the
PRODUCT's loader (JSON -> columns) and printers against the oracle."""

import json
import os
import sys

import numpy as np
import pytest

from oracle import cref, top_ref

HERE = os.path.dirname(os.path.abspath(__file__))
D = os.path.join(HERE, "golden", "config1")
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_fixtures as MF  # noqa: E402


def rd(name, mode="rb"):
    with open(os.path.join(D, name), mode) as f:
        return f.read()


def strip_trailing(text: str) -> str:
    return "\n".join(ln.rstrip() for ln in text.split("\n"))


# ---- the oracle itself ---------------------------------------------------------------
def test_fixture_files_are_what_the_generator_writes():
    podlist, nodelist, joblist, tfjoblist, helm = MF.config1()
    assert json.loads(rd("pods.json")) == podlist and json.loads(rd("nodes.json")) == nodelist
    assert json.loads(rd("jobs.json")) == joblist and json.loads(rd("tfjobs.json")) == tfjoblist
    assert rd("helm_list.txt", "r") == helm
    assert len(podlist["items"]) == 200 and len(nodelist["items"]) == 16


def test_oracle_matches_hand_derived_node_values():
    text, table = top_ref.top_node(rd("pods.json"), rd("nodes.json"))
    got = {name: (t, a) for name, t, a in table}
    for name, want in MF.HAND_NODE_TABLE_EDGE.items():
        assert got[name] == want, name
    assert text == rd("expected_top_node.txt", "r")
    assert [list(r) for r in table] == json.loads(rd("expected_node_table.json"))
    # summary vs details pick different addresses (top_node.go:120-129 vs 172-175)
    details, _ = top_ref.top_node(rd("pods.json"), rd("nodes.json"), details=True)
    assert details == rd("expected_top_node_details.txt", "r")
    assert "node-01  10.0.0.11" in text and "IPADDRESS:  node-01" in details
    assert "node-04  unknown" in text and "node-06  unknown" in text and "IPADDRESS:  1.2.3.4" in details


def test_oracle_matches_hand_derived_job_values():
    text, rows = top_ref.top_job(rd("pods.json"), rd("jobs.json"), rd("tfjobs.json"), rd("helm_list.txt", "r"), MF.NOW)
    got = {name: (trainer, req, alloc) for name, trainer, req, alloc in rows}
    assert got == MF.HAND_JOB_TABLE
    assert [r[0] for r in rows] == ["hvd", "solo", "tf-dist", "hvd-tie"]         # requested desc
    assert text == rd("expected_top_job.txt", "r")


def test_oracle_layout_matches_the_reference_doc_sample():
    """docs/userguide/1-tfjob-standalone.md:88-98 (6 nodes, 1/3 GPUs): the only
    output sample the reference ships; pins the tabwriter layout (the docs have
    trailing blanks trimmed)."""
    sample = """NAME                    IPADDRESS      ROLE    GPU(Total)  GPU(Allocated)
i-j6c68vrtpvj708d9x6j0  192.168.1.116  master  0           0
i-j6c8ef8d9sqhsy950x7x  192.168.1.119  worker  1           0
i-j6c8ef8d9sqhsy950x7y  192.168.1.120  worker  1           1
i-j6c8ef8d9sqhsy950x7z  192.168.1.118  worker  1           0
i-j6ccue91mx9n2qav7qsm  192.168.1.115  master  0           0
i-j6ce09gzdig6cfcy1lwr  192.168.1.117  master  0           0
-----------------------------------------------------------------------------------------
Allocated/Total GPUs In Cluster:
1/3 (33%)
"""
    names = ["i-j6c68vrtpvj708d9x6j0", "i-j6c8ef8d9sqhsy950x7x", "i-j6c8ef8d9sqhsy950x7y", "i-j6c8ef8d9sqhsy950x7z",
             "i-j6ccue91mx9n2qav7qsm", "i-j6ce09gzdig6cfcy1lwr"]
    ips = [116, 119, 120, 118, 115, 117]
    master = [1, 0, 0, 0, 1, 1]
    nodes = [MF.node(n, {MF.G: "0" if m else "1"}, None, [("InternalIP", f"192.168.1.{ip}")],
                     {"node-role.kubernetes.io/master": ""} if m else None) for n, ip, m in zip(names, ips, master)]
    pods = [MF.pod("tf-git-tfjob-worker-0", names[2], "Running", [MF.ctr({MF.G: "1"})])]
    text, _ = top_ref.top_node(json.dumps({"items": pods}), json.dumps({"items": nodes}))
    assert strip_trailing(text) == sample
    # and the job sample, docs/userguide/1-tfjob-standalone.md:73-82
    job_sample = """NAME    STATUS   TRAINER  AGE  NODE           GPU(Requests)  GPU(Allocated)
tf-git  RUNNING  TFJOB    17s  192.168.1.120  1              1


Total Allocated GPUs of Training Job:
1

Total Requested GPUs of Training Job:
1
"""
    lab = {"app": "tfjob", "release": "tf-git", "group_name": "kubeflow.org", "tf-replica-type": "worker",
           "tf-replica-index": "0"}
    pods = [MF.pod("tf-git-tfjob-worker-0", names[2], "Running", [MF.ctr({MF.G: "1"})], labels=lab, host_ip="192.168.1.120")]
    tfjobs = {"items": [{"metadata": {"name": "tf-git-tfjob", "namespace": "default", "labels": {"app": "tfjob", "release": "tf-git"}},
                         "status": {"startTime": "2018-07-23T00:00:00Z", "conditions": [{"type": "Running", "status": "True"}]}}]}
    helm = "NAME  \tREVISION\tUPDATED\tSTATUS\tCHART\tNAMESPACE\ntf-git\t1\tMon\tDEPLOYED\ttfjob-0.1.0\tdefault\n"
    jtext, _ = top_ref.top_job(json.dumps({"items": pods}), None, json.dumps(tfjobs), helm, 1532304000 + 17)
    assert strip_trailing(jtext) == job_sample


# ---- the product's host side (no GPU): loader and printers -------------------------------
@pytest.fixture(scope="module")
def host_node(built):
    from arena_b200 import HostSnapshot
    return HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"))


def test_loader_columns_give_the_expected_node_table(host_node):
    h = host_node
    assert (h.n_pods, h.n_nodes) == (200, 16)
    assert h.n_containers == sum(len(p["spec"]["containers"]) for p in json.loads(rd("pods.json"))["items"])
    pn = h.cols["pod_node"]
    assert np.all(pn[:-1] <= pn[1:]), "rows must be grouped by node"
    order = h.pod_order()
    assert sorted(order.tolist()) == list(range(200))
    for n in range(16):                                   # stable: list order kept inside a node
        rows = np.nonzero(pn == n)[0]
        assert np.all(np.diff(order[rows].astype(np.int64)) > 0)
    tot, alloc, cl, pod_gpu = cref.top_node_columnar(h.cols, True)
    want = json.loads(rd("expected_node_table.json"))
    assert [[h.node_name(n), int(tot[n]), int(alloc[n])] for n in range(16)] == want
    assert cl.tolist() == [100, 3201]


def test_printers_reproduce_the_expected_text(host_node):
    from arena_b200 import TopNodeResult, format_top_node
    h = host_node
    tot, alloc, cl, pod_gpu = cref.top_node_columnar(h.cols, True)
    res = TopNodeResult(tot, alloc, cl, pod_gpu)
    assert format_top_node(h, res, details=False) == rd("expected_top_node.txt", "r")
    assert format_top_node(h, res, details=True) == rd("expected_top_node_details.txt", "r")
    assert res.usage_percent() == 3201


def test_loader_and_printer_for_top_job(built):
    from arena_b200 import HostSnapshot, TopJobResult, format_top_job
    h = HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"), rd("helm_list.txt"),
                               group_by_job=True)
    J = h.n_jobs
    names = [h.job_name(j) for j in range(J)]
    assert names == sorted(MF.HAND_JOB_TABLE)             # ids in name order; "other-chart" has no row
    pj = h.cols["pod_job"]
    assert np.all(pj[:-1] <= pj[1:]), "rows must be grouped by job"
    req, alloc, latest, totals = cref.top_job_columnar(h.cols, J)
    got = {names[j]: (h.job_trainer(j), int(req[j]), int(alloc[j])) for j in range(J)}
    assert got == MF.HAND_JOB_TABLE
    order = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
    text = format_top_job(h, TopJobResult(req, alloc, order, latest, totals), MF.NOW)
    assert text == rd("expected_top_job.txt", "r")
    assert totals.tolist() == [14, 13]


def test_timestamp_tie_survives_node_grouping(built):
    """Two Job-owned launcher pods of `hvd-tie` carry the same creationTimestamp and sit on
    different nodes, the later one in the list on the LOWER node id.  `latest.Before(current)` is
    strict (trainer_mpi.go:263-278), so the first in LIST order counts; grouping the rows by node
    must not change the winner."""
    from arena_b200 import HostSnapshot, TopJobResult, format_top_job
    h = HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"), rd("helm_list.txt"),
                               group_by_job=False)
    pn = h.cols["pod_node"]
    assert np.all(pn[:-1] <= pn[1:]), "rows must be grouped by node"
    J = h.n_jobs
    req, alloc, latest, totals = cref.top_job_columnar(h.cols, J)
    got = {h.job_name(j): (h.job_trainer(j), int(req[j]), int(alloc[j])) for j in range(J)}
    assert got == MF.HAND_JOB_TABLE
    tie = [h.job_name(j) for j in range(J)].index("hvd-tie")
    assert h.pod_name(int(latest[tie])) == "tie-a"
    order = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
    assert format_top_job(h, TopJobResult(req, alloc, order, latest, totals), MF.NOW) == rd("expected_top_job.txt", "r")


def test_arena_list_text(built):
    """`arena list` (list.go:16-80): age order, no GPU columns.  The golden text, then variants where
    a job has no status.startTime (sorts first, AGE 0s) and where two jobs start in the same second
    (tie by name), product printer against the Python oracle."""
    from arena_b200 import HostSnapshot, TopJobResult, format_job_list
    docs = [rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"), rd("helm_list.txt")]

    def product(jobs_json, tfjobs_json):
        h = HostSnapshot.from_json(docs[0], docs[1], jobs_json, tfjobs_json, docs[4], group_by_job=True)
        req, alloc, latest, totals = cref.top_job_columnar(h.cols, h.n_jobs)
        return format_job_list(h, TopJobResult(req, alloc, np.zeros(h.n_jobs, np.uint32), latest, totals), MF.NOW)

    assert product(docs[2], docs[3]) == rd("expected_list.txt", "r")
    jobs, tfjobs = json.loads(docs[2]), json.loads(docs[3])
    for mutate in ("drop", "null", "same"):
        j2, t2 = json.loads(docs[2]), json.loads(docs[3])
        if mutate == "drop":
            del [it for it in j2["items"] if it["metadata"]["name"] == "hvd"][0]["status"]["startTime"]
        elif mutate == "null":
            t2["items"][0]["status"]["startTime"] = None
        else:
            for it in j2["items"]:
                it["status"]["startTime"] = "2018-07-23T01:00:00Z"
        J2, T2 = json.dumps(j2).encode(), json.dumps(t2).encode()
        want, _ = top_ref.job_list(docs[0], J2, T2, docs[4], MF.NOW)
        assert product(J2, T2) == want, mutate
        if mutate != "same":
            assert "0s" in want.splitlines()[1], "a job without a start time is listed first"


def test_loader_rejects_what_the_reference_decode_rejects(built):
    from arena_b200 import ArenaError, HostSnapshot
    nodes = rd("nodes.json")
    for bad in ['"1x"', '""', '"1.2.3"', "true", '{"a":1}']:
        pods = json.dumps({"items": [{"metadata": {"name": "p"}, "spec": {"nodeName": "node-00", "containers": [
            {"resources": {"limits": {"cpu": "1"}, "requests": {"memory": "__BAD__"}}}]}}]}).replace('"__BAD__"', bad)
        with pytest.raises(ArenaError) as ei:           # ANY bad quantity fails the List decode, not only GPU keys
            HostSnapshot.from_json(pods.encode(), nodes)
        assert ei.value.code == -6
    with pytest.raises(ArenaError):
        HostSnapshot.from_json(b'{"items": [', nodes)
    with pytest.raises(ArenaError):
        HostSnapshot.from_json(b'{"items": [{"spec": {"nodeName": 5}}]}', nodes)
    # empty lists are fine
    h = HostSnapshot.from_json(b'{"items": []}', b'{"items": null}')
    assert (h.n_pods, h.n_nodes, h.n_containers) == (0, 0, 0)


def test_config2_generated_fixture_loader_vs_python_oracle(built):
    """BASELINE configs[1]: 1e4 TFJob/Horovod pods on 1000 nodes (generated, seed
    20180723), `top job`: product loader + C-oracle sums + product printer against the
    pure-Python path (label scans, latest-pod rule, ties)."""
    from arena_b200 import HostSnapshot, TopJobResult, format_top_job
    podlist, nodelist, joblist, tfjoblist, helm = MF.config2(n_pods_target=3000, n_nodes=300)
    P, N, Jl, T = (json.dumps(x).encode() for x in (podlist, nodelist, joblist, tfjoblist))
    text, rows = top_ref.top_job(P, Jl, T, helm, MF.NOW)
    h = HostSnapshot.from_json(P, N, Jl, T, helm.encode(), group_by_job=True)
    J = h.n_jobs
    assert J == len(rows)
    req, alloc, latest, totals = cref.top_job_columnar(h.cols, J)
    order = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
    assert format_top_job(h, TopJobResult(req, alloc, order, latest, totals), MF.NOW) == text
    # and `top node` on the same lists
    ntext, table = top_ref.top_node(P, N)
    hn = HostSnapshot.from_json(P, N)
    tot, al, cl, _ = cref.top_node_columnar(hn.cols)
    assert [(hn.node_name(n), int(tot[n]), int(al[n])) for n in range(hn.n_nodes)] == [tuple(r) for r in table]
