"""GPU: the JSON fixtures end to end through the C ABI and through the
arena-top CLI, against the expected texts the oracle produced."""

import json
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import top_ref

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
D = os.path.join(HERE, "golden", "config1")
CLI = os.path.join(ROOT, "arena_b200", "arena-top")
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_fixtures as MF  # noqa: E402


def rd(name, mode="rb"):
    with open(os.path.join(D, name), mode) as f:
        return f.read()


def test_config1_top_node_text(gpu):
    h = gpu.HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"))
    dev = gpu.DeviceSnapshot.upload(h)
    res = gpu.top_node(dev, details=True)
    assert gpu.format_top_node(h, res) == rd("expected_top_node.txt", "r")
    assert gpu.format_top_node(h, res, details=True) == rd("expected_top_node_details.txt", "r")
    table = [[h.node_name(n), int(res.node_total[n]), int(res.node_alloc[n])] for n in range(h.n_nodes)]
    assert table == json.loads(rd("expected_node_table.json"))
    got = {r[0]: (r[1], r[2]) for r in table}
    for name, want in MF.HAND_NODE_TABLE_EDGE.items():
        assert got[name] == want


def test_config1_top_job_text(gpu):
    h = gpu.HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"),
                                   rd("helm_list.txt"), group_by_job=True)
    res = gpu.top_job(gpu.DeviceSnapshot.upload(h))
    assert gpu.format_top_job(h, res, MF.NOW) == rd("expected_top_job.txt", "r")
    got = {h.job_name(j): (h.job_trainer(j), int(res.job_req[j]), int(res.job_alloc[j])) for j in range(h.n_jobs)}
    assert got == MF.HAND_JOB_TABLE


def test_config2_top_job_full_size(gpu):
    """BASELINE configs[1]: top job over 1e4 TFJob/Horovod pods on 1000 nodes."""
    podlist, nodelist, joblist, tfjoblist, helm = MF.config2()
    P, N, Jl, T = (json.dumps(x).encode() for x in (podlist, nodelist, joblist, tfjoblist))
    assert len(podlist["items"]) >= 10_000 and len(nodelist["items"]) == 1000
    text, rows = top_ref.top_job(P, Jl, T, helm, MF.NOW)
    h = gpu.HostSnapshot.from_json(P, N, Jl, T, helm.encode(), group_by_job=True)
    res = gpu.top_job(gpu.DeviceSnapshot.upload(h))
    assert gpu.format_top_job(h, res, MF.NOW) == text
    # node-grouped rows: the same table, also where two launcher pods tie on the timestamp (the
    # loader settles "first in LIST order" before it regroups the rows)
    h2 = gpu.HostSnapshot.from_json(P, N, Jl, T, helm.encode(), group_by_job=False)
    res2 = gpu.top_job(gpu.DeviceSnapshot.upload(h2))
    assert np.array_equal(res2.job_req, res.job_req) and np.array_equal(res2.job_alloc, res.job_alloc)
    assert np.array_equal(res2.job_order, res.job_order) and np.array_equal(res2.totals, res.totals)
    assert gpu.format_top_job(h2, res2, MF.NOW) == text
    ntext, _ = top_ref.top_node(P, N)
    r = gpu.top_node(gpu.DeviceSnapshot.upload(h2))
    assert gpu.format_top_node(h2, r) == ntext


def test_arena_list_on_the_gpu(gpu):
    """`arena list`: STATUS / NODE of mpijob and standalonejob rows come from the latest-pod rule the
    `top job` kernels evaluate; config 1 golden text and the generated config 2 against the oracle."""
    h = gpu.HostSnapshot.from_json(rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"),
                                   rd("helm_list.txt"), group_by_job=True)
    res = gpu.top_job(gpu.DeviceSnapshot.upload(h))
    assert gpu.format_job_list(h, res, MF.NOW) == rd("expected_list.txt", "r")
    podlist, nodelist, joblist, tfjoblist, helm = MF.config2(n_pods_target=3000, n_nodes=300)
    P, N, Jl, T = (json.dumps(x).encode() for x in (podlist, nodelist, joblist, tfjoblist))
    want, _ = top_ref.job_list(P, Jl, T, helm, MF.NOW)
    h = gpu.HostSnapshot.from_json(P, N, Jl, T, helm.encode(), group_by_job=False)
    res = gpu.top_job(gpu.DeviceSnapshot.upload(h))
    assert gpu.format_job_list(h, res, MF.NOW) == want


def run_cli(*args):
    p = subprocess.run([CLI, *args], capture_output=True, text=True, timeout=120)
    return p.returncode, p.stdout, p.stderr


def test_cli_matches_expected_output(gpu):
    pods, nodes = os.path.join(D, "pods.json"), os.path.join(D, "nodes.json")
    rc, out, err = run_cli("node", "--pods", pods, "--nodes", nodes, "--time")     # default: the GPU loader ...
    assert rc == 0 and out == rd("expected_top_node.txt", "r")
    assert "host loader took over" in err or "handed back" in err    # ... which hands this fixture's exotic items to the host
    rc, out, _ = run_cli("node", "--pods", pods, "--nodes", nodes, "--loader", "host")
    assert rc == 0 and out == rd("expected_top_node.txt", "r")
    rc, out, _ = run_cli("node", "-d", "--pods", pods, "--nodes", nodes)
    assert rc == 0 and out == rd("expected_top_node_details.txt", "r")
    rc, out, err = run_cli("job", "--pods", pods, "--nodes", nodes, "--jobs", os.path.join(D, "jobs.json"), "--tfjobs",
                           os.path.join(D, "tfjobs.json"), "--releases", os.path.join(D, "helm_list.txt"), "--now",
                           str(MF.NOW), "--time")
    assert rc == 0 and out == rd("expected_top_job.txt", "r") and "[arena-top]" in err
    rc, out, err = run_cli("list", "--pods", pods, "--nodes", nodes, "--jobs", os.path.join(D, "jobs.json"), "--tfjobs",
                           os.path.join(D, "tfjobs.json"), "--releases", os.path.join(D, "helm_list.txt"), "--now", str(MF.NOW))
    assert rc == 0 and out == rd("expected_list.txt", "r")


def test_cli_parses_an_api_server_shaped_list_on_the_gpu(gpu, tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import loader_timing as LT
    pods, nodes = LT.make_docs(2000, 40, seed=9)
    (tmp_path / "pods.json").write_bytes(pods)
    (tmp_path / "nodes.json").write_bytes(nodes)
    want, _ = top_ref.top_node(pods, nodes)
    for loader in ("device", "host"):
        rc, out, err = run_cli("node", "--pods", str(tmp_path / "pods.json"), "--nodes", str(tmp_path / "nodes.json"),
                               "--loader", loader, "--time")
        assert rc == 0 and out == want, loader
        assert ("parsed on the GPU" in err) == (loader == "device")
    # -d needs pod names and namespaces: the whole host snapshot, the PodList still decoded on the GPU
    want_d, _ = top_ref.top_node(pods, nodes, details=True)
    for loader in ("device", "host"):
        rc, out, err = run_cli("node", "-d", "--pods", str(tmp_path / "pods.json"), "--nodes", str(tmp_path / "nodes.json"),
                               "--loader", loader, "--time")
        assert rc == 0 and out == want_d, loader
        assert ("parsed on the GPU" in err) == (loader == "device")


def test_cli_top_job_decodes_the_podlist_on_the_gpu(gpu, tmp_path):
    podlist, nodelist, joblist, tfjoblist, helm = MF.config2(n_pods_target=3000, n_nodes=300)
    docs = {"pods.json": podlist, "nodes.json": nodelist, "jobs.json": joblist, "tfjobs.json": tfjoblist}
    for name, doc in docs.items():
        (tmp_path / name).write_text(json.dumps(doc))
    (tmp_path / "helm.txt").write_text(helm)
    P, Jl, T = (json.dumps(x).encode() for x in (podlist, joblist, tfjoblist))
    want_job, _ = top_ref.top_job(P, Jl, T, helm, MF.NOW)
    want_list, _ = top_ref.job_list(P, Jl, T, helm, MF.NOW)
    flags = ["--pods", str(tmp_path / "pods.json"), "--nodes", str(tmp_path / "nodes.json"), "--jobs", str(tmp_path / "jobs.json"),
             "--tfjobs", str(tmp_path / "tfjobs.json"), "--releases", str(tmp_path / "helm.txt"), "--now", str(MF.NOW), "--time"]
    for loader in ("device", "host"):
        rc, out, err = run_cli("job", *flags, "--loader", loader)
        assert rc == 0 and out == want_job, loader
        assert ("parsed on the GPU" in err) == (loader == "device"), err
        rc, out, err = run_cli("list", *flags, "--loader", loader)
        assert rc == 0 and out == want_list, loader


def test_cli_error_behaviour(gpu, tmp_path):
    """Errors go to stdout and the exit status is 1 (top_node.go:34-48)."""
    bad = tmp_path / "pods.json"
    bad.write_text('{"items":[{"spec":{"containers":[{"resources":{"limits":{"nvidia.com/gpu":"1x"}}}]}}]}')
    rc, out, _ = run_cli("node", "--pods", str(bad), "--nodes", os.path.join(D, "nodes.json"))
    assert rc == 1 and "invalid quantity" in out
    rc, out, _ = run_cli("node", "--pods", "/nonexistent.json", "--nodes", os.path.join(D, "nodes.json"))
    assert rc == 1 and "no such file" in out
    rc, out, _ = run_cli()
    assert rc == 0 and "Display Resource (GPU) usage." in out
