"""GPU: arena_load_json_device (PodList parsed on the device) against the host loader
and the Python oracle: same columns, same `top node` table, same errors."""

import json
import os
import random
import sys

import numpy as np
import pytest

from oracle import top_ref

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
D = os.path.join(HERE, "golden", "config1")
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, HERE)
import loader_timing as LT  # noqa: E402
from test_loader_scan import nasty_podlist  # noqa: E402
from test_pod_scan import host_rows_in_doc_order  # noqa: E402


def rd(name, mode="rb"):
    with open(os.path.join(D, name), mode) as f:
        return f.read()


def device_rows(dev):
    c = dev.download()
    cb = c["pod_ctr_begin"]
    return [(int(c["pod_node"][p]), int(c["pod_phase"][p]),
             [(int(c["ctr_val"][k]), int(c["ctr_scale"][k])) for k in range(cb[p], cb[p + 1])]) for p in range(dev.n_pods)]


def check_device_load(gpu, pods, nodes, expect_fallback=None):
    dev, stats = gpu.DeviceSnapshot.from_json_device(pods, nodes)
    h = gpu.HostSnapshot.from_json(pods, nodes)
    if expect_fallback is not None:
        assert bool(stats["fell_back"]) == expect_fallback, stats
    assert (dev.n_pods, dev.n_nodes, dev.n_containers) == (h.n_pods, h.n_nodes, h.n_containers)
    if not stats["fell_back"]:
        assert dev.host.n_pods == 0 and stats["n_items"] == h.n_pods
        assert device_rows(dev) == host_rows_in_doc_order(h)              # rows in PodList order
    c = dev.download()
    assert np.array_equal(c["node_cap_val"], h.cols["node_cap_val"]) and np.array_equal(c["node_cap_scale"], h.cols["node_cap_scale"])
    res = gpu.top_node(dev)
    _, table = top_ref.top_node(pods, nodes)
    got = [(dev.host.node_name(n), int(res.node_total[n]), int(res.node_alloc[n])) for n in range(dev.n_nodes)]
    assert got == [tuple(r) for r in table]
    text, _ = top_ref.top_node(pods, nodes)
    assert gpu.format_top_node(dev.host, res) == text
    return stats


@pytest.mark.parametrize("indent", [None, 2])
def test_api_server_shaped_podlist(gpu, indent):
    pods, nodes = LT.make_docs(3000, 64, seed=3)
    if indent:
        pods = json.dumps(json.loads(pods), indent=indent).encode()
    stats = check_device_load(gpu, pods, nodes, expect_fallback=False)
    assert stats["n_items"] == 3000


def test_strings_full_of_brackets_quotes_and_backslashes(gpu):
    for seed, indent in ((5, None), (6, 1)):
        pods, nodes = nasty_podlist(1200, 37, seed=seed, indent=indent)
        check_device_load(gpu, pods, nodes, expect_fallback=False)


def test_edge_case_fixture_hands_items_back_with_equal_results(gpu):
    """configs[0] is full of spellings the device scanner does not decide (slow-path quantities, escaped keys ...):
    those items alone go through the host's item code and are patched into the device columns."""
    stats = check_device_load(gpu, rd("pods.json"), rd("nodes.json"))
    assert stats["flag_reason"] != 0
    assert stats["fell_back"] or stats["n_handed_back"] > 0


def test_handed_back_items_keep_the_device_path(gpu):
    """A plain PodList with a few exotic items: a pod of 12 containers, a slow-path GPU quantity ("1.5Gi"), an
    escaped key, a timestamp with an offset.  The document stays on the device, the exotic rows come from the host."""
    pods, nodes = LT.make_docs(5_000, 50, seed=13)
    doc = json.loads(pods)
    items = doc["items"]
    ctr = {"name": "c", "resources": {"limits": {"nvidia.com/gpu": "1"}}}
    items[7]["spec"]["containers"] = [dict(ctr) for _ in range(12)]
    items[4_000]["spec"]["containers"][0].setdefault("resources", {}).setdefault("limits", {})["nvidia.com/gpu"] = "1.5Gi"
    items[4_999]["metadata"]["creationTimestamp"] = "2024-01-01T00:00:00+08:00"
    text = json.dumps(doc).replace('"nodeName"', '"node\\u004eame"', 1)
    stats = check_device_load(gpu, text.encode(), nodes, expect_fallback=False)
    assert 3 <= stats["n_handed_back"] <= 8, stats
    # an item the host rejects sends the whole document to the host, with the host's message
    items[100]["spec"]["containers"][0].setdefault("resources", {}).setdefault("limits", {})["nvidia.com/gpu"] = "1x"
    bad = json.dumps(doc).encode()
    with pytest.raises(gpu.ArenaError) as e1:
        gpu.HostSnapshot.from_json(bad, nodes)
    with pytest.raises(gpu.ArenaError) as e2:
        gpu.DeviceSnapshot.from_json_device(bad, nodes)
    assert str(e1.value) == str(e2.value)


def test_envelope_variants(gpu):
    pods, nodes = LT.make_docs(50, 5, seed=1)
    items = json.loads(pods)["items"]
    for doc in ({"items": []}, {"items": None}, {"kind": "PodList"}, {"items": items[:1]},
                {"metadata": {"x": [{"a": [1, {"b": "]"}]}]}, "items": items, "zz": [[{"k": "{"}]]}):
        check_device_load(gpu, json.dumps(doc).encode(), nodes, expect_fallback=False)
    # whitespace everywhere, items first / last
    text = json.dumps({"items": items[:7], "kind": "PodList"}, indent=4).replace("\n", "\r\n \t")
    check_device_load(gpu, text.encode(), nodes, expect_fallback=False)
    # not decidable on the device: duplicate "items", scalar element
    dup = b'{"items": [], ' + json.dumps({"items": items}).encode()[1:]
    check_device_load(gpu, dup, nodes, expect_fallback=True)
    for bad in (b'{"items": [1]}', b'{"items": [{}, 2]}', b'{"items": {"a": 1}}', b'[{"items": []}]', b'{"items": [[]]}'):
        with pytest.raises(gpu.ArenaError) as e1:
            gpu.HostSnapshot.from_json(bad, nodes)
        with pytest.raises(gpu.ArenaError) as e2:
            gpu.DeviceSnapshot.from_json_device(bad, nodes)
        assert str(e1.value) == str(e2.value)


def test_errors_are_the_host_loader_s(gpu):
    pods, nodes = LT.make_docs(400, 9, seed=2)
    rng = random.Random(4)
    cases = [pods[:cut] for cut in sorted(rng.sample(range(1, len(pods) - 1), 25))]
    broken = bytearray(pods)
    k = pods.index(b'"restartPolicy": "Never"', len(pods) // 2)
    broken[k + 17:k + 24] = b'nEver" '
    cases.append(bytes(broken))
    doc = json.loads(pods)
    doc["items"][123]["spec"]["containers"][0]["resources"]["requests"]["memory"] = "1x"
    cases.append(json.dumps(doc).encode())
    doc = json.loads(pods)
    doc["items"][7]["spec"]["nodeName"] = 5
    cases.append(json.dumps(doc).encode())
    for case in cases:
        with pytest.raises(gpu.ArenaError) as e1:
            gpu.HostSnapshot.from_json(case, nodes)
        with pytest.raises(gpu.ArenaError) as e2:
            gpu.DeviceSnapshot.from_json_device(case, nodes)
        assert str(e1.value) == str(e2.value)
    bad_nodes = json.loads(nodes)
    bad_nodes["items"][3]["status"]["capacity"]["nvidia.com/gpu"] = "1x"
    with pytest.raises(gpu.ArenaError) as e2:
        gpu.DeviceSnapshot.from_json_device(pods, json.dumps(bad_nodes).encode())
    assert "invalid quantity" in str(e2.value)


def test_mutated_documents_agree_with_the_host_loader(gpu):
    """The fuzz of tests/test_pod_scan.py, through the device this time."""
    import test_pod_scan as T
    rng = random.Random(77)
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(6)]}).encode()
    seen = {"device": 0, "handed": 0, "fallback": 0, "error": 0}
    for trial in range(150):
        items = []
        for i in range(rng.randrange(1, 30)):
            pod = LT.make_pod(rng.randrange(100), rng, 8)
            pod["metadata"]["managedFields"] = []
            if rng.random() < 0.05:
                T.set_path(pod, rng.choice(T.PATHS), rng.choice(T.WEIRD_VALUES), rng)
            items.append(pod)
        text = json.dumps({"kind": "PodList", "items": items}, indent=rng.choice([None, 1]))
        if rng.random() < 0.2:
            text = T.text_tricks(text, rng)
        case = text.encode()
        try:
            gpu.HostSnapshot.from_json(case, nodes)
        except gpu.ArenaError as e1:
            with pytest.raises(gpu.ArenaError) as e2:
                gpu.DeviceSnapshot.from_json_device(case, nodes)
            assert str(e1) == str(e2.value)
            seen["error"] += 1
            continue
        stats = check_device_load(gpu, case, nodes)
        seen["fallback" if stats["fell_back"] else "handed" if stats["n_handed_back"] else "device"] += 1
    assert seen["device"] > 20 and seen["handed"] + seen["fallback"] > 5 and seen["error"] > 5, seen


def test_many_nodes_and_unmatched_names(gpu):
    """Hash join against 20000 node names, a third of the pods naming nodes that do not exist."""
    n_nodes = 20000
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(n_nodes)]}).encode()
    rng = random.Random(11)
    items = []
    for i in range(5000):
        pod = LT.make_pod(i, rng, n_nodes * 3 // 2)
        pod["metadata"]["managedFields"] = []
        items.append(pod)
    check_device_load(gpu, json.dumps({"items": items}).encode(), nodes, expect_fallback=False)


def test_fractional_gpu_limits_stay_on_the_device(gpu):
    """GPU-sharing pods ask for "500m" / "250m" (BASELINE configs[4]); a few write "1.5", "1e1", "2k".  The device
    scanner parses the reference's fast-path quantity grammar itself, so such a PodList is NOT handed to the host,
    and its columns carry the same {value, scale} pairs as the host loader's."""
    pods, nodes = LT.make_docs(20_000, 200, seed=11)
    doc = json.loads(pods)
    rnd = __import__("random").Random(3)
    changed = 0
    for it in doc["items"]:
        for c in it["spec"]["containers"]:
            lim = c.get("resources", {}).get("limits")
            if lim and "nvidia.com/gpu" in lim and rnd.random() < 0.1:
                lim["nvidia.com/gpu"] = rnd.choice(["500m", "250m", "500m", "1500m", "1.5", "1e1", "2k", "0.5"])
                changed += 1
    assert changed > 200
    stats = check_device_load(gpu, json.dumps(doc).encode(), nodes, expect_fallback=False)
    assert stats["n_items"] == 20_000


# ---- arena_load_json_gpu: the whole host snapshot (names, labels, job table) with the PodList decoded on the GPU ----

def test_full_snapshot_from_the_gpu_equals_the_host_loader_on_the_fixture(gpu):
    from test_loader_full_scan import same_snapshot
    docs = (rd("pods.json"), rd("nodes.json"), rd("jobs.json"), rd("tfjobs.json"), rd("helm_list.txt"))
    for by_job in (False, True):
        want = gpu.HostSnapshot.from_json(*docs, group_by_job=by_job)
        got, st = gpu.HostSnapshot.from_json_gpu(*docs, group_by_job=by_job)
        assert not st["fell_back"] and st["n_items"] == want.n_pods and 0 < st["n_handed_back"] < want.n_pods // 2
        same_snapshot(got, want)
    # and the texts the user sees, computed on the GPU from that snapshot
    import make_fixtures as MF
    got, _ = gpu.HostSnapshot.from_json_gpu(*docs, group_by_job=True)
    res = gpu.top_job(gpu.DeviceSnapshot.upload(got))
    assert gpu.format_top_job(got, res, MF.NOW) == rd("expected_top_job.txt", "r")
    assert gpu.format_job_list(got, res, MF.NOW) == rd("expected_list.txt", "r")
    got, _ = gpu.HostSnapshot.from_json_gpu(docs[0], docs[1])
    assert gpu.format_top_node(got, gpu.top_node(gpu.DeviceSnapshot.upload(got), details=True), details=True) == rd("expected_top_node_details.txt", "r")


def test_full_snapshot_config2_from_pinned_memory(gpu):
    """BASELINE configs[2] shape (jobs of all three trainers) at 3000 pods, the document in page-locked memory."""
    from test_loader_full_scan import same_snapshot
    import make_fixtures as MF
    podlist, nodelist, joblist, tfjoblist, helm = MF.config2(n_pods_target=3000, n_nodes=300)
    P, N, Jl, T = (json.dumps(x).encode() for x in (podlist, nodelist, joblist, tfjoblist))
    want = gpu.HostSnapshot.from_json(P, N, Jl, T, helm.encode(), group_by_job=True)
    buf = gpu.PinnedBuffer(len(P)).fill(P)
    got, st = gpu.HostSnapshot.from_json_gpu(buf, N, Jl, T, helm.encode(), group_by_job=True)
    assert not st["fell_back"] and st["n_handed_back"] == 0 and st["n_items"] == want.n_pods
    same_snapshot(got, want)
    text, _ = top_ref.top_job(P, Jl, T, helm, MF.NOW)
    assert gpu.format_top_job(got, gpu.top_job(gpu.DeviceSnapshot.upload(got)), MF.NOW) == text


def test_full_snapshot_fuzz_and_errors(gpu):
    """The job-field fuzz of tests/test_loader_full_scan.py through the device; errors are the host loader's."""
    import test_loader_full_scan as F
    rng = random.Random(99)
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(5)]}).encode()
    helm = b"NAME\tNAMESPACE\nr0\t1\tdefault\nr1\t1\tdefault\njob-00000\t1\tdefault\n"
    jobs = json.dumps({"items": [
        {"metadata": {"namespace": "default", "labels": {"release": "r0", "app": "tf-horovod"}}, "status": {"active": 2}},
        {"metadata": {"namespace": "default", "labels": {"release": "r1", "app": "training"}}, "status": {"succeeded": 1}},
        {"metadata": {"namespace": "default", "labels": {"release": "job-00000", "app": "tf-horovod"}}, "status": {}}]}).encode()
    seen = {"device": 0, "handed": 0, "fallback": 0, "error": 0}
    for trial in range(120):
        items = []
        for i in range(rng.randrange(1, 40)):
            pod = LT.make_pod(rng.randrange(8), rng, 6)
            pod["metadata"]["managedFields"] = []
            if rng.random() < 0.3:
                F.mutate_meta(pod, rng)
            items.append(pod)
        case = json.dumps({"kind": "PodList", "items": items}, ensure_ascii=rng.random() < 0.5).replace("\\\\u00", "\\u00").encode()
        by_job = rng.random() < 0.5
        try:
            want = gpu.HostSnapshot.from_json(case, nodes, jobs, None, helm, group_by_job=by_job)
        except gpu.ArenaError as e1:
            with pytest.raises(gpu.ArenaError) as e2:
                gpu.HostSnapshot.from_json_gpu(case, nodes, jobs, None, helm, group_by_job=by_job)
            assert str(e1) == str(e2.value)
            seen["error"] += 1
            continue
        got, st = gpu.HostSnapshot.from_json_gpu(case, nodes, jobs, None, helm, group_by_job=by_job)
        F.same_snapshot(got, want)
        seen["fallback" if st["fell_back"] else "handed" if st["n_handed_back"] else "device"] += 1
    assert seen["device"] > 10 and seen["handed"] > 10 and seen["error"] > 3, seen
    # a bad JobList: the PodList was fine, the message is still arena_load_json's
    pods, _ = LT.make_docs(50, 5, seed=1)
    for bad in (b'{"items": 5}', b'{"items": [', b'{"items":[{"status":{"active":"x"}, "metadata":{"labels":{"release":"r0","app":"tf-horovod"},"namespace":"default"}}]}'):
        with pytest.raises(gpu.ArenaError) as e1:
            gpu.HostSnapshot.from_json(pods, nodes, bad, None, helm)
        with pytest.raises(gpu.ArenaError) as e2:
            gpu.HostSnapshot.from_json_gpu(pods, nodes, bad, None, helm)
        assert str(e1.value) == str(e2.value)
