"""csrc/pod_scan.cuh -- the per-pod scanner the GPU loader runs one thread per item --
compiled for the host and compared with the host loader (loader.cc), which itself is
pinned to the Python oracle (tests/test_fixtures.py, tests/test_loader_scan.py).

The scanner's contract: flag == 0  =>  node / phase / per-container GPU values equal the
host loader's; anything it does not decide itself is flagged (the whole document then
takes the host path).  So for ANY input: where the host loader fails, some item must be
flagged; where it succeeds, every unflagged item must agree with it."""

import json
import os
import random
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
import loader_timing as LT  # noqa: E402

D = os.path.join(HERE, "golden", "config1")


def rd(name):
    with open(os.path.join(D, name), "rb") as f:
        return f.read()


def host_rows_in_doc_order(h):
    """(node, phase, [ (val, scale) per container ]) of every pod, in PodList order."""
    order = h.pod_order()
    cb = h.cols["pod_ctr_begin"]
    out = [None] * h.n_pods
    for row in range(h.n_pods):
        ctrs = [(int(h.cols["ctr_val"][c]), int(h.cols["ctr_scale"][c])) for c in range(cb[row], cb[row + 1])]
        out[int(order[row])] = (int(h.cols["pod_node"][row]), int(h.cols["pod_phase"][row]), ctrs)
    return out


def check_against_host(pods: bytes, nodes: bytes, n_items: int):
    """Returns (host_ok, n_flagged).  Asserts the contract."""
    from arena_b200 import ArenaError, HostSnapshot
    hn = HostSnapshot.from_json(b'{"items":[]}', nodes)
    try:
        r = hn.scan_podlist(pods, n_items + 8)
    except ArenaError:
        r = None                       # the array could not even be split: the device path gives up as well
    try:
        h = HostSnapshot.from_json(pods, nodes)
    except ArenaError:
        assert r is None or int((r["flag"] != 0).sum()) > 0, "host loader fails but no item is flagged"
        return False, None
    if r is None:
        return True, None
    want = host_rows_in_doc_order(h)
    assert r["n"] == len(want)
    for k, (node, phase, ctrs) in enumerate(want):
        if r["flag"][k]:
            continue
        n_c = int(r["n_ctrs"][k])
        got = (int(r["node"][k]), int(r["phase"][k]), [(int(v), int(sc)) for v, sc in zip(r["ctr_val"][k][:n_c], r["ctr_scale"][k][:n_c])])
        assert got == (node, phase, ctrs), (k, got, (node, phase, ctrs))
    return True, int((r["flag"] != 0).sum())


def test_api_server_shaped_pods_are_never_flagged(built):
    pods, nodes = LT.make_docs(1500, 40, seed=7)
    ok, flagged = check_against_host(pods, nodes, 1500)
    assert ok and flagged == 0
    pods = json.dumps(json.loads(pods), indent=2).encode()           # kubectl -o json layout
    ok, flagged = check_against_host(pods, nodes, 1500)
    assert ok and flagged == 0


def test_edge_case_fixture(built):
    """configs[0]: deprecated key, null quantities, unmatched nodes, init containers,
    suffixed quantities ... whatever the scanner keeps must match; the rest is flagged."""
    n = len(json.loads(rd("pods.json"))["items"])
    ok, flagged = check_against_host(rd("pods.json"), rd("nodes.json"), n)
    assert ok
    assert flagged < n, "the scanner should decide most of the fixture itself"


def test_gpu_quantities_on_the_fast_path_are_decided_by_the_scanner(built):
    """The scanner parses the reference's fast-path grammar (quantity.go:276-343) itself: "500m", "1k", "1.5", "2Ki",
    "1e3" ... give the same int64Amount{value, scale} as the host loader and are NOT handed back; the inf.Dec slow
    path, out-of-range scales and malformed text are flagged (and then match the host by delegation)."""
    import random
    from arena_b200 import HostSnapshot
    rnd = random.Random(5)
    fast = ["0", "1", "8", "0008", "500m", "1500m", "2000m", "250m", "0.5", "1.5", ".5", "3.", "1k", "2M", "1G", "1e3", "1E3", "1e0",
            "12e-1", "1.5e1", "+3", "-3", "-500m", "1Ki", "2Ki", "1Mi", "3Gi", "100u", "5n", "123456789012345678", "0m", "00", "1T", "2P",
            "1E", "7e+2", "0.000000001", "999999999999999999"]
    slow_or_bad = ["1.5Gi", "0.0000000001", "1234567890123456789", "10E", "1e30", "1e19", "1x", "", "e", "1.2.3", " 4", "4 ", "-", "m",
                   "1Ei", "9Ei", "0.5Ki", "1e-10", "12345678901234567890"]
    extra = []
    for _ in range(300):
        body = rnd.choice(["", "-", "+"]) + str(rnd.randrange(0, 10 ** rnd.randrange(1, 19)))
        if rnd.random() < 0.4:
            body += "." + "".join(rnd.choice("0123456789") for _ in range(rnd.randrange(0, 6)))
        body += rnd.choice(["", "", "m", "k", "M", "Ki", "Gi", "n", "u", "e3", "E-2", "e+5", "G", "T", "P", "E", "Ti"])
        extra.append(body)
    allq = fast + slow_or_bad + extra
    nodes = json.dumps({"items": [{"metadata": {"name": "n0"}, "status": {"capacity": {"nvidia.com/gpu": "8"}}}]}).encode()
    hn = HostSnapshot.from_json(b'{"items":[]}', nodes)
    n_fast_unflagged = 0
    for q in allq:
        pod = {"metadata": {"name": "p"}, "spec": {"nodeName": "n0", "containers": [{"resources": {"limits": {"nvidia.com/gpu": q}}}]},
               "status": {"phase": "Running"}}
        doc = json.dumps({"items": [pod]}).encode()
        r = hn.scan_podlist(doc, 4)
        try:
            h = HostSnapshot.from_json(doc, nodes)
        except Exception:
            assert r["flag"][0] != 0, f"{q!r}: the host rejects it but the scanner decided it"
            continue
        if r["flag"][0]:
            assert q not in fast, f"{q!r} is on the fast path but was handed back"
            continue
        want = (int(h.cols["ctr_val"][0]), int(h.cols["ctr_scale"][0]))
        got = (int(r["ctr_val"][0][0]), int(r["ctr_scale"][0][0]))
        assert got == want, (q, got, want)
        n_fast_unflagged += q in fast
    assert n_fast_unflagged == len(fast)


# ---- mutation fuzz ------------------------------------------------------------------------
WEIRD_VALUES = [None, True, 5, -1, 1.5, "", "x", [], {}, [1], {"a": 1}, "8", "0008", "1k", "500m", " 4 ", "4 ",
                "1e3", "1E3", "2Ki", "+3", "-3", "3.", ".5", "0.0", "1x", "1.2.3", "12345678901234567890", 123456789012345678,
                "2024-02-30T00:00:00Z", "2024-02-29T00:00:00Z", "2023-02-29T00:00:00Z", "2024-01-01T00:00:00+08:00",
                "2024-01-01t00:00:00z", "2024-01-01T00:00:00.5Z", "2024-13-01T00:00:00Z", "Running", "Succeeded", "Failed",
                "node-0001", "node-9999", "nvidia.com/gpu"]
PATHS = [
    ("metadata",), ("spec",), ("status",), ("metadata", "name"), ("metadata", "namespace"), ("metadata", "labels"),
    ("metadata", "labels", "app"), ("metadata", "labels", "release"), ("metadata", "labels", "zzz"),
    ("metadata", "creationTimestamp"), ("metadata", "ownerReferences"), ("metadata", "ownerReferences", 0),
    ("metadata", "ownerReferences", 0, "kind"), ("spec", "nodeName"), ("spec", "containers"), ("spec", "initContainers"),
    ("spec", "containers", 0), ("spec", "containers", 1), ("spec", "containers", 0, "resources"),
    ("spec", "containers", 0, "resources", "limits"), ("spec", "containers", 0, "resources", "requests"),
    ("spec", "containers", 0, "resources", "limits", "nvidia.com/gpu"),
    ("spec", "containers", 0, "resources", "limits", "alpha.kubernetes.io/nvidia-gpu"),
    ("spec", "containers", 0, "resources", "limits", "cpu"), ("spec", "containers", 0, "resources", "requests", "memory"),
    ("spec", "containers", 1, "resources", "limits", "nvidia.com/gpu"), ("status", "phase"), ("status", "hostIP"),
]


def set_path(obj, path, value, rng):
    for k in path[:-1]:
        if isinstance(k, int):
            if not isinstance(obj, list) or k >= len(obj):
                return
            obj = obj[k]
        else:
            if not isinstance(obj, dict):
                return
            obj = obj.setdefault(k, {})
    k = path[-1]
    if isinstance(k, int):
        if isinstance(obj, list) and k < len(obj):
            obj[k] = value
    elif isinstance(obj, dict):
        if rng.random() < 0.15:
            obj.pop(k, None)
        else:
            obj[k] = value


def text_tricks(text: str, rng) -> str:
    """Spellings json.dumps never produces: escaped keys and values, duplicate keys."""
    r = rng.random()
    if r < 0.15:
        key = rng.choice(["nodeName", "phase", "limits", "spec", "containers", "nvidia.com/gpu", "resources", "kind"])
        esc = "\\u%04x" % ord(key[0]) + key[1:]
        return text.replace('"%s"' % key, '"%s"' % esc, rng.choice([1, 1000]))
    if r < 0.3:
        key = rng.choice(["nodeName", "phase", "limits", "spec", "containers", "status", "resources", "nvidia.com/gpu"])
        dup = rng.choice(['"%s": null, ' % key, '"%s": "x", ' % key, '"%s": {}, ' % key, '"%s": [], ' % key])
        return text.replace('"%s":' % key, dup + '"%s":' % key, rng.choice([1, 1000]))
    if r < 0.4:
        return text.replace('"Running"', '"Run\\u006eing"', 1).replace('"node-0001"', '"node-000\\u0031"', 1)
    if r < 0.5 and len(text) > 10:
        k = rng.randrange(len(text))
        return text[:k] + rng.choice(['"', "\\", "}", "]", "{", "[", ",", ":", " ", "x", "\n", "\x01"]) + text[k + 1:]
    return text


def test_mutation_fuzz_against_host_loader(built):
    rng = random.Random(20180723)
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(6)]}).encode()
    outcomes = {"host_ok": 0, "host_err": 0, "flagged": 0, "clean": 0}
    for trial in range(1500):
        n = rng.randrange(1, 6)
        items = []
        for i in range(n):
            pod = LT.make_pod(rng.randrange(100), rng, 8)          # node ids 6, 7 match nothing
            pod["metadata"]["managedFields"] = []                  # keep the documents small
            if rng.random() < 0.3:
                pod["spec"]["initContainers"] = [{"name": "init", "resources": {"limits": {"nvidia.com/gpu": "9", "cpu": "1"}}}]
            if rng.random() < 0.2:
                pod["spec"]["containers"] = pod["spec"]["containers"] * rng.randrange(0, 6)
            for _ in range(rng.randrange(0, 3)):
                set_path(pod, rng.choice(PATHS), rng.choice(WEIRD_VALUES), rng)
            items.append(pod)
        text = json.dumps({"kind": "PodList", "items": items}, indent=rng.choice([None, None, 1]))
        if rng.random() < 0.5:
            text = text_tricks(text, rng)
        ok, flagged = check_against_host(text.encode(), nodes, n)
        outcomes["host_ok" if ok else "host_err"] += 1
        if ok and flagged is not None:
            outcomes["flagged" if flagged else "clean"] += 1
    # the fuzz must exercise all three outcomes
    assert outcomes["host_err"] > 20 and outcomes["flagged"] > 20 and outcomes["clean"] > 20, outcomes


VALID_TOKENS = ["0", "-0", "1", "-12", "1.5", "-0.5e10", "1E+2", "1e-2", "0.0", "true", "false", "null", '""', '"a\\"b"',
                '"\\\\"', '"\\/\\b\\f\\n\\r\\t"', '"\\u00e9\\uD83D\\uDE00"', '"\\uabCD"', "[]", "{}", "[[],{}]", '{"a":{"b":[1,2,{"c":null}]}}',
                "[ 1 , 2 ]", '{ "k" : "v" }', '"}{]["', '"é"', "[1,[2,[3,[4,[5]]]]]", "1234567890123456789012345678901234567890"]
INVALID_TOKENS = ["01", "-", "1.", ".5", "1e", "1e+", "+1", "tru", "nul", "falsy", "True", "NULL", '"\\x"', '"\\u12G4"', '"\\u123"',
                  '"a\tb"', '"a\nb"', '"\x01"', "[1,]", "[,1]", "{,}", '{"a"}', '{"a":}', '{"a":1,}', "{1:2}", "[1 2]", "'a'",
                  "[}", "{]", '{"a":1]', "[1}", "1 2", "--1", "1e1.5", "0x10", "1_000", "NaN", "Infinity", '"unterminated', "@"]


def test_lexer_accepts_and_flags_what_the_host_parser_does(built):
    """Every valid JSON token shape in a place nobody looks at must stay on the device
    path; every invalid one must be handed back (the host then reports the error)."""
    from arena_b200 import ArenaError, HostSnapshot
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(3)]}).encode()
    rng = random.Random(5)
    base = LT.make_pod(1, rng, 3)
    base["metadata"]["managedFields"] = []
    for tokens, valid in ((VALID_TOKENS, True), (INVALID_TOKENS, False)):
        for tok in tokens:
            for where in ("annotation", "element", "item_member"):
                pod = json.loads(json.dumps(base))
                if where == "annotation":
                    pod["metadata"]["annotations"]["x"] = "@@"
                elif where == "element":
                    pod["spec"]["tolerations"].append("@@")
                else:
                    pod["zzz"] = "@@"
                for pad in ("", " ", "\n\t "):
                    text = json.dumps({"items": [pod, base]}).replace('"@@"', pad + tok + pad)
                    ok, flagged = check_against_host(text.encode(), nodes, 2)
                    assert ok == valid, (tok, where)
                    if valid:
                        assert flagged == 0, (tok, where)


def test_a_key_that_only_hashes_like_a_known_one_is_handed_back(built):
    """Keys are recognised by length + FNV-1a while walking and confirmed byte for byte
    afterwards: "enmfbqia" collides with "nodeName" and must not be taken for it."""
    from arena_b200 import HostSnapshot
    nodes = json.dumps({"items": [LT.make_node(n) for n in range(3)]}).encode()
    rng = random.Random(1)
    pod = LT.make_pod(1, rng, 3)
    pod["metadata"]["managedFields"] = []
    pod["spec"]["enmfbqia"] = "node-0002"
    del pod["spec"]["nodeName"]                            # (next to the real key it would count as a duplicate)
    text = json.dumps({"items": [pod]}).encode()
    hn = HostSnapshot.from_json(b'{"items":[]}', nodes)
    r = hn.scan_podlist(text, 4)
    assert r["flag"].tolist() == [11]                      # kFlagKeyHash
    ok, flagged = check_against_host(text, nodes, 1)
    assert ok and flagged == 1
    pod["metadata"]["annotations"]["enmfbqia"] = "x"       # where key names do not matter, nothing happens
    del pod["spec"]["enmfbqia"]
    ok, flagged = check_against_host(json.dumps({"items": [pod]}).encode(), nodes, 1)
    assert ok and flagged == 0
