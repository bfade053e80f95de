"""The loader's structural scan + parallel item parsing against the
Python oracle (oracle/top_ref.py): every way a PodList can be laid out must give the
columns a sequential decode gives, and the same first error."""

import json
import os
import random
import sys

import numpy as np
import pytest

from oracle import cref, top_ref

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
import loader_timing as LT  # noqa: E402


def node_table(h):
    tot, alloc, cl, _ = cref.top_node_columnar(h.cols)
    return [(h.node_name(n), int(tot[n]), int(alloc[n])) for n in range(h.n_nodes)], cl.tolist()


def load_with_threads(monkeypatch, threads, pods, nodes, **kw):
    from arena_b200 import HostSnapshot
    monkeypatch.setenv("ARENA_LOADER_THREADS", str(threads))
    return HostSnapshot.from_json(pods, nodes, **kw)


NASTY_STRINGS = [
    'a"]}', "\\", '\\"', "]]]]", "{{{{", '{"x":[', "tab\there", "éé\U0001F600", "\\\\\\", 'q\\"q\\\\"', "",
]


def nasty_podlist(n_pods, n_nodes, seed, indent=None):
    """Strings full of brackets, quotes and backslash runs of every length, so quote and
    escape state crosses the scan's 64-byte block boundaries at every alignment."""
    rng = random.Random(seed)
    items = []
    for i in range(n_pods):
        pod = LT.make_pod(i, rng, n_nodes)
        pad = "\\" * rng.randrange(0, 9) + '"' * rng.randrange(0, 3) + "x" * rng.randrange(0, 70)
        pod["metadata"]["annotations"] = {rng.choice(NASTY_STRINGS) + str(k): rng.choice(NASTY_STRINGS) + pad
                                          for k in range(rng.randrange(0, 4))}
        pod["spec"]["containers"][0]["command"] = [rng.choice(NASTY_STRINGS) for _ in range(rng.randrange(0, 5))]
        pod["metadata"]["name"] = pod["metadata"]["name"] + rng.choice(["", '-"q"', "-\\", "-]"])
        items.append(pod)
    nodes = {"items": [LT.make_node(n) for n in range(n_nodes)]}
    return json.dumps({"items": items, "kind": "PodList"}, indent=indent).encode(), json.dumps(nodes).encode()


@pytest.mark.parametrize("indent", [None, 1])
def test_parallel_split_matches_the_python_oracle(built, monkeypatch, indent):
    pods, nodes = nasty_podlist(2600, 37, seed=5, indent=indent)
    _, table = top_ref.top_node(pods, nodes)
    want = [tuple(r) for r in table]
    names = [p["metadata"]["name"] for p in json.loads(pods)["items"]]
    cols = None
    for threads in (1, 3, 8):
        h = load_with_threads(monkeypatch, threads, pods, nodes)
        got, _ = node_table(h)
        assert got == want
        order = h.pod_order()
        assert [h.pod_name(r) for r in range(h.n_pods)] == [names[k] for k in order]
        if cols is None:
            cols = {k: v.copy() for k, v in h.cols.items()}
        else:
            for k, v in h.cols.items():
                assert np.array_equal(v, cols[k]), k


def test_split_and_in_place_parse_agree(built, monkeypatch):
    """A scalar element makes the array unsplittable: the in-place (sequential) parse
    must then see exactly what the split path sees without it."""
    pods, nodes = nasty_podlist(600, 11, seed=9)
    doc = json.loads(pods)
    h = load_with_threads(monkeypatch, 4, pods, nodes)
    # same items behind a duplicate "items" key: the LAST one wins, the first is still validated
    dup = b'{"items": [{"spec": {"nodeName": "node-0000", "containers": [{"resources": {"limits": {"nvidia.com/gpu": "7"}}}]}}], ' \
        + pods[1:]
    h2 = load_with_threads(monkeypatch, 4, dup, nodes)
    for k, v in h.cols.items():
        assert np.array_equal(v, h2.cols[k]), k
    from arena_b200 import ArenaError
    bad_first = b'{"items": [{"spec": {"nodeName": 5}}], ' + pods[1:]      # dropped by last-wins, so no field error ...
    h3 = load_with_threads(monkeypatch, 4, bad_first, nodes)
    assert h3.n_pods == h.n_pods
    with pytest.raises(ArenaError) as ei:                                    # ... but damage in it is still a syntax error
        load_with_threads(monkeypatch, 4, b'{"items": [{"spec": tru}], ' + pods[1:], nodes)
    assert "JSON syntax error at byte 20" in str(ei.value)
    doc["items"].append(17)
    with pytest.raises(ArenaError) as ei:
        load_with_threads(monkeypatch, 4, json.dumps(doc).encode(), nodes)
    assert "PodList item is not an object" in str(ei.value)


def test_first_error_is_the_one_a_sequential_decode_reports(built, monkeypatch):
    from arena_b200 import ArenaError
    pods, nodes = nasty_podlist(3000, 11, seed=3)
    doc = json.loads(pods)
    body = json.dumps(doc).encode()
    # two damaged places: the earlier offset must be reported whatever thread finds it
    marks = [m for m in range(len(body)) if body.startswith(b'"restartPolicy": "Never"', m)]
    a, b = marks[700], marks[2500]
    broken = bytearray(body)
    for m in (a, b):
        broken[m + 17:m + 24] = b'nEver" '          # "restartPolicy": nEver"  -> invalid literal at m+17
    for threads in (1, 8):
        with pytest.raises(ArenaError) as ei:
            load_with_threads(monkeypatch, threads, bytes(broken), nodes)
        assert f"PodList: JSON syntax error at byte {a + 17}:" in str(ei.value), str(ei.value)
    # field errors: lowest item index wins; a syntax error anywhere outranks them
    doc["items"][2100]["spec"]["nodeName"] = 5
    doc["items"][400]["status"]["phase"] = ["Running"]
    for threads in (1, 8):
        with pytest.raises(ArenaError) as ei:
            load_with_threads(monkeypatch, threads, json.dumps(doc).encode(), nodes)
        assert 'field "phase" is not a string' in str(ei.value)
    with pytest.raises(ArenaError) as ei:
        load_with_threads(monkeypatch, 8, json.dumps(doc).encode()[:-1] + b"]", nodes)
    assert "JSON syntax error" in str(ei.value)
    # a bad field in a NODE is reported only if the PodList is at least well-formed
    bad_nodes = json.loads(nodes)
    bad_nodes["items"][3]["status"]["capacity"]["nvidia.com/gpu"] = "1x"
    with pytest.raises(ArenaError) as ei:
        load_with_threads(monkeypatch, 8, pods, json.dumps(bad_nodes).encode())
    assert "invalid quantity" in str(ei.value)
    with pytest.raises(ArenaError) as ei:
        load_with_threads(monkeypatch, 8, bytes(broken), json.dumps(bad_nodes).encode())
    assert "PodList: JSON syntax error" in str(ei.value)


def test_truncated_documents_fail_cleanly(built, monkeypatch):
    from arena_b200 import ArenaError
    pods, nodes = nasty_podlist(300, 5, seed=1)
    rng = random.Random(0)
    for cut in sorted(rng.sample(range(1, len(pods) - 1), 60)):
        with pytest.raises(ArenaError) as ei:
            load_with_threads(monkeypatch, 4, pods[:cut], nodes)
        assert ei.value.code == -6
