"""Multi-GPU parity on the box: tools/multigpu_check.py under torchrun with as many ranks as there
are GPUs (at most 8).  On a one-GPU box it still runs, with a world of one (the exchange talks to
itself); the gloo tests in test_multigpu_gloo.py cover the host logic at world size 2."""

import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_rank_agrees_with_the_oracle(gpu):
    import torch
    n = min(8, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "multigpu_check.py"), "3000000"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, (p.stdout[-3000:], p.stderr[-3000:])
    assert "FAIL" not in p.stdout and p.stdout.count("OK") >= 12


def _n_devices():
    import torch
    return min(8, torch.cuda.device_count())


def test_single_process_multi_device_top_node(gpu):
    """arena_acct_init_devices + arena_top_node_multi: one process drives every GPU of the box (one on
    a one-GPU box), node-range shards, against the C oracle: uniform, skewed, and the details view."""
    import numpy as np
    from oracle import cref
    gpu.init_devices(list(range(_n_devices())))
    assert gpu.lib().arena_acct_device_count() == _n_devices()
    for skewed, seed in ((False, 41), (True, 5)):
        host = gpu.HostSnapshot.synth(3_000_000, 30_000, seed=seed, skewed=skewed)
        tot, alloc, cl, pg = cref.top_node_columnar(host.cols, True)
        ms = gpu.MultiSnapshot(host)
        assert ms.n_shards == _n_devices() and int(ms.key_cut[-1]) == host.n_nodes and int(ms.pod_cut[-1]) == host.n_pods
        for details in (False, True, False):
            r = gpu.top_node_multi(ms, details=details)
            assert np.array_equal(r.node_total, tot) and np.array_equal(r.node_alloc, alloc) and np.array_equal(r.cluster, cl)
            if details:
                assert np.array_equal(r.pod_gpu, pg)
        ms.close()
    # the single-device API still works beside it (slot 0)
    r = gpu.top_node(gpu.DeviceSnapshot.upload(host))
    assert np.array_equal(r.node_alloc, alloc)


def test_single_process_multi_device_top_job(gpu):
    import numpy as np
    from oracle import cref
    gpu.init_devices(list(range(_n_devices())))
    J = 70_000
    host = gpu.HostSnapshot.synth(2_000_000, 20_000, seed=43, n_jobs=J)
    req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
    byjob = host.regroup(by_job=True)
    src = byjob.pod_order()
    ms = gpu.MultiSnapshot(byjob, by_job=True)
    for _ in range(2):
        r = gpu.top_job_multi(ms)
        assert np.array_equal(r.job_req, req) and np.array_equal(r.job_alloc, alloc) and np.array_equal(r.totals, totals)
        has = latest != 0xFFFFFFFF
        assert np.array_equal(r.job_latest != 0xFFFFFFFF, has) and np.array_equal(src[r.job_latest[has]], latest[has])
        assert np.array_equal(r.job_order, np.lexsort((np.arange(J), -req)).astype(np.uint32))
    ms.close()
    # a node-range snapshot is refused by the job call, and the other way round
    ms = gpu.MultiSnapshot(host)
    with pytest.raises(gpu.ArenaError):
        gpu.top_job_multi(ms)
    ms.close()


def test_cli_gpus_flag_prints_the_golden_text(gpu):
    """arena-top node|job --gpus N: the config-1 fixture through the multi-device path, byte for byte."""
    cli = os.path.join(ROOT, "arena_b200", "arena-top")
    g = os.path.join(ROOT, "tests", "golden", "config1")
    n = str(max(1, _n_devices()))
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_fixtures as MF
    for args, want in (
        (["node", "--gpus", n, "--loader", "host"], "expected_top_node.txt"),
        (["node", "-d", "--gpus", n], "expected_top_node_details.txt"),
        (["job", "--gpus", n, "--jobs", os.path.join(g, "jobs.json"), "--tfjobs", os.path.join(g, "tfjobs.json"),
          "--releases", os.path.join(g, "helm_list.txt"), "--now", str(MF.NOW)], "expected_top_job.txt"),
    ):
        p = subprocess.run([cli, *args, "--pods", os.path.join(g, "pods.json"), "--nodes", os.path.join(g, "nodes.json")],
                           capture_output=True, text=True, timeout=120)
        assert p.returncode == 0, p.stdout + p.stderr
        assert p.stdout == open(os.path.join(g, want)).read(), args


def test_shutdown_and_reinit_in_a_fresh_process(gpu):
    """arena_acct_shutdown drops every device context (and the multi-device driver's memory of them);
    a later arena_acct_init_devices starts from scratch.  In a child process: the session's own
    library state must stay as it is."""
    code = (
        "import numpy as np, arena_b200 as ab\n"
        "from oracle import cref\n"
        "h = ab.HostSnapshot.synth(400_000, 4_000, seed=3)\n"
        "tot, alloc, cl, _ = cref.top_node_columnar(h.cols)\n"
        "for round in range(2):\n"
        "    ab.init_devices([0])\n"
        "    ms = ab.MultiSnapshot(h)\n"
        "    r = ab.top_node_multi(ms)\n"
        "    assert np.array_equal(r.node_alloc, alloc) and np.array_equal(r.cluster, cl)\n"
        "    r = ab.top_node_packed(h.pack())\n"
        "    assert np.array_equal(r.node_alloc, alloc)\n"
        "    ms.close()\n"
        "    ab.lib().arena_acct_shutdown()\n"
        "    assert ab.lib().arena_acct_device_count() == 0\n"
        "print('reinit ok')\n")
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, cwd=ROOT,
                       env=dict(os.environ, PYTHONPATH=ROOT))
    assert p.returncode == 0 and "reinit ok" in p.stdout, (p.stdout[-1000:], p.stderr[-2000:])


def test_multi_device_edge_shapes(gpu):
    """No pods at all, fewer nodes than devices, every pod on one node, pods that name no node: the cut may
    leave devices with empty ranges; the table and the totals must not notice."""
    import numpy as np
    from oracle import cref
    gpu.init_devices(list(range(_n_devices())))
    z = lambda dt, n=0: np.zeros(n, dt)
    rng = np.random.default_rng(4)

    def snap(pod_node, n_nodes):
        P = len(pod_node)
        counts = rng.integers(0, 4, P)
        cb = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint32)
        C = int(cb[-1])
        return gpu.HostSnapshot.from_arrays(dict(
            pod_node=np.asarray(pod_node, np.uint32), pod_phase=rng.choice(np.array([1, 2, 3, 4], np.uint8), P) if P else z(np.uint8),
            pod_ctr_begin=cb, ctr_val=rng.integers(0, 9, C).astype(np.int64), ctr_scale=z(np.int8, C),
            node_cap_val=rng.integers(0, 9, n_nodes).astype(np.int64), node_cap_scale=z(np.int8, n_nodes)))

    cases = [snap([], 3), snap([], 0), snap([0] * 700, 1), snap([0] * 5 + [0xFFFFFFFF] * 9, 2),
             snap(np.sort(rng.integers(0, 3, 4000)), 3), snap([0xFFFFFFFF] * 50, 4)]
    for h in cases:
        tot, alloc, cl, pg = cref.top_node_columnar(h.cols, True)
        ms = gpu.MultiSnapshot(h)
        for details in (False, True):
            r = gpu.top_node_multi(ms, details=details)
            assert np.array_equal(r.node_total, tot) and np.array_equal(r.node_alloc, alloc) and np.array_equal(r.cluster, cl)
            if details:
                assert np.array_equal(r.pod_gpu, pg)
        ms.close()
