"""The N > 1 path on CPU: world_size-2 `gloo` processes run the sharding plan,
the totals all-reduce and the table gather of arena_b200.multigpu.  There is no
GPU here, so each rank's per-shard table comes from the C oracle -- the test is
about the plumbing: disjoint node ranges, no pod crossing ranks, one int64[2]
exchange, identical bits at any rank count."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, P, N, seed, skewed, out_dir):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from arena_b200 import HostSnapshot, multigpu
    from oracle import cref
    r, w, _ = multigpu.init_process_group("gloo")
    assert (r, w) == (rank, world)
    host = HostSnapshot.synth(P, N, seed=seed, skewed=skewed)        # every rank sees the same snapshot
    shard = host.shard(rank, world)
    tot, alloc, cl, _ = cref.top_node_columnar(shard.cols)
    g = multigpu.all_reduce_totals(torch.from_numpy(cl.copy()))
    full_total, full_alloc = multigpu.gather_tables(tot, alloc, dst=0)
    if rank == 0:
        np.savez(os.path.join(out_dir, "r0.npz"), cluster=g.numpy(), total=full_total, alloc=full_alloc,
                 pods=np.array([shard.n_pods]), nodes=np.array([shard.n_nodes]))
    else:
        assert full_total is None and full_alloc is None
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), cluster=g.numpy(), pods=np.array([shard.n_pods]),
                 nodes=np.array([shard.n_nodes]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("P,N,seed,skewed", [(200_000, 2_000, 1, False), (300_000, 500, 5, True)])
def test_two_rank_sharded_top_node(built, tmp_path, P, N, seed, skewed):
    from arena_b200 import HostSnapshot
    from oracle import cref
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), P, N, seed, skewed, str(tmp_path)), nprocs=world, join=True)
    host = HostSnapshot.synth(P, N, seed=seed, skewed=skewed)
    tot, alloc, cl, _ = cref.top_node_columnar(host.cols)
    r0 = np.load(tmp_path / "r0.npz")
    r1 = np.load(tmp_path / "r1.npz")
    assert np.array_equal(r0["cluster"], cl) and np.array_equal(r1["cluster"], cl)     # every rank has the totals
    assert np.array_equal(r0["total"], tot) and np.array_equal(r0["alloc"], alloc)     # gathered table == unsharded
    assert int(r0["pods"][0] + r1["pods"][0]) == P and int(r0["nodes"][0] + r1["nodes"][0]) == N


def test_shard_plan_properties(built):
    from arena_b200 import HostSnapshot
    host = HostSnapshot.synth(500_000, 3_000, seed=9, skewed=True)
    pn = host.cols["pod_node"]
    for G in (1, 2, 3, 4, 8):
        key_cut, pod_cut = host.shard_plan(G)
        assert key_cut[0] == 0 and key_cut[-1] == host.n_nodes and pod_cut[0] == 0 and pod_cut[-1] == host.n_pods
        assert np.all(np.diff(key_cut.astype(np.int64)) >= 0) and np.all(np.diff(pod_cut.astype(np.int64)) >= 0)
        for i in range(G):
            lo, hi = int(pod_cut[i]), int(pod_cut[i + 1])
            seg = pn[lo:hi]
            real = seg[seg != 0xFFFFFFFF]
            if len(real):                                  # no pod crosses a shard boundary
                assert real.min() >= key_cut[i] and real.max() < key_cut[i + 1]
            if i < G - 1:
                assert not np.any(seg == 0xFFFFFFFF)       # unmatched pods ride with the last shard
        # balance: no shard exceeds the ideal share by more than the largest node
        counts = np.bincount(pn[pn != 0xFFFFFFFF], minlength=host.n_nodes)
        sizes = np.diff(pod_cut.astype(np.int64))
        assert sizes[:-1].max(initial=0) <= host.n_pods / G + counts.max()


def test_shard_plan_rejects_ungrouped_input(built):
    from arena_b200 import ArenaError, HostSnapshot
    host = HostSnapshot.synth(100_000, 100, seed=3, grouped=False)
    with pytest.raises(ArenaError):
        host.shard_plan(2)


def test_job_range_shards_reproduce_top_job(built):
    """`top job` shards by job-id range (jobs span nodes): per-shard sums concatenate to the
    unsharded table and the totals add up."""
    from arena_b200 import HostSnapshot
    from oracle import cref
    J = 400
    host = HostSnapshot.synth(120_000, 800, seed=4, n_jobs=J)
    # group rows by job (stable) so that job ranges are contiguous pod ranges
    order = np.argsort(host.cols["pod_job"], kind="stable")
    cb = host.cols["pod_ctr_begin"]
    lens = (cb[1:] - cb[:-1])[order]
    new_cb = np.zeros(len(order) + 1, np.uint32)
    new_cb[1:] = np.cumsum(lens)
    idx = np.concatenate([np.arange(cb[p], cb[p + 1]) for p in order]) if len(order) else np.zeros(0, np.int64)
    cols = {k: host.cols[k][order] for k in ("pod_node", "pod_phase", "pod_job", "pod_role", "pod_created")}
    cols.update(pod_ctr_begin=new_cb, ctr_val=host.cols["ctr_val"][idx], ctr_scale=host.cols["ctr_scale"][idx],
                node_cap_val=host.cols["node_cap_val"], node_cap_scale=host.cols["node_cap_scale"])
    grouped = HostSnapshot.from_arrays(cols, n_jobs=J)
    req, alloc, latest, totals = cref.top_job_columnar(grouped.cols, J)
    parts = []
    for i in range(3):
        sh = grouped.shard(i, 3, by_job=True)
        r, a, l, t = cref.top_job_columnar(sh.cols, sh.n_jobs)
        parts.append((r, a, t))
    assert np.array_equal(np.concatenate([p[0] for p in parts]), req)
    assert np.array_equal(np.concatenate([p[1] for p in parts]), alloc)
    assert (sum(int(p[2][0]) for p in parts), sum(int(p[2][1]) for p in parts)) == (int(totals[0]), int(totals[1]))


def test_regroup_by_job_then_merge_orders(built):
    """The host logic of the multi-GPU `top job`: regroup a node-grouped snapshot by job, cut it into
    job-id ranges, and merge the per-shard orderings -- per-shard tables from the C oracle."""
    from arena_b200 import HostSnapshot, multigpu
    from oracle import cref
    J = 900
    host = HostSnapshot.synth(150_000, 700, seed=6, n_jobs=J)
    byjob = host.regroup(by_job=True)
    pj = byjob.cols["pod_job"].astype(np.int64)
    pj[pj == 0xFFFFFFFF] = J
    assert np.all(pj[:-1] <= pj[1:]), "rows must be grouped by job"
    # regrouping changes no sum and keeps list order inside a job (the latest-pod tie rule)
    req, alloc, latest, totals = cref.top_job_columnar(host.cols, J)
    req2, alloc2, latest2, totals2 = cref.top_job_columnar(byjob.cols, J)
    assert np.array_equal(req, req2) and np.array_equal(alloc, alloc2) and np.array_equal(totals, totals2)
    src = byjob.pod_order()
    has = latest != 0xFFFFFFFF
    assert np.array_equal(src[latest2[has]], latest[has])
    want_order = np.array(sorted(range(J), key=lambda j: (-int(req[j]), j)), np.uint32)
    for G in (1, 2, 5):
        key_cut, pod_cut = byjob.shard_plan(G, by_job=True)
        reqs, orders, tot = [], [], np.zeros(2, np.int64)
        for i in range(G):
            sh = byjob.shard(i, G, by_job=True)
            r, a, l, t = cref.top_job_columnar(sh.cols, sh.n_jobs)
            reqs.append(r)
            orders.append(np.array(sorted(range(sh.n_jobs), key=lambda j: (-int(r[j]), j)), np.uint32))
            tot += t
        assert np.array_equal(np.concatenate(reqs), req) and tot.tolist() == totals.tolist()
        assert np.array_equal(multigpu.merge_job_orders(reqs, orders, key_cut), want_order)
    # and back: grouped by node again, the node table is the original one
    again = byjob.regroup(by_job=False)
    assert np.array_equal(again.cols["pod_node"], host.cols["pod_node"])
    for x, y in zip(cref.top_node_columnar(again.cols, False)[:3], cref.top_node_columnar(host.cols, False)[:3]):
        assert np.array_equal(x, y)
