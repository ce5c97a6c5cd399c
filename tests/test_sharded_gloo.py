"""N>1 host logic on CPU: world_size-2 gloo.  The per-rank scan and the merge are injected test doubles
built from the oracle, so this exercises what oryx_b200.sharded owns: the shard plan, update routing,
exclusion splitting, the exchange layout and global row ids."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from oryx_b200.sharded import ShardPlan, allreduce_gram, exchange_and_merge


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _host_merge(g_scores, g_rows, g_n, plan):
    # the reference's final sorted(DESC).limit(howMany) (ALSServingModel.java:334) over the gathered lists
    pairs = []
    for r in range(plan.world):
        for j in range(int(g_n[r])):
            pairs.append((float(g_scores[r, j]), int(plan.to_global(r, int(g_rows[r, j])))))
    pairs.sort(key=lambda t: (-t[0], t[1]))
    pairs = pairs[: g_scores.shape[1]]
    return np.array([p[1] for p in pairs], dtype=np.int64), np.array([p[0] for p in pairs])


def _worker(rank, world, port, n, f, how_many, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(1234)
        y = rng.standard_normal((n, f), dtype=np.float32)
        q = rng.standard_normal(f, dtype=np.float32)
        excluded_global = np.array([3, 77, n - 1, n // 2, n // 2 + 1])
        plan = ShardPlan.for_rows(n, world)
        lo, hi = plan.local_range(rank, n)
        shard = y[lo:hi]
        excl_local = plan.split_exclusions(rank, excluded_global)
        rows, scores = oracle.topn(oracle.DOT, q, shard, how_many, excluded=excl_local if len(excl_local) else None)
        ls = torch.zeros(how_many, dtype=torch.float64)
        lr = torch.zeros(how_many, dtype=torch.int32)
        ls[: len(scores)] = torch.from_numpy(scores)
        lr[: len(rows)] = torch.from_numpy(rows.astype(np.int32))
        ln = torch.tensor([len(rows)], dtype=torch.int32)
        m_rows, m_scores = exchange_and_merge(dist, torch, plan, how_many, ls, lr, ln, _host_merge)
        gram = torch.from_numpy(oracle.transpose_times_self(shard)) if hi > lo else torch.zeros((f, f), dtype=torch.float64)
        gram = allreduce_gram(dist, torch, gram)
        if rank == 0:
            out_q.put((m_rows, m_scores, gram.numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,how_many", [(1001, 10), (7, 10)])
def test_sharded_query_and_gram_world2(n, how_many):
    f, world = 16, 2
    ctx = mp.get_context("spawn")
    out_q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, f, how_many, out_q)) for r in range(world)]
    for p in procs:
        p.start()
    m_rows, m_scores, gram = out_q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(1234)
    y = rng.standard_normal((n, f), dtype=np.float32)
    q = rng.standard_normal(f, dtype=np.float32)
    excl = np.unique(np.array([3, 77, n - 1, n // 2, n // 2 + 1]) % max(n, 1)) if n < 78 else np.array([3, 77, n - 1, n // 2, n // 2 + 1])
    excl = excl[excl < n]
    e_rows, e_scores = oracle.topn(oracle.DOT, q, y, how_many, excluded=np.unique(excl))
    if n >= 78:
        assert np.array_equal(m_rows, e_rows) and np.array_equal(m_scores, e_scores)
    g = oracle.transpose_times_self(y)
    assert np.max(np.abs(gram - g)) <= 1e-12 * np.max(np.abs(g))


def test_shard_plan_routing():
    plan = ShardPlan.for_rows(10, 4)  # 3 rows per rank
    assert plan.rows_per_rank == 3 and plan.owner(9) == 3 and plan.local(9) == 0 and plan.to_global(2, 1) == 7
    assert plan.local_range(3, 10) == (9, 10) and plan.local_range(0, 10) == (0, 3)
    routed = plan.route(np.array([9, 0, 4, 5, 2]))
    assert sorted(routed) == [0, 1, 3]
    pos, loc = routed[1]
    assert list(pos) == [2, 3] and list(loc) == [1, 2]
    assert list(plan.split_exclusions(1, [0, 4, 9])) == [1]
    assert plan.split_exclusions(2, []).size == 0


def test_merge_batch_topn_orders_like_the_device_merge():
    """Host merge of per-shard batched answers: score descending, global row ascending, ragged counts."""
    from oryx_b200.sharded import merge_batch_topn
    hm = 4
    s0 = (np.array([[5, 1, 9, 0], [2, 0, 0, 0]], np.int32), np.array([[9.0, 7.0, 7.0, 1.0], [3.0, 0, 0, 0]]), np.array([4, 1], np.int32))
    s1 = (np.array([[0, 3, 0, 0], [0, 0, 0, 0]], np.int32), np.array([[8.0, 7.0, 0, 0], [0, 0, 0, 0]]), np.array([2, 0], np.int32))
    rows, scores, cnt = merge_batch_topn([s0, s1], row_stride=100, how_many=hm)
    assert cnt.tolist() == [4, 1]
    assert rows[0].tolist() == [5, 100, 1, 9] and scores[0].tolist() == [9.0, 8.0, 7.0, 7.0]  # ties: lower global row first
    assert rows[1, 0] == 2 and scores[1, 0] == 3.0
    with pytest.raises(ValueError):
        merge_batch_topn([s0, (s1[0], s1[1], np.array([2, -1], np.int32))], 100, hm)
