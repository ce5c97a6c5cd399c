"""Row-sharded queries over peer memory (oryx_exchange_* / oryx_topn_sharded_device): the publish /
wait / merge kernel against the oracle's top-N over the concatenated shards.  World 1 runs on any GPU
box; the 2-GPU cases need two visible devices (gpurun --gpus 2) and are skipped otherwise.  The
inter-process (CUDA IPC) connection is exercised by `bench.py --gpus N`."""
import numpy as np
import pytest
import torch

import oracle
import oryx_b200 as ob
from als_harness import random_hash_vectors, random_model

pytestmark = pytest.mark.gpu


def _run(world, n_per, f, hm, steps, kind=ob.DOT, seed=0xB2000050, lsh=None, compact=False):
    ys = [random_model(n_per, f, seed + r) for r in range(world)]
    y_all = np.concatenate(ys)
    models, xs, outs, streams, dqs = [], [], [], [], []
    qs = random_model(steps, f, seed + 100)
    hv, bits, h = None, 0, 0
    if lsh is not None:                      # (sample_rate, cores): LSH candidate filtering on every shard
        h, bits = ob.lsh_plan(*lsh)
        hv = random_hash_vectors(h, f, seed=seed + 7)
    for r in range(world):
        with torch.cuda.device(r):
            m = ob.Model(f, device=r, hash_vectors=hv, max_bits_differing=bits)
            m.upsert(ob.Y, np.arange(n_per, dtype=np.int32), ys[r])
            if compact:
                m.compact(ob.Y)
            models.append(m)
            xs.append(ob.Exchange(r, world, r, max(hm, 16)))
            dev = torch.device("cuda", r)
            outs.append([(torch.zeros(hm, dtype=torch.int64, device=dev), torch.zeros(hm, dtype=torch.float64, device=dev),
                          torch.zeros(1, dtype=torch.int32, device=dev)) for _ in range(steps)])
            streams.append(torch.cuda.Stream(device=dev))
            dqs.append(torch.from_numpy(qs).to(dev))
    ob.Exchange.connect_local(xs)
    # every rank enqueues all its queries back to back (no host synchronisation in between): the
    # exchange kernels of different ranks meet on the device
    for i in range(steps):
        for r in range(world):
            rows, scores, n = outs[r][i]
            xs[r].topn(models[r], kind, dqs[r][i].data_ptr(), hm, n_per, rows.data_ptr(), scores.data_ptr(),
                       n.data_ptr(), streams[r].cuda_stream)
    for r in range(world):
        streams[r].synchronize()
    part_all = oracle.lsh_index_rows(hv, y_all) if lsh is not None else None
    for i in range(steps):
        cand = oracle.lsh_candidates(h, bits, oracle.lsh_index(hv, qs[i])) if lsh is not None else None
        want_rows, want_scores = oracle.topn(kind, qs[i], y_all, hm, part=part_all, cand=cand)
        for r in range(world):
            rows, scores, n = outs[r][i]
            k = int(n.item())
            assert k == len(want_rows), (i, r, k)
            assert rows[:k].cpu().numpy().tolist() == want_rows.tolist(), (i, r)
            assert scores[:k].cpu().numpy().tobytes() == want_scores.tobytes(), (i, r)
    for x in xs:
        x.close()
    for m in models:
        m.close()


def test_exchange_world_1_is_the_local_topn():
    _run(1, 30000, 32, 10, steps=6)


def test_exchange_world_1_short_shard_and_cosine():
    _run(1, 7, 16, 10, steps=3, kind=ob.COSINE)  # fewer rows than how_many


@pytest.mark.parametrize("compact", [False, True])
def test_exchange_world_1_with_lsh_filtering(compact):
    _run(1, 40000, 32, 10, steps=8, lsh=(0.3, 32), compact=compact)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("hm", [10, 100])
def test_exchange_two_gpus_matches_oracle(hm):
    _run(2, 50000, 64, hm, steps=40)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("compact", [False, True])
def test_exchange_two_gpus_with_lsh_filtering(compact):
    """LSH candidate filtering AND the sharded exchange together: every shard hashes the query itself and scans
    only its candidate partitions; the merged list equals the oracle's over the concatenated shards."""
    _run(2, 60000, 48, 10, steps=24, lsh=(0.3, 32), compact=compact)
    _run(2, 30000, 64, 10, steps=8, kind=ob.COSINE, lsh=(0.5, 8), compact=compact)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs four GPUs")
def test_exchange_four_gpus_matches_oracle():
    _run(4, 20000, 48, 10, steps=40)


def _sharded_fold_in(world):
    """X and Y sharded by rows over `world` GPUs of this process: the sharded fold-in (rows of other
    shards read over NVLink) must be bit-identical to the single-model one, and the sum of the shards'
    Grams must be the Gram of the whole."""
    f, nu, ni, n = 24, 900, 1200, 4000
    rng = np.random.default_rng(0xB20000B0)
    x = (rng.standard_normal((nu, f)) * 0.3).astype(np.float32)
    y = (rng.standard_normal((ni, f)) * 0.3).astype(np.float32)
    xs, ys = -(-nu // world), -(-ni // world)   # rows per shard
    whole = ob.Model(f, device=0)
    whole.upsert(ob.X, np.arange(nu, dtype=np.int32), x)
    whole.upsert(ob.Y, np.arange(ni, dtype=np.int32), y)
    shards = []
    for r in range(world):
        m = ob.Model(f, device=r)
        lo, hi = r * xs, min(nu, (r + 1) * xs)
        m.upsert(ob.X, np.arange(hi - lo, dtype=np.int32), x[lo:hi])
        lo, hi = r * ys, min(ni, (r + 1) * ys)
        m.upsert(ob.Y, np.arange(hi - lo, dtype=np.int32), y[lo:hi])
        shards.append(m)
    shards[-1].remove(ob.Y, np.array([3], np.int32))          # a hole in the last shard ...
    whole.remove(ob.Y, np.array([(world - 1) * ys + 3], np.int32))  # ... and the same row in the whole
    g_whole = whole.gram(ob.Y)[0]
    g_sum = sum(m.gram(ob.Y)[0] for m in shards)
    assert np.max(np.abs(g_sum - g_whole)) <= 1e-12 * np.max(np.abs(g_whole))
    yty = ob.Solver(g_whole, device=0)
    ur = rng.integers(-1, nu + 30, n).astype(np.int64)
    ir = rng.integers(0, ni + 30, n).astype(np.int64)
    ir[:3] = (world - 1) * ys + 3
    vals = rng.choice([1.0, 2.0, -1.0, 0.5], n)
    for implicit in (True, False):
        want, wv = yty.fold_in_rows(whole, implicit, ob.X, ur.astype(np.int32), ir.astype(np.int32), vals)
        got, gv = yty.fold_in_rows_sharded(shards, xs, ys, implicit, ob.X, ur, ir, vals)
        assert np.array_equal(gv, wv)
        assert np.array_equal(got[gv == 1], want[wv == 1])
    for m in shards + [whole]:
        m.close()


def test_sharded_fold_in_world_1():
    _sharded_fold_in(1)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_sharded_fold_in_two_gpus():
    _sharded_fold_in(2)


def test_sharded_batch_merges_to_the_unsharded_answer():
    """A page of queries over Y split in two shards (two models; both on GPU 0 if only one GPU is there):
    each shard answers through the tensor-core batch path, the host merges -- equal to the oracle over
    the concatenated rows."""
    from oryx_b200.sharded import merge_batch_topn
    n_per, f, hm = 40000, 64, 15
    ys = [random_model(n_per, f, 0xB2000120 + r) for r in range(2)]
    q = random_model(40, f, 0xB2000122)
    res = []
    for r in range(2):
        dev = r if torch.cuda.device_count() > 1 else 0
        with ob.Model(f, device=dev) as m:
            m.upsert(ob.Y, np.arange(n_per, dtype=np.int32), ys[r])
            res.append(m.topn_batch(ob.DOT, q, hm))
    rows, scores, cnt = merge_batch_topn(res, n_per, hm)
    y_all = np.concatenate(ys)
    for i in range(len(q)):
        wr, ws = oracle.topn(ob.DOT, q[i], y_all, hm)
        assert cnt[i] == len(wr) and rows[i, :cnt[i]].tolist() == wr.tolist() and scores[i, :cnt[i]].tobytes() == ws.tobytes()
    # the same merge on the device: one launch for the whole page (oryx_merge_topn_batch_device)
    from oryx_b200.sharded import merge_batch_topn_device
    d_rows, d_scores, d_cnt = merge_batch_topn_device(ob, torch, res, n_per, hm, torch.device("cuda", 0))
    torch.cuda.synchronize()
    assert np.array_equal(d_cnt.cpu().numpy(), cnt)
    for i in range(len(q)):
        k = int(cnt[i])
        assert np.array_equal(d_rows[i, :k].cpu().numpy(), rows[i, :k])
        assert d_scores[i, :k].cpu().numpy().tobytes() == scores[i, :k].tobytes()


def _gram_allreduce(world):
    """oryx_gram_sharded / oryx_gram_allreduce_device: per-shard Grams summed by one kernel over peer memory equal
    the oracle's Gram of the whole matrix; every rank holds the same bits."""
    import threading
    f, n_per = 48, 30000
    ys = [random_model(n_per, f, 0xB2000700 + r) for r in range(world)]
    models, xs = [], []
    for r in range(world):
        with torch.cuda.device(r):
            m = ob.Model(f, device=r)
            m.upsert(ob.Y, np.arange(n_per, dtype=np.int32), ys[r])
            models.append(m)
            xs.append(ob.Exchange(r, world, r, 16))
    ob.Exchange.connect_local(xs)
    want = oracle.transpose_times_self(np.concatenate(ys))
    for rep in range(3):   # parities alternate; the third call reuses the first one's slots
        got = [None] * world
        def run(r):
            got[r] = xs[r].gram_sharded(models[r], ob.Y)
        ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        for r in range(world):
            g, used = got[r]
            assert used == n_per
            assert np.max(np.abs(g - want)) <= 1e-12 * np.max(np.abs(want))
            assert g.tobytes() == got[0][0].tobytes()      # rank order of the additions: identical everywhere
    # the device form, on matrices of the caller
    d = [torch.full((f, f), float(r + 1), dtype=torch.float64, device=torch.device("cuda", r)) for r in range(world)]
    st = [torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", r)) for r in range(world)]
    for r in range(world):
        with torch.cuda.device(r):
            xs[r].gram_allreduce_device(d[r].data_ptr(), f, st[r].data_ptr(), torch.cuda.current_stream().cuda_stream)
    for r in range(world):
        with torch.cuda.device(r):
            torch.cuda.synchronize()
        assert int(st[r].item()) == 0
        assert torch.all(d[r] == world * (world + 1) / 2).item()
    for x in xs:
        x.close()
    for m in models:
        m.close()


def test_gram_allreduce_world_1():
    _gram_allreduce(1)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_gram_allreduce_two_gpus():
    _gram_allreduce(2)


@pytest.mark.skipif(torch.cuda.device_count() < 4, reason="needs four GPUs")
def test_gram_allreduce_four_gpus():
    _gram_allreduce(4)
