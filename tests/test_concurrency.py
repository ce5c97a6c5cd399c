"""The reference's threading model on the device path: many request threads call topN while exactly
one updater thread applies "UP" messages (ServingLayer.java:225; ModelManagerListener.java:135-145;
ALSServingModel.java:161-183).  Readers must always get a well-formed answer that is consistent with
SOME state of the store: every returned row exists, scores are in descending order, and each score is
the exact dot product of the query with either the old or the new vector of that row."""
import threading

import numpy as np
import pytest

import oracle
import oryx_b200 as ob
from als_harness import random_hash_vectors, random_model

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("lsh", [False, True])
def test_many_readers_one_updater(lsh):
    n, f, hm = 40000, 32, 10
    y_old = random_model(n, f, 0xB2000090)
    y_new = random_model(n, f, 0xB2000091)
    hv = random_hash_vectors(3, f, 0xB2000092) if lsh else None
    m = ob.Model(f, hash_vectors=hv, max_bits_differing=1 if lsh else 0)
    m.upsert(ob.Y, np.arange(n, dtype=np.int32), y_old)
    queries = random_model(64, f, 0xB2000093)
    errors, answers = [], []
    stop = threading.Event()

    def reader(t):
        try:
            i = 0
            while not stop.is_set() or i < 50:
                q = queries[(t * 17 + i) % len(queries)]
                rows, scores = m.topn(ob.DOT, q, hm, excluded=[1, 2, 3] if i % 3 == 0 else None)
                answers.append((q, rows, scores))
                i += 1
                if i > 400:
                    break
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    def updater():
        try:
            rng = np.random.default_rng(0xB2000094)
            for _ in range(60):
                rows = rng.choice(n, 500, replace=False).astype(np.int32)
                m.upsert(ob.Y, rows, y_new[rows])          # vectors change (and, with LSH, partitions)
                gone = rng.choice(n, 20, replace=False).astype(np.int32)
                m.remove(ob.Y, gone)
                m.upsert(ob.Y, gone, y_new[gone])          # and come back
            m.compact(ob.Y)
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))
        finally:
            stop.set()

    threads = [threading.Thread(target=reader, args=(t,)) for t in range(6)] + [threading.Thread(target=updater)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=300)
    assert not errors, errors[:3]
    assert len(answers) > 300
    for q, rows, scores in answers:
        assert len(rows) <= hm and (len(rows) == hm or lsh)
        assert np.all(np.diff(scores) <= 0)
        assert len(set(rows.tolist())) == len(rows)
        for r, s in zip(rows, scores):
            assert 0 <= r < n
            assert s == oracle.dot(q, y_old[r]) or s == oracle.dot(q, y_new[r]), (r, s)
    # after the dust settles the store equals "old, overwritten by new where touched"; a final query
    # must equal the oracle on the read-back matrix
    final = np.stack([m.get(ob.Y, int(r)) for r in range(0, n, 97)])
    assert all(np.array_equal(v, y_old[r]) or np.array_equal(v, y_new[r]) for v, r in zip(final, range(0, n, 97)))
    m.close()


def test_coalesced_callers_get_the_single_query_answers():
    """oryx_topn_coalesced from 24 threads (mixed how_many, kinds and exclusion lists): every caller gets
    exactly what oryx_topn returns for its own request, whether it rode in a page or ran alone."""
    n, f = 60000, 48
    y = random_model(n, f, 0xB20000D0)
    queries = random_model(96, f, 0xB20000D1)
    m = ob.Model(f)
    m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
    m.topn_batch(ob.DOT, queries[:32], 10)   # the bf16 shadow exists: pages of any size take the batch path
    rng = np.random.default_rng(0xB20000D2)
    plan = []
    for i in range(len(queries)):
        kind = ob.DOT if i % 5 else ob.COSINE
        hm = [10, 10, 10, 25][i % 4]
        excl = rng.choice(n, int(rng.integers(0, 30)), replace=False).astype(np.int32) if i % 3 == 0 else None
        plan.append((kind, hm, excl))
    want = [m.topn(k, queries[i], hm, ex) for i, (k, hm, ex) in enumerate(plan)]
    got, errors = [None] * len(queries), []

    def caller(t):
        try:
            for i in range(t, len(queries), 24):
                k, hm, ex = plan[i]
                got[i] = m.topn_coalesced(k, queries[i], hm, ex)
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    for _ in range(3):   # a few rounds: different interleavings
        threads = [threading.Thread(target=caller, args=(t,)) for t in range(24)]
        for th in threads:
            th.start()
        for th in threads:
            th.join(timeout=120)
        assert not errors, errors[:3]
        for i, (w, g) in enumerate(zip(want, got)):
            assert g is not None and g[0].tolist() == w[0].tolist() and g[1].tobytes() == w[1].tobytes(), i
    with pytest.raises(ob.OryxError):
        m.topn_coalesced(ob.DOT, queries[0], 0)
    m.close()


def test_coalesced_recommend_by_user_row():
    """oryx_recommend_coalesced from 16 threads: each caller gets oryx_recommend's answer for its user,
    unknown users included."""
    n_items, n_users, f, hm = 50000, 200, 32, 10
    y = random_model(n_items, f, 0xB20000F0)
    x = random_model(n_users, f, 0xB20000F1)
    m = ob.Model(f)
    m.upsert(ob.Y, np.arange(n_items, dtype=np.int32), y)
    m.upsert(ob.X, np.arange(n_users, dtype=np.int32), x)
    rng = np.random.default_rng(0xB20000F2)
    m.known_items_add(np.arange(n_users), [rng.integers(0, n_items, int(k)).astype(np.int32) for k in rng.poisson(20, n_users)])
    # no batch has run yet: the first page the combiner sees builds the bf16 shadow itself
    users = list(range(0, n_users, 3)) + [n_users + 5]
    want = {u: m.recommend(u, hm) for u in users}
    got, errors = {}, []

    def caller(t):
        try:
            for u in users[t::16]:
                got[u] = m.recommend_coalesced(u, hm)
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=caller, args=(t,)) for t in range(16)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=120)
    assert not errors, errors[:3]
    for u in users:
        if want[u] is None:
            assert got[u] is None
        else:
            assert got[u][0].tolist() == want[u][0].tolist() and got[u][1].tobytes() == want[u][1].tobytes(), u
    m.close()
