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
