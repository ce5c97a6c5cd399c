"""Known items resident on the device and /recommend by user row:
oryx_known_items_add / _get / _retain / oryx_known_counts / oryx_recommend against the reference's
goldens (RecommendTest.java, ALSServingModelTest.java:58-73) and the CPU oracle's top-N with the same
exclusion sets."""
import numpy as np
import pytest

import oryx_b200 as ob
from als_harness import load_golden, random_hash_vectors, random_model
import oracle

pytestmark = pytest.mark.gpu


def _fixed_model():
    g = load_golden("als_fixed_model")["model"]
    x = np.array(g["X"], dtype=np.float32)
    y = np.array(g["Y"], dtype=np.float32)
    known = np.array(g["known"])
    m = ob.Model(g["features"], implicit=g["implicit"])
    m.upsert(ob.X, np.arange(len(x), dtype=np.int32), x)
    m.upsert(ob.Y, np.arange(len(y), dtype=np.int32), y)
    m.known_items_add(np.arange(len(x)), [np.nonzero(known[u] > 0)[0] for u in range(len(x))])
    return m, x, y, known


def test_recommend_goldens_from_device_state():
    m, x, y, known = _fixed_model()
    for c in load_golden("als_fixed_model")["recommend"]:
        if "count" not in c or c.get("rescorer"):
            continue
        u = int(c["user"][1:])
        rows, scores = m.recommend(u, 10, consider_known_items=bool(c.get("consider_known")))
        assert len(rows) == c["count"], c["_cite"]
        assert "I%d" % rows[0] == c["top_id"], c["_cite"]
        assert abs(scores[0] - c["top_value"]) <= c["tol"], c["_cite"]
    # RecommendTest.java:46-50 testHowMany(/recommend/U5): 10 -> 2, 2 -> 2, 1 -> 1
    assert [len(m.recommend(5, h)[0]) for h in (10, 2, 1)] == [2, 2, 1]
    assert m.recommend(99, 10) is None  # unknown user: the reference's 404
    m.close()


def test_known_items_set_semantics_and_counts():
    # ALSServingModelTest.java:58-73 testUserItemCounts with populateKnownItems (:142-152)
    m = ob.Model(2)
    for i in range(10):
        for j in range(10):
            if abs(i - j) <= 1:
                m.known_items_add([i], [[j]])  # one singleton at a time, as the test does
    m.known_items_add([0, 0], [[1, 1, 0], [0]])  # duplicates are ignored (Set.addAll)
    assert m.known_items_get(0).tolist() == [0, 1]
    assert m.known_items_get(1).tolist() == [0, 1, 2]
    assert m.known_items_get(9).tolist() == [8, 9]
    assert m.known_items_get(1234).tolist() == []
    uc = m.known_counts(ob.X, 10)
    assert (uc[0], uc[1], uc[9]) == (2, 3, 2)
    ic = m.known_counts(ob.Y, 10)
    assert (ic[0], ic[1], ic[9]) == (2, 3, 2)
    # ALSServingModelTest.java:105-132 testRetainKnown, second half (no recent ids left)
    keep_u = np.zeros(10, np.uint8); keep_u[[4, 5, 6]] = 1
    keep_i = np.zeros(10, np.uint8); keep_i[[4, 5, 6]] = 1
    m.known_items_retain(keep_u, keep_i)
    assert m.known_items_get(3).tolist() == [] and m.known_items_get(2).tolist() == []
    assert m.known_items_get(4).tolist() == [4, 5] and m.known_items_get(6).tolist() == [5, 6]
    assert m.known_counts(ob.Y, 10).tolist() == [0, 0, 0, 0, 2, 3, 2, 0, 0, 0]
    m.close()


@pytest.mark.parametrize("lsh", [False, True])
def test_recommend_matches_oracle_with_device_resident_exclusions(lsh):
    n_items, n_users, f, hm = 60000, 300, 48, 25
    y = random_model(n_items, f, 0xB2000030)
    x = random_model(n_users, f, 0xB2000031)
    hv = random_hash_vectors(4, f, 0xB2000032) if lsh else None
    m = ob.Model(f, hash_vectors=hv, max_bits_differing=1 if lsh else 0)
    m.upsert(ob.Y, np.arange(n_items, dtype=np.int32), y)
    m.upsert(ob.X, np.arange(n_users, dtype=np.int32), x)
    part = oracle.lsh_index_rows(hv, y) if lsh else None
    rng = np.random.default_rng(0xB2000033)
    known = [set() for _ in range(n_users)]

    def add_round(max_len):
        users, lists = [], []
        for u in rng.choice(n_users, n_users // 2, replace=False):
            k = int(rng.integers(0, max_len))
            # bias half of the list towards the user's best items so exclusions actually bite
            top = np.argsort(-(y[:4000] @ x[u]))[: k // 2]
            items = np.concatenate([top, rng.integers(0, n_items + 50, k - k // 2)])  # some rows have no vector
            users.append(int(u))
            lists.append(items.astype(np.int32))
            known[u].update(int(i) for i in items)
        m.known_items_add(users, lists)

    def check(users):
        for u in users:
            assert m.known_items_get(u).tolist() == sorted(known[u])
            cand = None
            if lsh:
                cand = oracle.lsh_candidates(4, 1, oracle.lsh_index(hv, x[u]))
            dead = set(np.nonzero(~alive)[0].tolist())  # removed rows: the oracle sees them as excluded
            excl = np.array(sorted({i for i in known[u] if i < n_items} | dead), dtype=np.int32)
            want_rows, want_scores = oracle.topn(ob.DOT, x[u], y_now, hm, part=part_now, cand=cand, excluded=excl)
            rows, scores = m.recommend(u, hm)
            assert rows.tolist() == want_rows.tolist(), u
            assert scores.tobytes() == want_scores.tobytes(), u
            rows2, _ = m.recommend(u, hm, consider_known_items=True)
            w2, _ = oracle.topn(ob.DOT, x[u], y_now, hm, part=part_now, cand=cand,
                                excluded=np.array(sorted(dead), dtype=np.int32))
            assert rows2.tolist() == w2.tolist()

    alive = np.ones(n_items, bool)
    y_now, part_now = y, part
    add_round(40)
    add_round(400)       # relocations: segments outgrow their capacity
    known[7].update(range(0, 5000))  # longer than the shared-memory sort: host translation path
    m.known_items_add([7], [np.arange(5000, dtype=np.int32)])
    check([0, 7, 11] + [int(u) for u in rng.choice(n_users, 12, replace=False)])
    # tombstones and re-used slots: removed items drop out of the exclusion list and of the results
    gone = rng.choice(n_items, 3000, replace=False).astype(np.int32)
    m.remove(ob.Y, gone)
    alive[gone] = False
    check([7] + [int(u) for u in rng.choice(n_users, 6, replace=False)])
    back = gone[:1000]
    y_now = y.copy()
    y_now[back] = random_model(1000, f, 0xB2000034)
    m.upsert(ob.Y, back, y_now[back])
    alive[back] = True
    if lsh:
        part_now = oracle.lsh_index_rows(hv, y_now)
    check([int(u) for u in rng.choice(n_users, 6, replace=False)])
    m.compact(ob.Y)      # slots move: the device row->slot map must follow
    check([7] + [int(u) for u in rng.choice(n_users, 6, replace=False)])
    # counts against numpy
    ic = m.known_counts(ob.Y, n_items + 50)
    want = np.zeros(n_items + 50, np.int64)
    for s in known:
        for i in s:
            want[i] += 1
    assert ic.tolist() == want.tolist()
    assert m.known_counts(ob.X, n_users).tolist() == [len(s) for s in known]
    m.close()


def test_config_c1_recommend_parity():
    """BASELINE.json configs[0] -- the reference's own CPU-runnable case: 100 k users x 100 k items,
    50 features, /recommend howMany=10, LSH sample-rate 1.0, Poisson(20) known items per user
    (LoadTestALSModelFactory.java:38-51, 62-95).  Every sampled user's list must equal the oracle's,
    scores bit for bit."""
    n_users = n_items = 100_000
    f, hm = 50, 10
    y = random_model(n_items, f, 0xB2000080)
    x = random_model(n_users, f, 0xB2000081)
    hashes, bits = ob.lsh_plan(1.0, 8)  # sample-rate 1.0: every partition is a candidate
    hv = random_hash_vectors(hashes, f, 0xB2000082)
    m = ob.Model(f, hash_vectors=hv, max_bits_differing=bits)
    m.upsert(ob.Y, np.arange(n_items, dtype=np.int32), y)
    m.upsert(ob.X, np.arange(n_users, dtype=np.int32), x)
    rng = np.random.default_rng(0xB2000083)
    counts = rng.poisson(20, n_users)
    flat = rng.integers(0, n_items, int(counts.sum())).astype(np.int32)
    begin = np.concatenate([[0], np.cumsum(counts)])
    m.known_items_add(np.arange(n_users), [flat[begin[u]:begin[u + 1]] for u in range(n_users)])
    for u in rng.choice(n_users, 200, replace=False):
        known = np.unique(flat[begin[u]:begin[u + 1]])
        want_rows, want_scores = oracle.topn(ob.DOT, x[u], y, hm, excluded=known)
        rows, scores = m.recommend(int(u), hm)
        assert rows.tolist() == want_rows.tolist(), u
        assert scores.tobytes() == want_scores.tobytes(), u
    m.close()


def test_recommend_batch_equals_one_at_a_time():
    """oryx_recommend_batch (queries gathered from the device X store, scored on the tensor cores, each
    against its own device-resident known items) returns exactly what oryx_recommend returns user by
    user -- including users without a vector and users whose known items were removed from Y."""
    n_items, n_users, f, hm = 70000, 400, 64, 20
    y = random_model(n_items, f, 0xB20000A0)
    x = random_model(n_users, f, 0xB20000A1)
    m = ob.Model(f)
    m.upsert(ob.Y, np.arange(n_items, dtype=np.int32), y)
    m.upsert(ob.X, np.arange(n_users, dtype=np.int32), x)
    rng = np.random.default_rng(0xB20000A2)
    lists = []
    for u in range(n_users):
        top = np.argsort(-(y[:3000] @ x[u]))[: int(rng.integers(0, 15))]
        lists.append(np.concatenate([top, rng.integers(0, n_items, int(rng.poisson(20)))]).astype(np.int32))
    m.known_items_add(np.arange(n_users), lists)
    m.remove(ob.Y, rng.choice(n_items, 500, replace=False).astype(np.int32))
    m.remove(ob.X, np.array([5, 77], np.int32))
    users = np.concatenate([rng.permutation(n_users)[:300], [n_users + 9, -1]]).astype(np.int32)
    for consider in (False, True):
        rows, scores, cnt = m.recommend_batch(users, hm, consider_known_items=consider)
        for i, u in enumerate(users):
            one = m.recommend(int(u), hm, consider_known_items=consider)
            if one is None:
                assert cnt[i] == -1, u
                continue
            assert cnt[i] == len(one[0])
            assert rows[i, :cnt[i]].tolist() == one[0].tolist(), u
            assert scores[i, :cnt[i]].tobytes() == one[1].tobytes(), u
    m.close()
