"""GPU parity: the CUDA library, called through its C ABI, against the CPU oracle on the same
seeded inputs and against the reference's golden vectors.  Integer/index results (LSH partitions,
candidate lists, top-N row ids) must be identical; returned scores are produced by the exact
finishing pass and must be bit-identical to the oracle's doubles."""
import numpy as np
import pytest

import oracle
import oryx_b200 as ob
from als_harness import (ServingHarness, check_fixed_model_goldens, load_golden, random_hash_vectors,
                         random_model)

pytestmark = pytest.mark.gpu


class GpuBackend:
    """als_harness backend that routes every computation through liboryx_b200.so."""

    name = "gpu"

    def __init__(self, features, y, hash_vectors=None, max_bits=0, implicit=True):
        self.features = features
        self.n = len(y)
        self.model = ob.Model(features, implicit=implicit, hash_vectors=hash_vectors, max_bits_differing=max_bits)
        self.parts = self.model.upsert(ob.Y, np.arange(self.n, dtype=np.int32), y) if self.n else None

    def topn(self, kind, q, how_many, excluded=None):
        return self.model.topn(kind, q, how_many, excluded)

    def scores_all(self, kind, q):
        return self.model.scores_all(kind, q, self.n)

    def gram(self):
        return self.model.gram(ob.Y)[0]

    def solver(self, g):
        return ob.Solver(g)

    def updated_xu(self, solver, value, xu, yi, implicit):
        if yi is None:
            return None
        out, valid = solver.fold_in_batch(implicit, None if xu is None else xu[None, :],
                                          None if xu is not None else np.zeros(1, np.uint8),
                                          yi[None, :], [value])
        return out[0] if valid[0] else None

    def dot(self, x, y):
        # point lookup through the store is exercised in test_scores_rows; here a 1-row scratch model
        return _point(self.features, x, y, ob.DOT)

    def norm(self, x):
        import math
        return math.sqrt(_point(self.features, x, x, ob.DOT))

    def close(self):
        self.model.close()


def _point(features, q, y, kind):
    with ob.Model(features) as m:
        m.upsert(ob.Y, [0], np.asarray(y, dtype=np.float32)[None, :])
        return float(m.scores_rows(kind, q, [0])[0][0])


def _assert_same_topn(got, exp, scores_all=None):
    grows, gscores = got
    erows, escores = exp
    assert len(grows) == len(erows)
    # scores are exact doubles on both sides
    assert np.array_equal(gscores, escores), (gscores, escores)
    if not np.array_equal(grows, erows):
        # only exact score ties may be ordered differently (the reference leaves ties unspecified)
        for g, e, s in zip(grows, erows, escores):
            if g != e:
                assert scores_all is not None and scores_all[g] == s, (g, e, s)


def test_fixed_model_endpoint_goldens():
    g = load_golden("als_fixed_model")
    h = ServingHarness(g["model"], lambda f, y: GpuBackend(f, y))
    check_fixed_model_goldens(h)
    h.backend.close()


@pytest.mark.parametrize("n,f,how_many", [(1, 7, 3), (9, 2, 10), (1000, 50, 10), (4097, 128, 10),
                                           (20000, 20, 5), (3000, 256, 100), (777, 300, 7), (50000, 64, 1016)])
def test_topn_matches_oracle(n, f, how_many):
    y = random_model(n, f, seed=n * 31 + f)
    rng = np.random.default_rng(n + f)
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        for kind in (ob.DOT, ob.COSINE):
            for _ in range(3):
                q = rng.standard_normal(f, dtype=np.float32)
                exp = oracle.topn(kind, q, y, how_many)
                got = m.topn(kind, q, how_many)
                _assert_same_topn(got, exp, oracle.scores_all(kind, q, y))


@pytest.mark.parametrize("n,f,how_many,lsh", [(60000, 32, 2000, False), (50000, 50, 5000, True), (70000, 16, 65536, False),
                                             (3000, 8, 4000, False)])
def test_topn_deep_pages_match_oracle(n, f, how_many, lsh):
    """howMany + offset has no upper bound in the reference (Recommend.java:81-82,108-112).  Beyond the fused
    scan's 1016 the library scores every row exactly, radix-selects and sorts on the device; rows and
    binary64 scores must equal the oracle's list (LSH filtering, exclusions, removals, duplicate rows,
    fewer competing rows than requested)."""
    from oryx_b200 import native
    y = random_model(n, f, seed=n + how_many)
    y[100:140] = y[99]                      # a block of identical rows: equal scores, ranked by row
    rng = np.random.default_rng(how_many)
    hv, bits, h = None, 0, 0
    if lsh:
        h, bits = ob.lsh_plan(0.5, 8)
        hv = random_hash_vectors(h, f, seed=3)
    with ob.Model(f, hash_vectors=hv, max_bits_differing=bits) as m:
        parts = m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        gone = rng.choice(n, 50, replace=False).astype(np.int32)
        m.remove(ob.Y, gone)
        part = (parts if lsh else np.zeros(n, np.int32)).copy()
        part[gone] = -1
        for kind in (ob.DOT, ob.COSINE):
            q = rng.standard_normal(f, dtype=np.float32)
            excl = rng.integers(0, n, 300).astype(np.int32)
            cand = oracle.lsh_candidates(h, bits, oracle.lsh_index(hv, q)) if lsh else None
            e_rows, e_scores = oracle.topn(kind, q, y, how_many, part=part, cand=cand, excluded=np.unique(excl))
            rows, scores = m.topn(kind, q, how_many, excluded=excl)
            assert len(rows) == len(e_rows)
            assert np.array_equal(scores, e_scores)
            # equal scores are ordered by ascending row here; the oracle's heap leaves ties unspecified
            _assert_same_topn((rows, scores), (e_rows, e_scores), oracle.scores_all(kind, q, y))
            assert len(set(rows.tolist())) == len(rows)
            same = np.nonzero(np.diff(scores) == 0)[0]
            assert np.all(rows[same] < rows[same + 1])
            # the fused scan and the deep path agree on a page both can serve
            r1, s1 = m.topn(kind, q, native.MAX_HOW_MANY, excluded=excl)
            assert np.array_equal(s1, scores[:len(s1)])
    with pytest.raises(ob.OryxError):
        with ob.Model(f) as m2:
            m2.upsert(ob.Y, np.arange(10, dtype=np.int32), y[:10])
            m2.topn(ob.DOT, y[0], native.MAX_HOW_MANY_DEEP + 1)


def test_topn_with_exclusions_and_removals():
    n, f = 5000, 32
    y = random_model(n, f, seed=11)
    rng = np.random.default_rng(12)
    q = rng.standard_normal(f, dtype=np.float32)
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        full_rows, _ = oracle.topn(oracle.DOT, q, y, 20)
        excl = np.concatenate([full_rows[:7], rng.integers(0, n, 40)]).astype(np.int32)
        exp = oracle.topn(oracle.DOT, q, y, 10, excluded=np.unique(excl))
        got = m.topn(ob.DOT, q, 10, excluded=excl)  # unsorted, with duplicates: the library sorts a copy
        _assert_same_topn(got, exp)
        # remove the current best rows: they must disappear (FeatureVectors.removeVector)
        gone = full_rows[:5].astype(np.int32)
        m.remove(ob.Y, gone)
        keep = np.ones(n, dtype=bool)
        keep[gone] = False
        part = np.where(keep, 0, -1).astype(np.int32)
        exp = oracle.topn(oracle.DOT, q, y, 10, part=part)
        _assert_same_topn(m.topn(ob.DOT, q, 10), exp)
        assert m.get(ob.Y, int(gone[0])) is None
        # re-insert one with a new vector (setItemVector overwrite), update another in place
        y2 = y.copy()
        y2[gone[0]] = 10.0 * q  # becomes the clear winner
        y2[123] = -y[123]
        m.upsert(ob.Y, [int(gone[0]), 123], y2[[gone[0], 123]])
        keep[gone[0]] = True
        part = np.where(keep, 0, -1).astype(np.int32)
        exp = oracle.topn(oracle.DOT, q, y2, 10, part=part)
        got = m.topn(ob.DOT, q, 10)
        _assert_same_topn(got, exp)
        assert got[0][0] == gone[0]
        assert np.array_equal(m.get(ob.Y, 123), y2[123])
        st = m.stats()
        assert st["rows"][ob.Y] == n - 4


def test_empty_and_tiny_models():
    with ob.Model(5) as m:
        rows, scores = m.topn(ob.DOT, np.ones(5, np.float32), 10)
        assert len(rows) == 0  # ALSServingModel.java:307-310 Stream.empty()
        m.upsert(ob.Y, [3], np.ones((1, 5), np.float32))
        rows, scores = m.topn(ob.DOT, np.ones(5, np.float32), 10)
        assert list(rows) == [3] and list(scores) == [5.0]
        rows, scores = m.topn(ob.DOT, np.ones(5, np.float32), 10, excluded=[3])
        assert len(rows) == 0
        with pytest.raises(ob.OryxError):
            m.topn(ob.DOT, np.ones(5, np.float32), 0)
        rows, scores = m.topn(ob.DOT, np.ones(5, np.float32), 5000)   # deep page over a one-row model
        assert list(rows) == [3] and list(scores) == [5.0]
        with pytest.raises(ob.OryxError):
            m.topn(ob.DOT, np.ones(5, np.float32), 70000)              # beyond ORYX_MAX_HOW_MANY_DEEP


def test_nan_and_ties():
    f = 4
    y = np.zeros((6, f), np.float32)
    y[0] = [1, 0, 0, 0]
    y[1] = [1, 0, 0, 0]      # exact tie with row 0
    y[2] = [np.nan, 0, 0, 0]  # NaN score never enters (TopNConsumer.java:60)
    y[3] = [0.5, 0, 0, 0]
    y[4] = [0, 0, 0, 0]       # cosine: 0/0 = NaN
    y[5] = [-1, 0, 0, 0]
    q = np.array([1, 0, 0, 0], np.float32)
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(6, dtype=np.int32), y)
        rows, scores = m.topn(ob.DOT, q, 10)
        assert list(scores) == [1.0, 1.0, 0.5, 0.0, -1.0]
        assert sorted(rows[:2]) == [0, 1] and list(rows[2:]) == [3, 4, 5]
        rows, scores = m.topn(ob.COSINE, q, 10)
        assert list(scores) == [1.0, 1.0, 1.0, -1.0] and sorted(rows[:3]) == [0, 1, 3] and rows[3] == 5


@pytest.mark.parametrize("f", [7, 64, 250, 256])
def test_exact_rescore_over_wide_dynamic_ranges(f):
    """The finishing kernel computes VectorMath.dot's value with the whole warp when the binary64 chain provably
    never rounds, and walks the chain otherwise (common.cuh warp_exact_dot).  Rows whose terms span 2^-70 .. 2^70
    (chains that DO round, int64 headroom exceeded), subnormal products, exact cancellations, zero rows and signed
    zeros must all come back with the oracle's bits."""
    rng = np.random.default_rng(0xB2000900 + f)
    n = 4000
    y = rng.standard_normal((n, f)).astype(np.float32)
    scale = np.exp2(rng.integers(-35, 36, (n, f))).astype(np.float32)
    y[:1500] *= scale[:1500]                                  # wide spread: most of these chains round
    y[1500:1600] *= np.float32(2.0 ** -80)                    # products deep in the subnormal range
    y[1600:1700, : f // 2] = -y[1600:1700, f - f // 2:][:, ::-1] if f % 2 == 0 else y[1600:1700, : f // 2]
    y[1700:1710] = 0.0
    y[1710:1720] = -0.0
    y[1720:1800] = np.round(y[1720:1800] * 8) / 8              # short significands: nothing ever rounds
    qs = [rng.standard_normal(f).astype(np.float32),
          (rng.standard_normal(f) * np.exp2(rng.integers(-35, 36, f))).astype(np.float32),
          np.full(f, 2.0 ** -70, np.float32),
          np.where(np.arange(f) % 2 == 0, 1.0, -1.0).astype(np.float32)]
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        for q in qs:
            for kind in (ob.DOT, ob.COSINE):
                for hm in (10, 200):
                    with np.errstate(all="ignore"):
                        exp = oracle.topn(kind, q, y, hm)
                        allsc = oracle.scores_all(kind, q, y) if hasattr(oracle, "scores_all") else None
                    _assert_same_topn(m.topn(kind, q, hm), exp, allsc)


@pytest.mark.parametrize("rate,cores,f", [(0.5, 32, 20), (0.3, 32, 50), (0.1, 8, 128), (1.0, 8, 10), (0.5, 1, 16)])
def test_lsh_bit_exact_and_filtered_topn(rate, cores, f):
    h, bits = ob.lsh_plan(rate, cores)
    assert (h, bits) == oracle.lsh_plan(rate, cores)
    hv = random_hash_vectors(h, f, seed=h * 7 + f)
    n = 20000
    y = random_model(n, f, seed=3 * f)
    rng = np.random.default_rng(f)
    with ob.Model(f, hash_vectors=hv, max_bits_differing=bits) as m:
        parts = m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        exp_parts = oracle.lsh_index_rows(hv, y)
        assert np.array_equal(parts, exp_parts)  # bit-exact partition index per item
        assert np.array_equal(m.lsh_index_rows(y[:1000]), exp_parts[:1000])
        for _ in range(4):
            q = rng.standard_normal(f, dtype=np.float32)
            main = oracle.lsh_index(hv, q)
            assert m.lsh_index(q) == main
            cand = oracle.lsh_candidates(h, bits, main)
            assert np.array_equal(m.lsh_candidates(q), cand)  # same list, same order
            for kind in (ob.DOT, ob.COSINE):
                exp = oracle.topn(kind, q, y, 10, part=exp_parts, cand=cand)
                _assert_same_topn(m.topn(kind, q, 10), exp)
        # compaction changes layout only
        q = rng.standard_normal(f, dtype=np.float32)
        before = m.topn(ob.DOT, q, 10)
        m.compact(ob.Y)
        after = m.topn(ob.DOT, q, 10)
        assert np.array_equal(before[0], after[0]) and np.array_equal(before[1], after[1])
        assert np.array_equal(m.get(ob.Y, 77), y[77])
        # an update that changes a row's partition is honoured (ALSServingModel.java:165-176)
        newv = -y[5]
        p = m.upsert(ob.Y, [5], newv[None, :])
        assert p[0] == oracle.lsh_index(hv, newv)


@pytest.mark.parametrize("compact", [False, True])
def test_lsh_pipelined_device_queries_match_oracle(compact):
    """oryx_topn_device calls queued back to back on one stream, LSH on: every query is hashed and its candidate
    blocks are listed by one kernel, and the scan that follows is released without waiting for the previous
    query's finishing kernel (scan_topn.cu).  Each answer must still be the oracle's for ITS query."""
    import torch
    f, n, hm, nq = 64, 300_000, 10, 96
    h, bits = ob.lsh_plan(0.3, 32)
    hv = random_hash_vectors(h, f, seed=41)
    y = random_model(n, f, seed=42)
    qs = np.random.default_rng(43).standard_normal((nq, f)).astype(np.float32)
    # a few queries that sit on a hash hyperplane (the exact fallback decides their bit)
    qs[5] = hv[1]
    qs[6] = np.float32(1e-30)
    parts = oracle.lsh_index_rows(hv, y)
    dev = torch.device("cuda", 0)
    with ob.Model(f, hash_vectors=hv, max_bits_differing=bits, capacity_hint=n) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        if compact:
            m.compact(ob.Y)
        d_q = torch.from_numpy(qs).to(dev)
        d_rows = torch.full((nq, hm), -1, dtype=torch.int32, device=dev)
        d_scores = torch.zeros((nq, hm), dtype=torch.float64, device=dev)
        d_n = torch.zeros(nq, dtype=torch.int32, device=dev)
        st = torch.cuda.current_stream(dev)
        torch.cuda.synchronize()
        for rep in range(3):
            for i in range(nq):
                m.topn_device(ob.DOT, d_q[i].data_ptr(), hm, d_rows[i].data_ptr(), d_scores[i].data_ptr(),
                              d_n[i:].data_ptr(), st.cuda_stream)
        torch.cuda.synchronize()
        rows, scores, cnt = d_rows.cpu().numpy(), d_scores.cpu().numpy(), d_n.cpu().numpy()
        for i in range(nq):
            main = oracle.lsh_index(hv, qs[i])
            cand = oracle.lsh_candidates(h, bits, main)
            exp = oracle.topn(ob.DOT, qs[i], y, hm, part=parts, cand=cand)
            _assert_same_topn((rows[i, :cnt[i]], scores[i, :cnt[i]]), exp)


@pytest.mark.parametrize("f,h", [(128, 5), (50, 8), (33, 16), (256, 3)])
def test_lsh_index_near_orthogonal_and_degenerate_rows(f, h):
    """The hashing kernel calls a sign from fp32 arithmetic only when the value clears a rigorous error
    bound and otherwise repeats the reference's exact float-multiply / double-add chain.  Rows built to sit
    on (or within rounding of) a hash hyperplane, zero / denormal / huge / NaN / infinite rows and exact
    cancellations must all get the oracle's index."""
    rng = np.random.default_rng(f * 31 + h)
    hv = random_hash_vectors(h, f, seed=f + h)
    rows = []
    w = rng.standard_normal((4000, f))
    for i in range(len(w)):                       # remove the component along one hash vector (in binary64),
        k = i % h                                 # then round to binary32: the dot product is ~1e-8 of |w|
        hk = hv[k].astype(np.float64)
        rows.append(w[i] - (w[i] @ hk) / (hk @ hk) * hk)
    rows = np.asarray(rows)
    rows[::7] *= 1e-3
    rows[1::7] *= 1e4
    extra = np.zeros((12, f))
    extra[1, :] = 1e-42                            # denormals
    extra[2, :] = 3e37
    extra[3, :] = -3e37
    extra[4, 0] = np.nan
    extra[5, 0] = np.inf
    extra[6, :2] = [np.inf, -np.inf]
    extra[7] = hv[0]                               # on the normal itself
    extra[8] = -hv[0]
    extra[9, :2] = [1.0, -1.0]                     # exact cancellation against a constant hash vector (below)
    extra[10, :] = 1e-30
    extra[11, :] = -1e-30
    y = np.concatenate([rows, extra]).astype(np.float32)
    hv2 = hv.copy()
    hv2[h - 1, :] = 0.0
    hv2[h - 1, :2] = 0.5                           # dot with row 9 is exactly 0 -> bit must be 0
    for vecs in (hv, hv2):
        with ob.Model(f, hash_vectors=vecs, max_bits_differing=0) as m:
            with np.errstate(all="ignore"):
                exp = oracle.lsh_index_rows(vecs, y)
            assert np.array_equal(m.lsh_index_rows(y), exp)
            assert np.array_equal(m.upsert(ob.Y, np.arange(len(y), dtype=np.int32), y), exp)


def test_lsh_edge_vectors():
    # zero vector -> index 0; candidates lists of LocalitySensitiveHashTest.java:85-139
    L = load_golden("lsh")
    for key in ("candidates_no_sample", "candidates_one_bit", "candidates_bands"):
        c = L[key]
        h, bits = ob.lsh_plan(c["sample_rate"], c["cores"])
        hv = random_hash_vectors(h, c["features"], seed=5)
        with ob.Model(c["features"], hash_vectors=hv, max_bits_differing=bits) as m:
            zero = np.zeros(c["features"], np.float32)
            one = np.ones(c["features"], np.float32)
            assert m.lsh_index(zero) == 0
            for v in (zero, one):
                exp = oracle.lsh_candidates(h, bits, oracle.lsh_index(hv, v))
                assert np.array_equal(m.lsh_candidates(v), exp)
            if key == "candidates_bands":
                cand = m.lsh_candidates(one)
                assert len(cand) == c["count"]
                for lo, hi, b in c["bands"]:
                    assert all(bin(int(cand[0]) ^ int(cand[i])).count("1") == b for i in range(lo, hi))


def test_scores_all_and_rows_exact():
    n, f = 3000, 77
    y = random_model(n, f, seed=21)
    q = np.random.default_rng(22).standard_normal(f, dtype=np.float32)
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        m.remove(ob.Y, [10])
        for kind in (ob.DOT, ob.COSINE):
            got = m.scores_all(kind, q, n + 5)
            exp = oracle.scores_all(kind, q, y)
            assert np.isnan(got[10]) and np.all(np.isnan(got[n:]))
            mask = np.ones(n, bool)
            mask[10] = False
            assert np.array_equal(got[:n][mask], exp[mask])
        rows = [0, 5, 10, 2999, 5000]
        got, found = m.scores_rows(ob.DOT, q, rows)
        assert list(found) == [True, True, False, True, False]
        assert got[0] == oracle.dot(q, y[0]) and got[3] == oracle.dot(q, y[2999]) and got[2] == 0.0
        got, _ = m.scores_rows(ob.COSINE, q, [7])
        assert got[0] == oracle.dot(q, y[7]) / (oracle.norm(q) * oracle.norm(y[7]))


@pytest.mark.parametrize("n,f", [(3, 3), (2, 3), (1000, 50), (5000, 128), (700, 256), (100, 70)])
def test_gram_matches_oracle(n, f):
    y = random_model(n, f, seed=n + f)
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        g, used = m.gram(ob.Y)
        exp = oracle.transpose_times_self(y)
        assert used == n
        # same products, same binary64 adds, different row order only
        assert np.max(np.abs(g - exp)) <= 1e-12 * max(1.0, np.max(np.abs(exp)))
        k = min(5, n)
        m.upsert(ob.X, np.arange(k, dtype=np.int32), y[:k])
        gx, used = m.gram(ob.X)
        assert used == k and np.max(np.abs(gx - oracle.transpose_times_self(y[:k]))) <= 1e-12 * np.max(np.abs(exp))
    with ob.Model(f) as m:
        g, used = m.gram(ob.Y)
        assert g is None and used == 0  # VectorMath.java:65-67 null for empty


def test_gram_goldens():
    for c in load_golden("math")["transpose_times_self"]:
        rows = np.asarray(c["rows"], dtype=np.float32)
        with ob.Model(rows.shape[1]) as m:
            m.upsert(ob.Y, np.arange(len(rows), dtype=np.int32), rows)
            g, _ = m.gram(ob.Y)
            assert np.max(np.abs(g - np.asarray(c["value"]))) <= max(c["tol"], 1e-15), c["_cite"]


def test_solver_goldens_and_random():
    M = load_golden("math")["solver"]
    s = M["solve_f2f"]
    x = ob.Solver(s["a"]).solve_f2f(s["b"])
    assert np.array_equal(x, np.asarray(s["x"], dtype=np.float32)), x
    for c in M["non_singular"]:
        try:
            ob.Solver(c["a"])
            ok = True
        except ob.SingularMatrixSolverException:
            ok = False
        assert ok == c["value"], c["_cite"]
    for c in M["apparent_rank"]:
        with pytest.raises(ob.SingularMatrixSolverException) as ei:
            ob.Solver(c["a"])
        assert ei.value.apparent_rank == c["rank"], c["_cite"]
    rng = np.random.default_rng(4)
    for f in (2, 17, 64, 128, 256):
        y = rng.standard_normal((4 * f + 10, f), dtype=np.float32)
        g = oracle.transpose_times_self(y)
        so, sg = oracle.Solver(g), ob.Solver(g)
        for _ in range(3):
            b = rng.standard_normal(f, dtype=np.float32)
            xo, xg = so.solve_f2f(b), sg.solve_f2f(b)
            # binary64 inside, narrowed to float: at most a last-place difference
            assert np.max(np.abs(xo - xg)) <= 2e-7 * np.max(np.abs(xo))


def test_fold_in_goldens_and_random():
    u = load_golden("math")["updated_xu"]
    g = oracle.transpose_times_self(np.asarray(u["rows"], dtype=np.float32))
    sg = ob.Solver(g)
    for c in u["cases"]:
        if c["yi"] is None:
            continue
        xu = None if c["xu"] is None else np.asarray(c["xu"], np.float32)[None, :]
        has = None if xu is not None else np.zeros(1, np.uint8)
        out, valid = sg.fold_in_batch(c["implicit"], xu, has, np.asarray(c["yi"], np.float32)[None, :], [c["value"]])
        assert valid[0] == 1 and np.array_equal(out[0], np.asarray(c["out"], np.float32)), (out, c)
    # speed-layer goldens (ALSSpeedIT.java:80-113)
    s = load_golden("speed")
    X = {k: np.asarray(v, np.float32) for k, v in s["X"].items()}
    Y = {k: np.asarray(v, np.float32) for k, v in s["Y"].items()}
    yty = ob.Solver(oracle.transpose_times_self(np.stack(list(Y.values()))))
    xtx = ob.Solver(oracle.transpose_times_self(np.stack(list(X.values()))))
    for user, item, value in s["interactions"]:
        if user in s["expected_X"]:
            out, valid = yty.fold_in_batch(True, None, np.zeros(1, np.uint8), Y[item][None, :], [value])
            assert valid[0] and np.max(np.abs(out[0] - np.asarray(s["expected_X"][user], np.float32))) <= s["tol"]
        else:
            out, valid = xtx.fold_in_batch(True, None, np.zeros(1, np.uint8), X[user][None, :], [value])
            assert valid[0] and np.max(np.abs(out[0] - np.asarray(s["expected_Y"][item], np.float32))) <= s["tol"]
    # random batch, implicit and explicit, mixed null Xu, values that yield "no change"
    rng = np.random.default_rng(9)
    for f in (8, 50, 128):
        n = 1000
        ymat = rng.standard_normal((3000, f), dtype=np.float32) * 0.3
        gram = oracle.transpose_times_self(ymat)
        so, sgpu = oracle.Solver(gram), ob.Solver(gram)
        xu = rng.standard_normal((n, f), dtype=np.float32) * 0.3
        yi = ymat[rng.integers(0, 3000, n)]
        has = (rng.random(n) < 0.7).astype(np.uint8)
        vals = rng.choice([1.0, 2.5, -1.0, 0.0, 0.5, -0.3], n)
        for implicit in (True, False):
            eo, ev = oracle.fold_in_batch(so, implicit, xu, has, yi, vals)
            go, gv = sgpu.fold_in_batch(implicit, xu, has, yi, vals)
            assert np.array_equal(ev, gv)
            ok = ev.astype(bool)
            scale = np.maximum(np.abs(eo[ok]).max(axis=1, keepdims=True), 1e-6)
            assert np.max(np.abs(eo[ok] - go[ok]) / scale) <= 5e-7


def test_cross_shard_merge_kernels():
    """oryx_merge_topn_device / oryx_merge_topn_packed: per-shard exact lists -> the global list."""
    import torch
    n, f, how_many, shards = 9000, 24, 10, 3
    y = random_model(n, f, seed=61)
    q = np.random.default_rng(62).standard_normal(f, dtype=np.float32)
    per = n // shards
    dev = torch.device("cuda")
    rec = ob.packed_record_bytes(how_many)
    g_rec = torch.zeros(shards * rec, dtype=torch.uint8, device=dev)
    g_scores = torch.zeros((shards, how_many), dtype=torch.float64, device=dev)
    g_rows = torch.zeros((shards, how_many), dtype=torch.int32, device=dev)
    g_n = torch.zeros(shards, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)
    with ob.Model(f) as m:
        for sidx in range(shards):
            # shard 2 holds only 4 rows: a short list must merge correctly
            lo, hi = sidx * per, (sidx * per + 4 if sidx == 2 else (sidx + 1) * per)
            if sidx > 0:
                m.remove(ob.Y, np.arange(0, per, dtype=np.int32))
            m.upsert(ob.Y, np.arange(hi - lo, dtype=np.int32), y[lo:hi])
            r = g_rec[sidx * rec:(sidx + 1) * rec]
            d_q = torch.from_numpy(q).to(dev)
            m.topn_device(ob.DOT, d_q.data_ptr(), how_many, r[how_many * 8:].data_ptr(), r.data_ptr(),
                          r[how_many * 12:].data_ptr(), stream.cuda_stream)
            stream.synchronize()
            g_scores[sidx] = r[: how_many * 8].view(torch.float64)
            g_rows[sidx] = r[how_many * 8: how_many * 12].view(torch.int32)
            g_n[sidx] = r[how_many * 12: how_many * 12 + 4].view(torch.int32)[0]
    keep = np.ones(n, bool)
    keep[2 * per + 4:] = False
    part = np.where(keep, 0, -1).astype(np.int32)
    e_rows, e_scores = oracle.topn(oracle.DOT, q, y, how_many, part=part)
    for packed in (False, True):
        o_rows = torch.zeros(how_many, dtype=torch.int64, device=dev)
        o_scores = torch.zeros(how_many, dtype=torch.float64, device=dev)
        o_n = torch.zeros(1, dtype=torch.int32, device=dev)
        if packed:
            ob.merge_topn_packed(g_rec.data_ptr(), shards, rec, how_many, per, o_rows.data_ptr(),
                                 o_scores.data_ptr(), o_n.data_ptr(), stream.cuda_stream)
        else:
            ob.merge_topn_device(g_scores.data_ptr(), g_rows.data_ptr(), g_n.data_ptr(), shards, how_many, per,
                                 o_rows.data_ptr(), o_scores.data_ptr(), o_n.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        assert int(o_n.item()) == how_many
        assert np.array_equal(o_rows.cpu().numpy(), e_rows) and np.array_equal(o_scores.cpu().numpy(), e_scores)


def _batch_counters(m):
    """(batches answered by the tcgen05 path, full bf16 conversions of Y) of a model so far."""
    import ctypes as C
    from oryx_b200 import native
    L = native.lib()
    L.oryx_debug_batch_timings.restype = C.c_int32
    L.oryx_debug_batch_timings.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    t = (C.c_double * 9)()
    L.oryx_debug_batch_timings(m._h, t)
    return int(t[7]), int(t[8])


@pytest.mark.parametrize("n,f,nq,how_many", [(20000, 128, 64, 10), (9000, 50, 130, 100), (33000, 256, 40, 10),
                                             (5000, 64, 32, 1)])
def test_topn_batch_tensor_core_path_is_exact(n, f, nq, how_many):
    """oryx_topn_batch (tcgen05 scoring + exact pruning/refinement) == one exact topN per query."""
    y = random_model(n, f, seed=n + 7 * f)
    rng = np.random.default_rng(nq + f)
    queries = rng.standard_normal((nq, f), dtype=np.float32)
    queries[3] *= 0.01          # a low-norm query
    queries[5] = -queries[5]
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        m.remove(ob.Y, [17, 4000])  # tombstones must be ignored by the tensor-core pass as well
        keep = np.ones(n, bool)
        keep[[17, 4000]] = False
        part = np.where(keep, 0, -1).astype(np.int32)
        for kind in (ob.DOT, ob.COSINE):
            excl = [rng.integers(0, n, 5).astype(np.int32) if i % 3 == 0 else np.zeros(0, np.int32) for i in range(nq)]
            runs_before = _batch_counters(m)[0]
            rows, scores, cnt = m.topn_batch(kind, queries, how_many, excluded_lists=excl)
            # the page must have been answered by the tensor-core kernels, not by the one-by-one fallback
            assert _batch_counters(m)[0] == runs_before + 1
            for i in range(nq):
                e_rows, e_scores = oracle.topn(kind, queries[i], y, how_many, part=part,
                                               excluded=np.unique(excl[i]) if len(excl[i]) else None)
                assert cnt[i] == len(e_rows), (i, cnt[i], len(e_rows))
                _assert_same_topn((rows[i, :cnt[i]], scores[i, :cnt[i]]), (e_rows, e_scores),
                                  oracle.scores_all(kind, queries[i], y))


def test_topn_batch_in_passes_when_scratch_is_bounded():
    """The sub-block maxima need nq x N / 32 floats of scratch; a page that would exceed the budget is answered
    in passes of fewer queries.  Same answers, exclusions sliced per pass."""
    import os
    n, f, nq, how_many = 40000, 64, 300, 10
    y = random_model(n, f, seed=0xB2000400)
    rng = np.random.default_rng(5)
    q = rng.standard_normal((nq, f), dtype=np.float32)
    excl = [rng.integers(0, n, int(rng.integers(0, 9))).astype(np.int32) for _ in range(nq)]
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        whole = m.topn_batch(ob.DOT, q, how_many, excluded_lists=excl)
        os.environ["ORYX_BATCH_SCRATCH_MB"] = "1"      # 1 MB: passes of 128 queries
        try:
            runs0 = _batch_counters(m)[0]
            parts = m.topn_batch(ob.DOT, q, how_many, excluded_lists=excl)
            assert _batch_counters(m)[0] == runs0 + 1
        finally:
            del os.environ["ORYX_BATCH_SCRATCH_MB"]
        for a, b in zip(whole, parts):
            assert np.array_equal(a, b)
        for i in (0, 127, 128, 299):
            e = oracle.topn(oracle.DOT, q[i], y, how_many, excluded=np.unique(excl[i]) if len(excl[i]) else None)
            _assert_same_topn((parts[0][i, :parts[2][i]], parts[1][i, :parts[2][i]]), e)


@pytest.mark.parametrize("rate,cores", [(0.3, 32), (0.5, 8), (0.1, 8)])
def test_topn_batch_with_lsh_filtering(rate, cores):
    """With LSH filtering on, a page of queries still rides on the tensor-core path: every query only
    competes for rows of its own candidate partitions (LocalitySensitiveHash.getCandidateIndices).  Batch ==
    one query at a time == oracle, before compaction (every 256-slot tile mixes partitions) and after it
    (tiles are pure and non-candidate tiles are skipped), with removals and exclusions, through
    oryx_topn_batch, oryx_recommend_batch and the combining entry point."""
    n, f, nq, how_many = 60000, 64, 70, 10
    h, bits = ob.lsh_plan(rate, cores)
    hv = random_hash_vectors(h, f, seed=h + 100)
    y = random_model(n, f, seed=0xB2000300 + h)
    rng = np.random.default_rng(h)
    q = rng.standard_normal((nq, f), dtype=np.float32)
    gone = rng.choice(n, 300, replace=False).astype(np.int32)
    with ob.Model(f, hash_vectors=hv, max_bits_differing=bits) as m:
        parts = m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        m.upsert(ob.X, np.arange(nq, dtype=np.int32), q)
        m.remove(ob.Y, gone)
        part = parts.copy()
        part[gone] = -1
        excl = [rng.integers(0, n, 7).astype(np.int32) if i % 2 else np.zeros(0, np.int32) for i in range(nq)]
        m.known_items_add(np.arange(nq, dtype=np.int32), excl)
        for phase in ("as loaded", "compacted"):
            for kind in (ob.DOT, ob.COSINE):
                runs0 = _batch_counters(m)[0]
                rows, scores, cnt = m.topn_batch(kind, q, how_many, excluded_lists=excl)
                assert _batch_counters(m)[0] == runs0 + 1, "LSH filtering switched the tensor-core path off"
                for i in range(nq):
                    cand = oracle.lsh_candidates(h, bits, oracle.lsh_index(hv, q[i]))
                    e = oracle.topn(kind, q[i], y, how_many, part=part, cand=cand,
                                    excluded=np.unique(excl[i]) if len(excl[i]) else None)
                    assert cnt[i] == len(e[0]), (phase, kind, i)
                    _assert_same_topn((rows[i, :cnt[i]], scores[i, :cnt[i]]), e, oracle.scores_all(kind, q[i], y))
                    if i < 6:
                        _assert_same_topn(m.topn(kind, q[i], how_many, excl[i]), e, oracle.scores_all(kind, q[i], y))
            r_rows, r_scores, r_cnt = m.recommend_batch(np.arange(nq, dtype=np.int32), how_many)
            for i in (0, 1, nq - 1):
                cand = oracle.lsh_candidates(h, bits, oracle.lsh_index(hv, q[i]))
                e = oracle.topn(oracle.DOT, q[i], y, how_many, part=part, cand=cand,
                                excluded=np.unique(excl[i]) if len(excl[i]) else None)
                _assert_same_topn((r_rows[i, :r_cnt[i]], r_scores[i, :r_cnt[i]]), e, oracle.scores_all(oracle.DOT, q[i], y))
                _assert_same_topn(m.topn_coalesced(ob.DOT, q[i], how_many, excluded=excl[i]), e,
                                  oracle.scores_all(oracle.DOT, q[i], y))
            m.compact(ob.Y)


@pytest.mark.parametrize("n_known", [20, 60, 100])
def test_batch_paths_with_top_scoring_rows_excluded(n_known):
    """A user's known items are typically the rows that score highest for that user.  The batched
    pruning cut must count them: every query here excludes its own global top `n_known` rows, through
    oryx_topn_batch, oryx_recommend_batch and both combining entry points, against the oracle."""
    n, f, nq, how_many = 70000, 64, 48, 10
    y = random_model(n, f, seed=0xB2000200)
    x = random_model(nq, f, seed=0xB2000201)
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        m.upsert(ob.X, np.arange(nq, dtype=np.int32), x)
        known = [oracle.topn(oracle.DOT, x[i], y, n_known)[0].astype(np.int32) for i in range(nq)]
        m.known_items_add(np.arange(nq, dtype=np.int32), known)
        expect = [oracle.topn(oracle.DOT, x[i], y, how_many, excluded=np.sort(known[i])) for i in range(nq)]
        runs0 = _batch_counters(m)[0]
        rows, scores, cnt = m.topn_batch(ob.DOT, x, how_many, excluded_lists=known)
        r_rows, r_scores, r_cnt = m.recommend_batch(np.arange(nq, dtype=np.int32), how_many)
        assert _batch_counters(m)[0] == runs0 + 2
        for i in range(nq):
            assert cnt[i] == how_many and r_cnt[i] == how_many
            _assert_same_topn((rows[i, :cnt[i]], scores[i, :cnt[i]]), expect[i])
            _assert_same_topn((r_rows[i, :r_cnt[i]], r_scores[i, :r_cnt[i]]), expect[i])
        for i in (0, 7, nq - 1):
            _assert_same_topn(m.topn_coalesced(ob.DOT, x[i], how_many, excluded=known[i]), expect[i])
            got = m.recommend_coalesced(i, how_many)
            _assert_same_topn((got[0], got[1]), expect[i])


@pytest.mark.parametrize("n,f", [(70000, 100), (66000, 256)])
def test_gram_large_store_fast_path(n, f):
    """Stores of >= 65536 slots use fp32-FMA tiles with binary64 flushes; agreement with the exact
    float-multiply / double-accumulate Gram must stay far below what Solver.solveFToF's float output
    can resolve (tolerance 1e-8 relative to the largest entry)."""
    y = random_model(n, f, seed=n + f) * 0.3
    with ob.Model(f) as m:
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        m.remove(ob.Y, [5, 60000])
        g, used = m.gram(ob.Y)
        keep = np.ones(n, bool)
        keep[[5, 60000]] = False
        exp = oracle.transpose_times_self(y[keep])
        assert used == n - 2
        assert np.max(np.abs(g - exp)) <= 1e-8 * np.max(np.abs(exp))
        assert np.array_equal(g, g.T)


@pytest.mark.parametrize("n,f", [(1_100_000, 256), (1_048_576 + 777, 200), (1_200_000, 128)])
def test_gram_tensor_core_path(n, f):
    """Stores of >= 2^20 rows with 128..256 padded features take the exact-sum tcgen05 kernel (int8 digit
    planes, int32 TMEM accumulators).  Against the oracle's float-multiply / double-accumulate Gram the only
    difference is the 24-bit fixed-point quantisation of the inputs: tolerance 1e-8 relative to the largest
    entry (the bound the floating-point large-store kernel is held to); the result is exactly symmetric and
    tombstones contribute nothing."""
    import os
    y = random_model(n, f, seed=n + f) * 0.3
    gone = np.array([5, 60000, n - 1, n // 2], dtype=np.int32)
    with ob.Model(f, capacity_hint=n) as m:
        for off in range(0, n, 250_000):
            m.upsert(ob.Y, np.arange(off, min(n, off + 250_000), dtype=np.int32), y[off:off + 250_000])
        m.remove(ob.Y, gone)
        g, used = m.gram(ob.Y)
        ms, on_tensor_cores = m.debug_gram()
        assert on_tensor_cores, "the large store did not take the tensor-core Gram"
        assert used == n - len(gone)
        keep = np.ones(n, bool)
        keep[gone] = False
        exp = oracle.transpose_times_self_mt(y[keep], os.cpu_count() or 4)
        err = np.max(np.abs(g - exp)) / np.max(np.abs(exp))
        assert err <= 1e-8, err
        assert np.array_equal(g, g.T)
        # a second call (partials and counters are reused) gives the same matrix up to the order in which
        # the CTAs' binary64 partials were formed
        g2, _ = m.gram(ob.Y)
        assert np.max(np.abs(g2 - g)) <= 1e-13 * np.max(np.abs(g))
        # The int32 accumulators are drained every 32768 rows per CTA, which a store of this size never
        # reaches.  Lower the interval so that every CTA drains many times -- including intervals that end
        # exactly where a CTA's work ends (2048-row units) -- and the protocol is covered here too.
        for drain_rows in (2048, 4096, 1024, 6144):
            os.environ["ORYX_GRAM_TC_DRAIN_ROWS"] = str(drain_rows)
            try:
                for _ in range(2):
                    g3, _ = m.gram(ob.Y)
                    assert m.debug_gram()[1]
                    assert np.max(np.abs(g3 - g)) <= 1e-13 * np.max(np.abs(g)), drain_rows
            finally:
                del os.environ["ORYX_GRAM_TC_DRAIN_ROWS"]


def test_gram_tensor_core_falls_back_on_heavy_tails():
    """One huge outlier sets the fixed-point scale so coarse that the error estimate of the tensor-core Gram
    exceeds its tolerance: oryx_gram must then answer from the floating-point kernel."""
    import os
    n, f = 1_050_000, 128
    y = random_model(n, f, seed=77) * 0.1
    y[12345, 7] = 3.0e4
    with ob.Model(f, capacity_hint=n) as m:
        for off in range(0, n, 350_000):
            m.upsert(ob.Y, np.arange(off, min(n, off + 350_000), dtype=np.int32), y[off:off + 350_000])
        g, _ = m.gram(ob.Y)
        assert not m.debug_gram()[1]
        # off the outlier's row and column the entries must still be accurate to the usual tolerance of the
        # floating-point kernel, measured against the largest of THOSE entries
        exp = oracle.transpose_times_self_mt(y, os.cpu_count() or 4)
        sel = np.ones(f, bool)
        sel[7] = False
        sub, esub = g[np.ix_(sel, sel)], exp[np.ix_(sel, sel)]
        assert np.max(np.abs(sub - esub)) <= 1e-7 * np.max(np.abs(esub))
        assert abs(g[7, 7] - exp[7, 7]) <= 1e-7 * exp[7, 7]


def test_fold_in_rows_equals_host_buffer_path():
    """oryx_fold_in_rows (Xu / Yi gathered from the device stores) gives bit-identical results to
    oryx_fold_in_batch on the same vectors, including null Xu, null Yi and removed rows."""
    rng = np.random.default_rng(0xB2000060)
    f, nu, ni, n = 40, 500, 700, 3000
    x = rng.standard_normal((nu, f), dtype=np.float32) * 0.3
    y = rng.standard_normal((ni, f), dtype=np.float32) * 0.3
    with ob.Model(f) as m:
        m.upsert(ob.X, np.arange(nu, dtype=np.int32), x)
        m.upsert(ob.Y, np.arange(ni, dtype=np.int32), y)
        gone_x, gone_y = np.array([3, 77], np.int32), np.array([5, 600], np.int32)
        m.remove(ob.X, gone_x)
        m.remove(ob.Y, gone_y)
        yty = ob.Solver(m.gram(ob.Y)[0])
        xtx = ob.Solver(m.gram(ob.X)[0])
        ur = rng.integers(-1, nu + 20, n).astype(np.int32)   # -1 and rows >= nu: unknown users
        ir = rng.integers(0, ni + 20, n).astype(np.int32)    # rows >= ni: unknown items
        ur[:4] = [3, 77, 10, 11]
        ir[:4] = [1, 2, 5, 600]
        vals = rng.choice([1.0, 2.5, -1.0, 0.0, 0.5], n)
        has_x = ((ur >= 0) & (ur < nu) & ~np.isin(ur, gone_x)).astype(np.uint8)
        has_y = ((ir < ni) & ~np.isin(ir, gone_y)).astype(np.uint8)
        xu = np.where(has_x[:, None] == 1, x[np.clip(ur, 0, nu - 1)], 0).astype(np.float32)
        yi = np.where(has_y[:, None] == 1, y[np.clip(ir, 0, ni - 1)], 0).astype(np.float32)
        for implicit in (True, False):
            # new user vectors: solver of Y, Xu from X rows, Yi from Y rows
            want, wv = yty.fold_in_batch(implicit, xu, has_x, yi, vals)
            got, gv = yty.fold_in_rows(m, implicit, ob.X, ur, ir, vals)
            wv = wv & has_y
            assert np.array_equal(gv, wv)
            assert np.array_equal(got[gv == 1], want[wv == 1])
            # new item vectors: solver of X, "Xu" from Y rows, "Yi" from X rows
            want, wv = xtx.fold_in_batch(implicit, yi, has_y, xu, vals)
            got, gv = xtx.fold_in_rows(m, implicit, ob.Y, ir, ur, vals)
            wv = wv & has_x
            assert np.array_equal(gv, wv)
            assert np.array_equal(got[gv == 1], want[wv == 1])


def test_batch_shadow_follows_updates():
    """The bf16 shadow of Y that the batched path scores against is kept current by upserts and
    removals of existing rows (no full rebuild), small pages of queries ride on it, and the answers
    stay identical to the single-query path."""
    counters = _batch_counters   # batches answered by the tensor-core path, full conversions of Y

    n, f, hm = 50000, 64, 10
    rng = np.random.default_rng(0xB20000C0)
    y = random_model(n, f, 0xB20000C1)
    q = random_model(64, f, 0xB20000C2)
    for kind in (ob.DOT, ob.COSINE):
        with ob.Model(f) as m:
            m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
            m.topn_batch(kind, q, hm)                      # builds the shadow
            runs0, rebuilds0 = counters(m)
            assert rebuilds0 == 1
            y_now = y.copy()
            for step in range(3):
                rows = rng.choice(n, 300, replace=False).astype(np.int32)
                y_now[rows] = random_model(300, f, 0xB20000C3 + step) * (3.0 if step == 1 else 1.0)  # larger norms too
                m.upsert(ob.Y, rows, y_now[rows])
                gone = rng.choice(n, 40, replace=False).astype(np.int32)
                m.remove(ob.Y, gone)
                back = gone[:20]
                m.upsert(ob.Y, back, y_now[back])          # re-uses freed slots
                dead = set(gone[20:].tolist())
                page = q[step * 5: step * 5 + 5]           # a page of 5: only taken while the shadow is current
                b_rows, b_scores, b_cnt = m.topn_batch(kind, page, hm)
                for i in range(len(page)):
                    r1, s1 = m.topn(kind, page[i], hm)
                    assert b_cnt[i] == len(r1) and b_rows[i, :b_cnt[i]].tolist() == r1.tolist()
                    assert b_scores[i, :b_cnt[i]].tobytes() == s1.tobytes()
                    assert not (set(r1.tolist()) & dead)
                y_now[list(dead)] = 0
            runs1, rebuilds1 = counters(m)
            assert runs1 == runs0 + 3 and rebuilds1 == rebuilds0   # the tensor-core path, without a rebuild
            # growing the store invalidates the shadow: a small page then goes one by one, a large one rebuilds
            # (more new rows than the 60 freed slots, so the slot array really grows)
            m.upsert(ob.Y, np.arange(n, n + 200, dtype=np.int32), random_model(200, f, 0xB20000C9))
            m.topn_batch(kind, q[:5], hm)
            assert counters(m) == (runs1, rebuilds1)
            m.topn_batch(kind, q, hm)
            assert counters(m) == (runs1 + 1, rebuilds1 + 1)
            # the rebuilt shadow has room for the store's new slot capacity: further new rows are followed
            extra = random_model(100, f, 0xB20000CA)
            m.upsert(ob.Y, np.arange(n + 200, n + 300, dtype=np.int32), extra * 4.0)   # would win most queries
            b_rows, b_scores, b_cnt = m.topn_batch(kind, q[:5], hm)
            assert counters(m) == (runs1 + 2, rebuilds1 + 1)
            for i in range(5):
                r1, s1 = m.topn(kind, q[i], hm)
                assert b_rows[i, :b_cnt[i]].tolist() == r1.tolist() and b_scores[i, :b_cnt[i]].tobytes() == s1.tobytes()
            assert any(r >= n + 200 for r in b_rows[0, :b_cnt[0]].tolist()) or kind == ob.COSINE


def test_estimate_pairs_exact():
    """oryx_estimate_pairs: VectorMath.dot of (X row, Y row) pairs bit for bit (Estimate.java:62-71; golden
    /estimate/U0/I0/I1/I2 of EstimateTest.java:29-34), and MLlib's binary64-product arithmetic."""
    g = load_golden("als_fixed_model")
    x = np.array(g["model"]["X"], np.float32)
    y = np.array(g["model"]["Y"], np.float32)
    with ob.Model(2) as m:
        m.upsert(ob.X, np.arange(len(x), dtype=np.int32), x)
        m.upsert(ob.Y, np.arange(len(y), dtype=np.int32), y)
        s, f = m.estimate_pairs([0, 0, 0, 0, 99], [0, 1, 2, 77, 1])
        assert f.tolist() == [1, 1, 1, 0, 0] and s[3] == 0.0 and s[4] == 0.0   # unknown item / user -> 0
        for got, want in zip(s[:3], (0.38761318, 0.4653969, -0.0801478)):
            assert abs(got - want) <= 1e-6
    rng = np.random.default_rng(0xB2000100)
    f_, nu, ni, n = 37, 300, 500, 5000
    x = rng.standard_normal((nu, f_)).astype(np.float32)
    y = rng.standard_normal((ni, f_)).astype(np.float32)
    with ob.Model(f_) as m:
        m.upsert(ob.X, np.arange(nu, dtype=np.int32), x)
        m.upsert(ob.Y, np.arange(ni, dtype=np.int32), y)
        u = rng.integers(0, nu, n)
        it = rng.integers(0, ni, n)
        s, f = m.estimate_pairs(u, it)
        assert f.all()
        for k in range(0, n, 97):
            assert s[k] == oracle.dot(x[u[k]], y[it[k]])
        w, _ = m.estimate_pairs(u, it, wide_products=True)
        for k in range(0, n, 97):   # sequential binary64 products and sums
            acc = 0.0
            for a, b in zip(x[u[k]].astype(np.float64), y[it[k]].astype(np.float64)):
                acc += a * b
            assert w[k] == acc


@pytest.mark.parametrize("n,f", [(70000, 48), (5000, 130)])
def test_compaction_in_place_equals_out_of_place(n, f, monkeypatch):
    """oryx_items_compact when a second copy of the store does not fit (forced here): the rows are permuted
    in place by segments of the permutation's paths and cycles.  Every row reads back unchanged, LSH
    queries return what they returned before, and the slot order equals the out-of-place result."""
    rng = np.random.default_rng(0xB2000110)
    y = random_model(n, f, 0xB2000111)
    h, bits = ob.lsh_plan(0.3, 16)
    hv = random_hash_vectors(h, f, 0xB2000112)
    queries = random_model(6, f, 0xB2000113)

    def build():
        m = ob.Model(f, hash_vectors=hv, max_bits_differing=bits)
        m.upsert(ob.Y, np.arange(n, dtype=np.int32), y)
        gone = rng_local.choice(n, n // 10, replace=False).astype(np.int32)   # holes for the paths to end in
        m.remove(ob.Y, gone)
        back = gone[: n // 40]
        m.upsert(ob.Y, back, y[back])                                        # re-used slots
        return m, set(gone[n // 40:].tolist())

    rng_local = np.random.default_rng(7)
    m1, dead = build()
    rng_local = np.random.default_rng(7)
    m2, _ = build()
    before = [m1.topn(ob.DOT, q, 10) for q in queries]
    m1.compact(ob.Y)                                   # out of place
    monkeypatch.setenv("ORYX_COMPACT_IN_PLACE", "1")
    m2.compact(ob.Y)                                   # in place
    monkeypatch.delenv("ORYX_COMPACT_IN_PLACE")
    for q, (r0, s0) in zip(queries, before):
        for m in (m1, m2):
            r, s = m.topn(ob.DOT, q, 10)
            assert r.tolist() == r0.tolist() and s.tobytes() == s0.tobytes()
        assert np.array_equal(m1.scores_all(ob.DOT, q, n), m2.scores_all(ob.DOT, q, n), equal_nan=True)
    for row in list(range(0, n, 211)) + [n - 1]:
        v1, v2 = m1.get(ob.Y, row), m2.get(ob.Y, row)
        if row in dead:
            assert v1 is None and v2 is None
        else:
            assert np.array_equal(v1, y[row]) and np.array_equal(v2, y[row])
    assert m1.stats()["slots"][1] == m2.stats()["slots"][1] == n - len(dead)
    # the compacted store keeps working: updates, another (no-op) compaction in place
    m2.upsert(ob.Y, np.array([1, 2, 3], np.int32), y[[4, 5, 6]])
    monkeypatch.setenv("ORYX_COMPACT_IN_PLACE", "1")
    m2.compact(ob.Y)
    monkeypatch.delenv("ORYX_COMPACT_IN_PLACE")
    assert np.array_equal(m2.get(ob.Y, 2), y[5])
    m1.close(); m2.close()
