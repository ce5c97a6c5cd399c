"""Pins the CPU oracle (oracle/oryx_oracle.c) against the reference's own golden vectors.

Every fixture under tests/golden/ is transcribed from a reference JUnit test (citation inside the
JSON).  CPU-only: runs under -m "not gpu".
"""
import math

import numpy as np
import pytest

import oracle
from als_harness import (OracleBackend, ServingHarness, check_fixed_model_goldens, load_golden,
                         random_hash_vectors)

M = load_golden("math")
L = load_golden("lsh")


def test_dot_goldens():
    for c in M["dot"]:
        expected = c["value"] if "value" in c else 1e-24 * 1e-24
        got = oracle.dot(c["x"], c["y"])
        if "value_expr" in c:
            # VectorMathTest.testSmall: the binary32 product 1e-24f*1e-24f underflows to 0, the
            # expectation is the double 1e-48; OryxTest.assertEquals(double,double) compares with
            # DOUBLE_EPSILON = 1e-12 (OryxTest.java:41-42), which both satisfy.
            assert abs(got - expected) <= 1e-12
        else:
            assert abs(got - expected) <= c["tol"]


def test_norm_goldens():
    for c in M["norm"]:
        assert abs(oracle.norm(c["x"]) - c["value"]) <= c["tol"]


def test_transpose_times_self_goldens():
    for c in M["transpose_times_self"]:
        got = oracle.transpose_times_self(np.asarray(c["rows"], dtype=np.float32))
        exp = np.asarray(c["value"])
        if c["tol"] == 0.0:
            assert np.array_equal(got, exp), c["_cite"]
        else:
            assert np.max(np.abs(got - exp)) <= c["tol"], c["_cite"]


def test_solver_goldens():
    s = M["solver"]["solve_f2f"]
    x = oracle.Solver(s["a"]).solve_f2f(s["b"])
    assert np.array_equal(x, np.asarray(s["x"], dtype=np.float32)), (x, s["x"])
    for c in M["solver"]["non_singular"]:
        assert oracle.is_non_singular(c["a"]) == c["value"], c["_cite"]
    for c in M["solver"]["apparent_rank"]:
        with pytest.raises(oracle.SingularMatrixSolverException) as ei:
            oracle.Solver(c["a"])
        assert ei.value.apparent_rank == c["rank"], c["_cite"]


def test_target_qui_goldens():
    for implicit, value, cur, exp in M["target_qui"]["cases"]:
        got = oracle.compute_target_qui(implicit, value, cur)
        if exp == "nan":
            assert math.isnan(got)
        else:
            assert got == exp


def test_updated_xu_goldens():
    u = M["updated_xu"]
    solver = oracle.Solver(oracle.transpose_times_self(np.asarray(u["rows"], dtype=np.float32)))
    for c in u["cases"]:
        got = oracle.compute_updated_xu(solver, c["value"], c["xu"], c["yi"], c["implicit"])
        if c["out"] is None:
            assert got is None
        else:
            assert np.array_equal(got, np.asarray(c["out"], dtype=np.float32)), (got, c["out"])


def test_topn_consumer_golden():
    t = M["topn_consumer"]
    rng = np.random.default_rng(1234)
    for _ in range(20):
        vals = rng.permutation(t["num_values"]).astype(np.float64)
        rows, scores = oracle.topn_values(vals, t["how_many"])
        assert list(scores) == [float(v) for v in t["expected"]]
        assert [vals[r] for r in rows] == [float(v) for v in t["expected"]]


def test_topn_rejects_nan_and_is_strict():
    # TopNConsumer.java:60-64: strict '>' admission; NaN (rescorer's exclude signal) never enters.
    vals = np.array([1.0, float("nan"), 3.0, 3.0, 2.0, float("nan")])
    rows, scores = oracle.topn_values(vals, 10)
    assert list(scores) == [3.0, 3.0, 2.0, 1.0]
    rows, scores = oracle.topn_values(np.array([5.0, 5.0, 5.0]), 2)
    assert list(scores) == [5.0, 5.0] and list(rows) == [0, 1]  # third equal value is not admitted


def test_lsh_plan_goldens():
    for rate, cores, hashes, bits in L["plan"]["cases"]:
        assert oracle.lsh_plan(rate, cores) == (hashes, bits), (rate, cores)


def test_lsh_candidates_goldens():
    c = L["candidates_no_sample"]
    h, b = oracle.lsh_plan(c["sample_rate"], c["cores"])
    cand = oracle.lsh_candidates(h, b, 0)
    assert list(cand) == list(range(1 << h))

    c = L["candidates_one_bit"]
    h, b = oracle.lsh_plan(c["sample_rate"], c["cores"])
    assert b == 1
    hv = random_hash_vectors(h, c["features"], seed=7)
    zero = np.zeros(c["features"], dtype=np.float32)
    assert oracle.lsh_index(hv, zero) == 0  # exact zero dot is not > 0
    zc = oracle.lsh_candidates(h, b, oracle.lsh_index(hv, zero))
    assert len(zc) == 1 + h and zc[0] == 0
    assert all(zc[i] == 1 << (i - 1) for i in range(1, len(zc)))
    one = np.ones(c["features"], dtype=np.float32)
    oc = oracle.lsh_candidates(h, b, oracle.lsh_index(hv, one))
    assert all(oc[i] == oc[0] ^ (1 << (i - 1)) for i in range(1, len(oc)))

    c = L["candidates_bands"]
    h, b = oracle.lsh_plan(c["sample_rate"], c["cores"])
    assert (h, b) == (7, 3)
    hv = random_hash_vectors(h, c["features"], seed=8)
    cand = oracle.lsh_candidates(h, b, oracle.lsh_index(hv, np.ones(c["features"], dtype=np.float32)))
    assert len(cand) == c["count"]
    for lo, hi, bits in c["bands"]:
        for i in range(lo, hi):
            assert bin(int(cand[0]) ^ int(cand[i])).count("1") == bits


def test_lsh_hash_distribution():
    d = L["hash_distribution"]
    rng = np.random.default_rng(99)
    for features, rate, cores in d["cases"]:
        h, _ = oracle.lsh_plan(rate, cores)
        hv = random_hash_vectors(h, features, seed=features * 131 + cores)
        vecs = rng.standard_normal((d["trials"], features), dtype=np.float32)
        idx = oracle.lsh_index_rows(hv, vecs)
        counts = np.bincount(idx, minlength=1 << h)
        assert counts.sum() == d["trials"]
        assert counts.max() <= 2 * counts.min(), (features, rate, cores, counts.min(), counts.max())


def test_fixed_model_endpoint_goldens():
    g = load_golden("als_fixed_model")
    h = ServingHarness(g["model"], lambda f, y: OracleBackend(f, y))
    check_fixed_model_goldens(h)


def test_speed_fold_in_goldens():
    s = load_golden("speed")
    X = {k: np.asarray(v, dtype=np.float32) for k, v in s["X"].items()}
    Y = {k: np.asarray(v, dtype=np.float32) for k, v in s["Y"].items()}
    # ALSSpeedModelManager.buildUpdates hot section (ALSSpeedModelManager.java:173-204)
    xtx = oracle.Solver(oracle.transpose_times_self(np.stack(list(X.values()))))
    yty = oracle.Solver(oracle.transpose_times_self(np.stack(list(Y.values()))))
    for user, item, value in s["interactions"]:
        xu, yi = X.get(user), Y.get(item)
        new_xu = oracle.compute_updated_xu(yty, value, xu, yi, s["implicit"])
        new_yi = oracle.compute_updated_xu(xtx, value, yi, xu, s["implicit"])
        if user in s["expected_X"]:
            assert new_yi is None
            assert np.max(np.abs(new_xu - np.asarray(s["expected_X"][user], dtype=np.float32))) <= s["tol"]
        else:
            assert new_xu is None
            assert np.max(np.abs(new_yi - np.asarray(s["expected_Y"][item], dtype=np.float32))) <= s["tol"]


def test_numpy_cross_check():
    """Independent numpy restatement of the float-multiply / double-accumulate recipe."""
    rng = np.random.default_rng(5)
    y = rng.standard_normal((500, 37), dtype=np.float32)
    q = rng.standard_normal(37, dtype=np.float32)
    acc = np.zeros(500, dtype=np.float64)
    for i in range(37):
        acc += (y[:, i] * q[i]).astype(np.float64)  # fp32 product, fp64 add, index order
    assert np.array_equal(acc, oracle.scores_all(oracle.DOT, q, y))
    nrm = np.zeros(500, dtype=np.float64)
    for i in range(37):
        nrm += (y[:, i] * y[:, i]).astype(np.float64)
    assert np.array_equal(acc / np.sqrt(nrm), oracle.scores_all(oracle.COSINE, q, y))
    g = np.zeros((37, 37))
    for r in range(500):
        g += np.outer(y[r], y[r]).astype(np.float64)  # np.outer on float32 gives float32 products
    assert np.array_equal(g, oracle.transpose_times_self(y))
    rows, scores = oracle.topn(oracle.DOT, q, y, 10)
    order = np.argsort(-acc, kind="stable")[:10]
    assert list(rows) == list(order) and np.array_equal(scores, acc[order])
    r2, s2 = oracle.topn_mt(oracle.DOT, q, y, 10, threads=4)
    assert list(r2) == list(rows) and np.array_equal(s2, scores)


def test_faithful_layout_arm_agrees_with_the_row_ordered_oracle():
    """The "faithful" CPU-baseline arm (String-keyed partition maps, String-set predicate, task per
    partition) returns what the plain restatement returns -- with and without LSH candidates."""
    from als_harness import random_hash_vectors, random_model
    y = random_model(30000, 24, 0xB2000070)
    hv = random_hash_vectors(3, 24, 0xB2000071)
    part = oracle.lsh_index_rows(hv, y)
    fm = oracle.FaithfulModel(y, part, 8)
    rng = np.random.default_rng(0xB2000072)
    for i in range(6):
        q = rng.standard_normal(24).astype(np.float32)
        known = rng.integers(0, 30000, 25)
        for kind in (oracle.DOT, oracle.COSINE):
            r1, s1 = oracle.topn(kind, q, y, 10, excluded=np.unique(known))
            r2, s2 = fm.topn(kind, q, 10, known_rows=known, threads=3)
            assert r1.tolist() == r2.tolist() and s1.tobytes() == s2.tobytes()
        cand = oracle.lsh_candidates(3, 1, oracle.lsh_index(hv, q))
        r1, s1 = oracle.topn(oracle.DOT, q, y, 10, part=part, cand=cand)
        r2, s2 = fm.topn(oracle.DOT, q, 10, cand=cand, threads=3)
        assert r1.tolist() == r2.tolist() and s1.tobytes() == s2.tobytes()
    fm.close()
