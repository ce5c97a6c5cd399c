"""CPU oracle for the ALS serving/speed hot path -- TEST INFRASTRUCTURE ONLY.

ctypes binding over ``oracle/liboryx_oracle.so`` (built from ``oryx_oracle.c``, a plain-C
restatement of the reference's Java arithmetic; every C function cites the reference file:line it
follows).  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; the product (``oryx_b200``) never does.

Parity status: pinned -- ``tests/test_oracle_golden.py`` replays the reference's own golden
vectors (``tests/golden/``) against these functions.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboryx_oracle.so")

DOT, COSINE = 0, 1


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oryx_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboryx_oracle.so"])
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)
        i32p, i64p, u8p = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_uint8)
        L.oro_dot.restype = C.c_double
        L.oro_dot.argtypes = [f32p, f32p, C.c_int]
        L.oro_norm.restype = C.c_double
        L.oro_norm.argtypes = [f32p, C.c_int]
        L.oro_transpose_times_self.argtypes = [f32p, C.c_int64, C.c_int, f64p]
        L.oro_lsh_plan.argtypes = [C.c_double, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.oro_lsh_index.restype = C.c_int
        L.oro_lsh_index.argtypes = [f32p, C.c_int, C.c_int, f32p]
        L.oro_lsh_prototype.argtypes = [C.c_int, i32p]
        L.oro_lsh_candidates.restype = C.c_int
        L.oro_lsh_candidates.argtypes = [C.c_int, C.c_int, C.c_int, i32p]
        L.oro_lsh_index_rows.argtypes = [f32p, C.c_int, C.c_int, f32p, C.c_int64, i32p]
        L.oro_dots_mean.argtypes = [f32p, C.c_int, C.c_int, f32p]
        L.oro_cosine_average.argtypes = [f32p, C.c_int, C.c_int, C.c_int, f32p]
        L.oro_cosine_score.restype = C.c_double
        L.oro_cosine_score.argtypes = [f32p, f32p, C.c_int]
        L.oro_scores_all.argtypes = [C.c_int, f32p, f32p, C.c_int64, C.c_int, f64p]
        L.oro_topn_rows.restype = C.c_int
        L.oro_topn_rows.argtypes = [C.c_int, f32p, f32p, C.c_int64, C.c_int, i32p, i32p, C.c_int,
                                    i32p, C.c_int, C.c_int, i64p, f64p]
        L.oro_topn_values.restype = C.c_int
        L.oro_topn_values.argtypes = [f64p, C.c_int64, C.c_int, i64p, f64p]
        L.oro_faithful_create.restype = C.c_void_p
        L.oro_faithful_create.argtypes = [C.c_int, C.c_int]
        L.oro_faithful_put_rows.restype = None
        L.oro_faithful_put_rows.argtypes = [C.c_void_p, i32p, C.c_char_p, f32p, C.c_int64]
        L.oro_faithful_destroy.restype = None
        L.oro_faithful_destroy.argtypes = [C.c_void_p]
        L.oro_faithful_topn.restype = C.c_int
        L.oro_faithful_topn.argtypes = [C.c_void_p, C.c_int, f32p, C.c_int, i32p, C.c_int, C.POINTER(C.c_char_p),
                                        C.c_int, C.c_int, C.POINTER(C.c_int64), f64p]
        L.oro_topn_mt.restype = C.c_int
        L.oro_topn_mt.argtypes = [C.c_int, f32p, f32p, C.c_int64, C.c_int, i32p, C.c_int, C.c_int,
                                  C.c_int, i64p, f64p]
        L.oro_solver_create.restype = C.c_int
        L.oro_solver_create.argtypes = [f64p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int)]
        L.oro_solver_destroy.argtypes = [C.c_void_p]
        L.oro_is_non_singular.restype = C.c_int
        L.oro_is_non_singular.argtypes = [f64p, C.c_int]
        L.oro_solver_solve.argtypes = [C.c_void_p, f64p, f64p]
        L.oro_solve_f2f.argtypes = [C.c_void_p, f32p, f32p]
        L.oro_compute_target_qui.restype = C.c_double
        L.oro_compute_target_qui.argtypes = [C.c_int, C.c_double, C.c_double]
        L.oro_compute_updated_xu.restype = C.c_int
        L.oro_compute_updated_xu.argtypes = [C.c_void_p, C.c_double, f32p, f32p, C.c_int, C.c_int,
                                             f32p]
        L.oro_fold_in_batch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int64, f32p, u8p, f32p,
                                        f64p, f32p, u8p]
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def dot(x, y) -> float:
    x, y = _f32(x), _f32(y)
    return lib().oro_dot(_p(x, C.c_float), _p(y, C.c_float), x.size)


def norm(x) -> float:
    x = _f32(x)
    return lib().oro_norm(_p(x, C.c_float), x.size)


def transpose_times_self(m) -> np.ndarray:
    m = _f32(m)
    n, f = m.shape
    out = np.zeros((f, f), dtype=np.float64)
    lib().oro_transpose_times_self(_p(m, C.c_float), n, f, _p(out, C.c_double))
    return out


def transpose_times_self_mt(m, threads: int) -> np.ndarray:
    """getYTYSolver's sum of per-partition transposeTimesSelf results (ALSServingModel.java:357-373):
    contiguous row slices, one per thread, partials added in slice order."""
    m = _f32(m)
    n, f = m.shape
    out = np.zeros((f, f), dtype=np.float64)
    L = lib()
    L.oro_transpose_times_self_mt.restype = None
    L.oro_transpose_times_self_mt.argtypes = [C.POINTER(C.c_float), C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_double)]
    L.oro_transpose_times_self_mt(_p(m, C.c_float), n, f, threads, _p(out, C.c_double))
    return out


def lsh_plan(sample_rate: float, num_cores: int):
    h, b = C.c_int(), C.c_int()
    lib().oro_lsh_plan(sample_rate, num_cores, C.byref(h), C.byref(b))
    return h.value, b.value


def lsh_index(hash_vectors, v) -> int:
    hv = _f32(hash_vectors)
    v = _f32(v)
    h = hv.shape[0] if hv.ndim == 2 else 0
    if h == 0:
        return 0
    return lib().oro_lsh_index(_p(hv, C.c_float), h, v.size, _p(v, C.c_float))


def lsh_index_rows(hash_vectors, rows) -> np.ndarray:
    hv = _f32(hash_vectors)
    rows = _f32(rows)
    n, f = rows.shape
    out = np.zeros(n, dtype=np.int32)
    h = hv.shape[0] if hv.ndim == 2 else 0
    if h:
        lib().oro_lsh_index_rows(_p(hv, C.c_float), h, f, _p(rows, C.c_float), n, _p(out, C.c_int32))
    return out


def lsh_prototype(num_hashes: int) -> np.ndarray:
    out = np.zeros(1 << num_hashes, dtype=np.int32)
    lib().oro_lsh_prototype(num_hashes, _p(out, C.c_int32))
    return out


def lsh_candidates(num_hashes: int, max_bits: int, main_index: int) -> np.ndarray:
    out = np.zeros(1 << num_hashes, dtype=np.int32)
    n = lib().oro_lsh_candidates(num_hashes, max_bits, main_index, _p(out, C.c_int32))
    return out[:n].copy()


def dots_mean(vecs) -> np.ndarray:
    vecs = _f32(vecs)
    n, f = vecs.shape
    out = np.zeros(f, dtype=np.float32)
    lib().oro_dots_mean(_p(vecs, C.c_float), n, f, _p(out, C.c_float))
    return out


def cosine_average(vecs, single_ctor: bool | None = None) -> np.ndarray:
    vecs = _f32(vecs)
    if vecs.ndim == 1:
        vecs = vecs[None, :]
        single_ctor = True if single_ctor is None else single_ctor
    n, f = vecs.shape
    if single_ctor is None:
        single_ctor = False
    out = np.zeros(f, dtype=np.float32)
    lib().oro_cosine_average(_p(vecs, C.c_float), n, f, int(single_ctor), _p(out, C.c_float))
    return out


def cosine_score(q, y) -> float:
    q, y = _f32(q), _f32(y)
    return lib().oro_cosine_score(_p(q, C.c_float), _p(y, C.c_float), q.size)


def scores_all(kind: int, q, y) -> np.ndarray:
    q, y = _f32(q), _f32(y)
    n, f = y.shape
    out = np.zeros(n, dtype=np.float64)
    lib().oro_scores_all(kind, _p(q, C.c_float), _p(y, C.c_float), n, f, _p(out, C.c_double))
    return out


def topn(kind: int, q, y, how_many: int, part=None, cand=None, excluded=None):
    """ALSServingModel.topN over rows of ``y``; returns (rows int64[], scores float64[])."""
    q, y = _f32(q), _f32(y)
    n, f = y.shape if y.ndim == 2 else (0, q.size)
    part_a = None if part is None else np.ascontiguousarray(part, dtype=np.int32)
    cand_a = None if cand is None else np.ascontiguousarray(cand, dtype=np.int32)
    excl_a = None if excluded is None else np.ascontiguousarray(np.sort(np.asarray(excluded)), dtype=np.int32)
    rows = np.zeros(max(how_many, 1), dtype=np.int64)
    scores = np.zeros(max(how_many, 1), dtype=np.float64)
    k = lib().oro_topn_rows(kind, _p(q, C.c_float), _p(y, C.c_float), n, f,
                            _p(part_a, C.c_int32), _p(cand_a, C.c_int32),
                            0 if cand_a is None else cand_a.size,
                            _p(excl_a, C.c_int32), 0 if excl_a is None else excl_a.size,
                            how_many, _p(rows, C.c_int64), _p(scores, C.c_double))
    return rows[:k].copy(), scores[:k].copy()


def topn_mt(kind: int, q, y, how_many: int, threads: int, excluded=None):
    q, y = _f32(q), _f32(y)
    n, f = y.shape
    excl_a = None if excluded is None else np.ascontiguousarray(np.sort(np.asarray(excluded)), dtype=np.int32)
    rows = np.zeros(max(how_many, 1), dtype=np.int64)
    scores = np.zeros(max(how_many, 1), dtype=np.float64)
    k = lib().oro_topn_mt(kind, _p(q, C.c_float), _p(y, C.c_float), n, f,
                          _p(excl_a, C.c_int32), 0 if excl_a is None else excl_a.size,
                          how_many, threads, _p(rows, C.c_int64), _p(scores, C.c_double))
    return rows[:k].copy(), scores[:k].copy()


class FaithfulModel:
    """The "faithful" arm of the CPU baseline (oryx_oracle.c, section "faithful CPU baseline"): one
    String-keyed hash map of heap-allocated float[] per LSH partition, String-set known-items
    predicate, one task per candidate partition on a thread pool -- the structure of
    ALSServingModel.topN (ALSServingModel.java:289-335).  Item ids are ``"I<row>"``."""

    def __init__(self, y, part=None, n_partitions: int = 1):
        y = _f32(y)
        self.n, self.f = y.shape
        self.n_partitions = n_partitions
        self._h = C.c_void_p(lib().oro_faithful_create(self.f, n_partitions))
        part_a = None if part is None else np.ascontiguousarray(part, dtype=np.int32)
        lib().oro_faithful_put_rows(self._h, _p(part_a, C.c_int32), b"I", _p(y, C.c_float), self.n)

    def topn(self, kind: int, q, how_many: int, cand=None, known_rows=(), threads: int = 1):
        q = _f32(q)
        cand_a = np.ascontiguousarray(np.arange(self.n_partitions) if cand is None else cand, dtype=np.int32)
        ids = [b"I%d" % int(r) for r in known_rows]
        arr = (C.c_char_p * max(len(ids), 1))(*ids)
        rows = np.zeros(max(how_many, 1), dtype=np.int64)
        scores = np.zeros(max(how_many, 1), dtype=np.float64)
        k = lib().oro_faithful_topn(self._h, kind, _p(q, C.c_float), how_many, _p(cand_a, C.c_int32), cand_a.size,
                                    arr, len(ids), threads, _p(rows, C.c_int64), _p(scores, C.c_double))
        return rows[:k].copy(), scores[:k].copy()

    def close(self):
        if self._h:
            lib().oro_faithful_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def topn_values(scores, how_many: int):
    s = np.ascontiguousarray(scores, dtype=np.float64)
    rows = np.zeros(max(how_many, 1), dtype=np.int64)
    out = np.zeros(max(how_many, 1), dtype=np.float64)
    k = lib().oro_topn_values(_p(s, C.c_double), s.size, how_many, _p(rows, C.c_int64), _p(out, C.c_double))
    return rows[:k].copy(), out[:k].copy()


class SingularMatrixSolverException(Exception):
    """Mirror of com.cloudera.oryx.common.math.SingularMatrixSolverException."""

    def __init__(self, apparent_rank: int):
        super().__init__(f"Apparent rank: {apparent_rank}")
        self.apparent_rank = apparent_rank


class Solver:
    """LinearSystemSolver.getSolver(M) -> Solver (LinearSystemSolver.java:35-55)."""

    def __init__(self, m):
        m = np.ascontiguousarray(m, dtype=np.float64)
        assert m.ndim == 2 and m.shape[0] == m.shape[1]
        self.n = m.shape[0]
        h = C.c_void_p()
        rank = C.c_int()
        rc = lib().oro_solver_create(_p(m, C.c_double), self.n, C.byref(h), C.byref(rank))
        if rc != 0:
            raise SingularMatrixSolverException(rank.value)
        self._h = h

    def solve(self, b) -> np.ndarray:
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(self.n, dtype=np.float64)
        lib().oro_solver_solve(self._h, _p(b, C.c_double), _p(x, C.c_double))
        return x

    def solve_f2f(self, b) -> np.ndarray:
        b = _f32(b)
        x = np.zeros(self.n, dtype=np.float32)
        lib().oro_solve_f2f(self._h, _p(b, C.c_float), _p(x, C.c_float))
        return x

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and _lib is not None:
            _lib.oro_solver_destroy(h)
            self._h = None


def is_non_singular(m) -> bool:
    m = np.ascontiguousarray(m, dtype=np.float64)
    return bool(lib().oro_is_non_singular(_p(m, C.c_double), m.shape[0]))


def compute_target_qui(implicit: bool, value: float, current: float) -> float:
    return lib().oro_compute_target_qui(int(implicit), value, current)


def compute_updated_xu(solver: Solver, value: float, xu, yi, implicit: bool):
    if yi is None:
        return None
    yi = _f32(yi)
    xu_a = None if xu is None else _f32(xu)
    out = np.zeros(yi.size, dtype=np.float32)
    ok = lib().oro_compute_updated_xu(solver._h, value, _p(xu_a, C.c_float), _p(yi, C.c_float),
                                      yi.size, int(implicit), _p(out, C.c_float))
    return out if ok else None


def fold_in_batch(solver: Solver, implicit: bool, xu, has_xu, yi, values):
    yi = _f32(yi)
    n, f = yi.shape
    xu = _f32(xu if xu is not None else np.zeros_like(yi))
    has = None if has_xu is None else np.ascontiguousarray(has_xu, dtype=np.uint8)
    vals = np.ascontiguousarray(values, dtype=np.float64)
    out = np.zeros((n, f), dtype=np.float32)
    valid = np.zeros(n, dtype=np.uint8)
    lib().oro_fold_in_batch(solver._h, int(implicit), f, n, _p(xu, C.c_float), _p(has, C.c_uint8),
                            _p(yi, C.c_float), _p(vals, C.c_double), _p(out, C.c_float),
                            _p(valid, C.c_uint8))
    return out, valid
