"""ctypes binding over the C ABI in include/oryx_b200.h (see that header for the reference
counterpart of every entry point)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboryx_b200.so")

OK, E_INVALID, E_CUDA, E_SINGULAR, E_NOMEM, E_UNSUPPORTED = 0, 1, 2, 3, 4, 5
DOT, COSINE = 0, 1
X, Y = 0, 1
MAX_HOW_MANY = 1016          # fused scan
MAX_HOW_MANY_DEEP = 65536    # oryx_topn / oryx_recommend (exact select-and-sort path above MAX_HOW_MANY)
STATIC_MAX_HOW_MANY = 56


class OryxError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"oryx_b200 status {status}: {message}")
        self.status = status


class SingularMatrixSolverException(OryxError):
    """ORYX_E_SINGULAR, mirror of com.cloudera.oryx.common.math.SingularMatrixSolverException."""

    def __init__(self, apparent_rank: int, message: str):
        super().__init__(E_SINGULAR, message)
        self.apparent_rank = apparent_rank


class _Config(C.Structure):
    _fields_ = [("features", C.c_int32), ("implicit", C.c_int32), ("num_hashes", C.c_int32),
                ("max_bits_differing", C.c_int32), ("hash_vectors", C.POINTER(C.c_float)),
                ("device", C.c_int32), ("reserved_sms", C.c_int32), ("capacity_hint", C.c_int64)]


class _Stats(C.Structure):
    _fields_ = [("rows", C.c_int64 * 2), ("slots", C.c_int64 * 2), ("device_bytes", C.c_int64),
                ("queries", C.c_int64), ("kernel_launches", C.c_int64), ("last_query_ms", C.c_double)]


def build(verbose: bool = False) -> str:
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j", str(os.cpu_count() or 4)]
    if not verbose:
        cmd.insert(1, "-s")
    subprocess.check_call(cmd)
    return _SO


_lib = None

# name -> (restype, argtypes); the list every test checks against include/oryx_b200.h
_f32p, _f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)
_i32p, _i64p, _u8p = C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_uint8)
SIGNATURES = {
    "oryx_lsh_plan": [C.c_double, C.c_int32, _i32p, _i32p],
    "oryx_model_create": [C.POINTER(_Config), C.POINTER(C.c_void_p)],
    "oryx_model_destroy": [C.c_void_p],
    "oryx_items_upsert": [C.c_void_p, C.c_int32, _i32p, _f32p, C.c_int32, _i32p],
    "oryx_items_upsert_device": [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64],
    "oryx_items_remove": [C.c_void_p, C.c_int32, _i32p, C.c_int32],
    "oryx_items_get": [C.c_void_p, C.c_int32, C.c_int32, _f32p, _i32p],
    "oryx_items_compact": [C.c_void_p, C.c_int32],
    "oryx_lsh_index": [C.c_void_p, _f32p, _i32p],
    "oryx_lsh_index_rows": [C.c_void_p, _f32p, C.c_int32, _i32p],
    "oryx_lsh_candidates": [C.c_void_p, _f32p, _i32p, _i32p],
    "oryx_topn": [C.c_void_p, C.c_int32, _f32p, C.c_int32, _i32p, C.c_int32, _i32p, _f64p, _i32p],
    "oryx_known_items_add": [C.c_void_p, _i32p, _i64p, _i32p, C.c_int32],
    "oryx_known_items_get": [C.c_void_p, C.c_int32, _i32p, C.c_int32, _i32p],
    "oryx_known_items_retain": [C.c_void_p, _u8p, C.c_int32, _u8p, C.c_int32],
    "oryx_known_counts": [C.c_void_p, C.c_int32, _i32p, C.c_int32],
    "oryx_recommend": [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, _i32p, _f64p, _i32p, _i32p],
    "oryx_recommend_batch": [C.c_void_p, _i32p, C.c_int32, C.c_int32, C.c_int32, _i32p, _f64p, _i32p],
    "oryx_recommend_coalesced": [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, _i32p, _f64p, _i32p, _i32p],
    "oryx_topn_coalesced": [C.c_void_p, C.c_int32, _f32p, C.c_int32, _i32p, C.c_int32, _i32p, _f64p, _i32p],
    "oryx_topn_batch": [C.c_void_p, C.c_int32, _f32p, C.c_int32, C.c_int32, _i32p, _i32p, _i32p, _f64p, _i32p],
    "oryx_scores_all": [C.c_void_p, C.c_int32, _f32p, C.c_int32, _f64p],
    "oryx_estimate_pairs": [C.c_void_p, _i32p, _i32p, C.c_int64, C.c_int32, _f64p, _u8p],
    "oryx_scores_rows": [C.c_void_p, C.c_int32, _f32p, _i32p, C.c_int32, _f64p, _u8p],
    "oryx_topn_device": [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "oryx_topn_device_static": [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "oryx_merge_topn_device": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "oryx_merge_topn_batch_device": [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p],
    "oryx_merge_topn_packed": [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p],
    "oryx_exchange_create": [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)],
    "oryx_exchange_handle": [C.c_void_p, C.c_char_p],
    "oryx_exchange_connect": [C.c_void_p, C.c_char_p],
    "oryx_exchange_connect_local": [C.POINTER(C.c_void_p), C.c_int32],
    "oryx_exchange_destroy": [C.c_void_p],
    "oryx_topn_sharded_device": [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p],
    "oryx_fold_in_rows": [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, _i32p, _i32p, _f64p, C.c_int64, _f32p, _u8p],
    "oryx_fold_in_rows_sharded": [C.c_void_p, C.POINTER(C.c_void_p), C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                  _i64p, _i64p, _f64p, C.c_int64, _f32p, _u8p],
    "oryx_peer_stores_create": [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)],
    "oryx_peer_stores_handle": [C.c_void_p, C.c_char_p],
    "oryx_peer_stores_connect": [C.c_void_p, C.c_char_p],
    "oryx_peer_stores_destroy": [C.c_void_p],
    "oryx_fold_in_rows_peers": [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32, _i64p, _i64p, _f64p,
                                C.c_int64, _f32p, _u8p],
    "oryx_gram": [C.c_void_p, C.c_int32, _f64p, _i64p],
    "oryx_gram_allreduce_device": [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p],
    "oryx_gram_sharded": [C.c_void_p, C.c_void_p, C.c_int32, _f64p, _i64p],
    "oryx_solver_create": [_f64p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p), _i32p],
    "oryx_solver_destroy": [C.c_void_p],
    "oryx_solve_f2f": [C.c_void_p, _f32p, _f32p],
    "oryx_fold_in_batch": [C.c_void_p, C.c_int32, _f32p, _u8p, _f32p, _f64p, C.c_int64, _f32p, _u8p],
    "oryx_aggregate_interactions": [C.c_void_p, _i32p, _i32p, _f64p, _i64p, C.c_int64, C.c_int32, _i32p, _i32p, _f64p, _i64p],
    "oryx_last_error": [C.c_char_p, C.c_int32],
    "oryx_stats_get": [C.c_void_p, C.POINTER(_Stats)],
}


# development aids exported by the library but not part of include/oryx_b200.h (timings for bench.py)
DEBUG_SIGNATURES = {
    "oryx_debug_batch_timings": [C.c_void_p, _f64p],
    "oryx_debug_gram_ms": [C.c_void_p, _f64p],
    "oryx_debug_fold_in_timings": [C.c_void_p, _f64p],
}


def lib() -> C.CDLL:
    """Load liboryx_b200.so.  There is no fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise ImportError(
                f"{_SO} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C oryx_b200/csrc`).  oryx_b200 has no CPU fallback.")
        L = C.CDLL(_SO)
        for name, argtypes in list(SIGNATURES.items()) + list(DEBUG_SIGNATURES.items()):
            fn = getattr(L, name)
            fn.restype = C.c_int32
            fn.argtypes = argtypes
        _lib = L
    return _lib


def _last_error() -> str:
    buf = C.create_string_buffer(512)
    lib().oryx_last_error(buf, 512)
    return buf.value.decode("utf-8", "replace")


def _check(status: int):
    if status != OK:
        raise OryxError(status, _last_error())


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def lsh_plan(sample_rate: float, num_cores: int):
    h, b = C.c_int32(), C.c_int32()
    _check(lib().oryx_lsh_plan(sample_rate, num_cores, C.byref(h), C.byref(b)))
    return h.value, b.value


class Model:
    """Handle over oryx_model: the device-resident X/Y stores plus the LSH geometry."""

    def __init__(self, features: int, implicit: bool = True, hash_vectors=None, max_bits_differing: int = 0,
                 device: int = 0, capacity_hint: int = 0, reserved_sms: int = 0):
        hv = None if hash_vectors is None or len(hash_vectors) == 0 else _f32(hash_vectors)
        cfg = _Config()
        cfg.features = features
        cfg.implicit = int(implicit)
        cfg.num_hashes = 0 if hv is None else hv.shape[0]
        cfg.max_bits_differing = max_bits_differing
        cfg.hash_vectors = _ptr(hv, C.c_float)
        cfg.device = device
        cfg.reserved_sms = reserved_sms
        cfg.capacity_hint = capacity_hint
        self.features = features
        self.num_hashes = cfg.num_hashes
        self.max_bits = max_bits_differing
        self._h = C.c_void_p()
        _check(lib().oryx_model_create(C.byref(cfg), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            lib().oryx_model_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- store
    def upsert(self, which: int, rows, vecs) -> np.ndarray:
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        vecs = _f32(vecs).reshape(rows.size, self.features)
        parts = np.zeros(rows.size, dtype=np.int32)
        _check(lib().oryx_items_upsert(self._h, which, _ptr(rows, C.c_int32), _ptr(vecs, C.c_float),
                                       rows.size, _ptr(parts, C.c_int32)))
        return parts

    def upsert_device(self, which: int, first_row: int, device_ptr: int, n: int):
        _check(lib().oryx_items_upsert_device(self._h, which, first_row, C.c_void_p(device_ptr), n))

    def remove(self, which: int, rows):
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        _check(lib().oryx_items_remove(self._h, which, _ptr(rows, C.c_int32), rows.size))

    def get(self, which: int, row: int):
        out = np.zeros(self.features, dtype=np.float32)
        found = C.c_int32()
        _check(lib().oryx_items_get(self._h, which, row, _ptr(out, C.c_float), C.byref(found)))
        return out if found.value else None

    def compact(self, which: int = Y):
        _check(lib().oryx_items_compact(self._h, which))

    # ---- LSH
    def lsh_index(self, vec) -> int:
        v = _f32(vec)
        out = C.c_int32()
        _check(lib().oryx_lsh_index(self._h, _ptr(v, C.c_float), C.byref(out)))
        return out.value

    def lsh_index_rows(self, vecs) -> np.ndarray:
        v = _f32(vecs).reshape(-1, self.features)
        out = np.zeros(v.shape[0], dtype=np.int32)
        _check(lib().oryx_lsh_index_rows(self._h, _ptr(v, C.c_float), v.shape[0], _ptr(out, C.c_int32)))
        return out

    def lsh_candidates(self, vec) -> np.ndarray:
        v = _f32(vec)
        out = np.zeros(1 << self.num_hashes, dtype=np.int32)
        n = C.c_int32()
        _check(lib().oryx_lsh_candidates(self._h, _ptr(v, C.c_float), _ptr(out, C.c_int32), C.byref(n)))
        return out[:n.value].copy()

    # ---- queries
    def topn(self, kind: int, query, how_many: int, excluded=None):
        q = _f32(query)
        ex = None if excluded is None or len(excluded) == 0 else np.ascontiguousarray(excluded, dtype=np.int32)
        rows = np.zeros(max(how_many, 1), dtype=np.int32)
        scores = np.zeros(max(how_many, 1), dtype=np.float64)
        n = C.c_int32()
        _check(lib().oryx_topn(self._h, kind, _ptr(q, C.c_float), how_many, _ptr(ex, C.c_int32),
                               0 if ex is None else ex.size, _ptr(rows, C.c_int32),
                               _ptr(scores, C.c_double), C.byref(n)))
        return rows[:n.value].copy(), scores[:n.value].copy()

    # ---- known items on the device (ALSServingModel.java:189-258) and /recommend by user row
    def known_items_add(self, user_rows, item_lists):
        users = np.ascontiguousarray(user_rows, dtype=np.int32)
        begin = np.zeros(users.size + 1, dtype=np.int64)
        if users.size:
            np.cumsum([len(l) for l in item_lists], out=begin[1:])
        items = np.ascontiguousarray(np.concatenate([np.asarray(l, dtype=np.int32) for l in item_lists])
                                     if users.size and begin[-1] else np.zeros(1, np.int32), dtype=np.int32)
        _check(lib().oryx_known_items_add(self._h, _ptr(users, C.c_int32), _ptr(begin, C.c_int64),
                                          _ptr(items, C.c_int32), users.size))

    def known_items_get(self, user_row: int) -> np.ndarray:
        n = C.c_int32()
        lib().oryx_known_items_get(self._h, user_row, None, 0, C.byref(n))
        out = np.zeros(max(n.value, 1), dtype=np.int32)
        _check(lib().oryx_known_items_get(self._h, user_row, _ptr(out, C.c_int32), out.size, C.byref(n)))
        return out[:n.value].copy()

    def known_items_retain(self, keep_users=None, keep_items=None):
        ku = None if keep_users is None else np.ascontiguousarray(keep_users, dtype=np.uint8)
        ki = None if keep_items is None else np.ascontiguousarray(keep_items, dtype=np.uint8)
        _check(lib().oryx_known_items_retain(self._h, _ptr(ku, C.c_uint8), 0 if ku is None else ku.size,
                                             _ptr(ki, C.c_uint8), 0 if ki is None else ki.size))

    def known_counts(self, which: int, n_rows: int) -> np.ndarray:
        out = np.zeros(max(n_rows, 1), dtype=np.int32)
        _check(lib().oryx_known_counts(self._h, which, _ptr(out, C.c_int32), n_rows))
        return out[:n_rows]

    def recommend(self, user_row: int, how_many: int, consider_known_items: bool = False):
        """-> (rows, scores) or None when the user has no vector (the reference's 404)."""
        rows = np.zeros(max(how_many, 1), dtype=np.int32)
        scores = np.zeros(max(how_many, 1), dtype=np.float64)
        n, found = C.c_int32(), C.c_int32()
        _check(lib().oryx_recommend(self._h, user_row, how_many, int(consider_known_items), _ptr(rows, C.c_int32),
                                    _ptr(scores, C.c_double), C.byref(n), C.byref(found)))
        if not found.value:
            return None
        return rows[:n.value].copy(), scores[:n.value].copy()

    def topn_coalesced(self, kind: int, query, how_many: int, excluded=None):
        """topn() for many caller threads: concurrent calls are answered together by one batch pass."""
        q = _f32(query)
        ex = None if excluded is None or len(excluded) == 0 else np.ascontiguousarray(excluded, dtype=np.int32)
        rows = np.zeros(max(how_many, 1), dtype=np.int32)
        scores = np.zeros(max(how_many, 1), dtype=np.float64)
        n = C.c_int32()
        _check(lib().oryx_topn_coalesced(self._h, kind, _ptr(q, C.c_float), how_many, _ptr(ex, C.c_int32),
                                         0 if ex is None else ex.size, _ptr(rows, C.c_int32),
                                         _ptr(scores, C.c_double), C.byref(n)))
        return rows[:n.value].copy(), scores[:n.value].copy()

    def recommend_coalesced(self, user_row: int, how_many: int, consider_known_items: bool = False):
        """recommend() for many caller threads (answered together by one batch pass)."""
        rows = np.zeros(max(how_many, 1), dtype=np.int32)
        scores = np.zeros(max(how_many, 1), dtype=np.float64)
        n, found = C.c_int32(), C.c_int32()
        _check(lib().oryx_recommend_coalesced(self._h, user_row, how_many, int(consider_known_items),
                                              _ptr(rows, C.c_int32), _ptr(scores, C.c_double), C.byref(n),
                                              C.byref(found)))
        if not found.value:
            return None
        return rows[:n.value].copy(), scores[:n.value].copy()

    def recommend_batch(self, user_rows, how_many: int, consider_known_items: bool = False):
        """-> (rows [n x how_many], scores [n x how_many], counts [n]); counts[i] == -1: no such user."""
        users = np.ascontiguousarray(user_rows, dtype=np.int32)
        n = users.size
        rows = np.zeros((max(n, 1), how_many), dtype=np.int32)
        scores = np.zeros((max(n, 1), how_many), dtype=np.float64)
        cnt = np.zeros(max(n, 1), dtype=np.int32)
        _check(lib().oryx_recommend_batch(self._h, _ptr(users, C.c_int32), n, how_many, int(consider_known_items),
                                          _ptr(rows, C.c_int32), _ptr(scores, C.c_double), _ptr(cnt, C.c_int32)))
        return rows[:n], scores[:n], cnt[:n]

    def topn_batch(self, kind: int, queries, how_many: int, excluded_lists=None):
        q = _f32(queries).reshape(-1, self.features)
        nq = q.shape[0]
        offs = rows_ex = None
        if excluded_lists is not None:
            offs = np.zeros(nq + 1, dtype=np.int32)
            offs[1:] = np.cumsum([len(e) for e in excluded_lists])
            rows_ex = np.ascontiguousarray(np.concatenate([np.asarray(e, dtype=np.int32) for e in excluded_lists])
                                           if offs[-1] else np.zeros(1, dtype=np.int32), dtype=np.int32)
        rows = np.zeros((nq, how_many), dtype=np.int32)
        scores = np.zeros((nq, how_many), dtype=np.float64)
        n = np.zeros(nq, dtype=np.int32)
        _check(lib().oryx_topn_batch(self._h, kind, _ptr(q, C.c_float), nq, how_many, _ptr(offs, C.c_int32),
                                     _ptr(rows_ex, C.c_int32), _ptr(rows, C.c_int32), _ptr(scores, C.c_double),
                                     _ptr(n, C.c_int32)))
        return rows, scores, n

    def scores_all(self, kind: int, query, n_rows: int) -> np.ndarray:
        q = _f32(query)
        out = np.zeros(n_rows, dtype=np.float64)
        _check(lib().oryx_scores_all(self._h, kind, _ptr(q, C.c_float), n_rows, _ptr(out, C.c_double)))
        return out

    def scores_rows(self, kind: int, query, rows):
        q = _f32(query)
        r = np.ascontiguousarray(rows, dtype=np.int32)
        out = np.zeros(r.size, dtype=np.float64)
        found = np.zeros(r.size, dtype=np.uint8)
        _check(lib().oryx_scores_rows(self._h, kind, _ptr(q, C.c_float), _ptr(r, C.c_int32), r.size,
                                      _ptr(out, C.c_double), _ptr(found, C.c_uint8)))
        return out, found.astype(bool)

    def estimate_pairs(self, user_rows, item_rows, wide_products: bool = False):
        """dot(X[user], Y[item]) for many pairs -> (scores float64[n], found uint8[n])."""
        u = np.ascontiguousarray(user_rows, dtype=np.int32)
        it = np.ascontiguousarray(item_rows, dtype=np.int32)
        out = np.zeros(max(u.size, 1), dtype=np.float64)
        found = np.zeros(max(u.size, 1), dtype=np.uint8)
        _check(lib().oryx_estimate_pairs(self._h, _ptr(u, C.c_int32), _ptr(it, C.c_int32), u.size, int(wide_products),
                                         _ptr(out, C.c_double), _ptr(found, C.c_uint8)))
        return out[:u.size], found[:u.size]

    def topn_device(self, kind: int, d_query: int, how_many: int, d_rows: int, d_scores: int, d_n: int, stream: int):
        _check(lib().oryx_topn_device(self._h, kind, C.c_void_p(d_query), how_many, C.c_void_p(d_rows),
                                      C.c_void_p(d_scores), C.c_void_p(d_n), C.c_void_p(stream)))

    def topn_device_static(self, kind: int, d_query: int, how_many: int, d_rows: int, d_scores: int, d_n: int,
                           stream: int):
        """oryx_topn_device_static: same query as topn_device, scratch in the module's fixed array, no
        allocation.  how_many <= STATIC_MAX_HOW_MANY; calls that overlap on one GPU race."""
        _check(lib().oryx_topn_device_static(self._h, kind, C.c_void_p(d_query), how_many, C.c_void_p(d_rows),
                                             C.c_void_p(d_scores), C.c_void_p(d_n), C.c_void_p(stream)))

    def gram(self, which: int = Y):
        out = np.zeros((self.features, self.features), dtype=np.float64)
        used = C.c_int64()
        _check(lib().oryx_gram(self._h, which, _ptr(out, C.c_double), C.byref(used)))
        return (out if used.value > 0 else None), used.value

    def aggregate_interactions(self, user_rows, item_rows, strengths, timestamps, implicit: bool):
        """ALSSpeedModelManager.buildUpdates' aggregation step on the GPU -> (user_rows, item_rows, strengths)."""
        u = np.ascontiguousarray(user_rows, dtype=np.int32)
        it = np.ascontiguousarray(item_rows, dtype=np.int32)
        st = np.ascontiguousarray(strengths, dtype=np.float64)
        ts = np.ascontiguousarray(timestamps, dtype=np.int64)
        n = u.size
        ou, oi, os_ = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.float64)
        kept = C.c_int64()
        _check(lib().oryx_aggregate_interactions(self._h, _ptr(u, C.c_int32), _ptr(it, C.c_int32), _ptr(st, C.c_double),
                                                 _ptr(ts, C.c_int64), n, int(implicit), _ptr(ou, C.c_int32),
                                                 _ptr(oi, C.c_int32), _ptr(os_, C.c_double), C.byref(kept)))
        k = kept.value
        return ou[:k].copy(), oi[:k].copy(), os_[:k].copy()

    def debug_batch_timings(self) -> dict:
        """Phase timings (device ms) of the last tensor-core batch answered for this model."""
        t = np.zeros(9, dtype=np.float64)
        _check(lib().oryx_debug_batch_timings(self._h, _ptr(t, C.c_double)))
        keys = ("convert_ms", "score_ms", "threshold_ms", "refine_ms", "padded_flops", "cand_mean", "cand_max", "runs",
                "shadow_rebuilds")
        return dict(zip(keys, t.tolist()))

    def debug_gram_ms(self) -> float:
        """Device time (ms) of the last gram() on this model."""
        return self.debug_gram()[0]

    def debug_gram(self):
        """(device ms of the last gram(), True when the tensor-core kernel produced it)."""
        t = np.zeros(2, dtype=np.float64)
        _check(lib().oryx_debug_gram_ms(self._h, _ptr(t, C.c_double)))
        return float(t[0]), bool(t[1])

    def stats(self) -> dict:
        s = _Stats()
        _check(lib().oryx_stats_get(self._h, C.byref(s)))
        return {"rows": list(s.rows), "slots": list(s.slots), "device_bytes": s.device_bytes,
                "queries": s.queries, "kernel_launches": s.kernel_launches, "last_query_ms": s.last_query_ms}


class Exchange:
    """One rank's end of the peer-memory exchange of a row-sharded model (oryx_exchange_*)."""

    HANDLE_BYTES = 64

    def __init__(self, device: int, world: int, rank: int, max_how_many: int):
        self._h = C.c_void_p()
        self.world, self.rank = world, rank
        _check(lib().oryx_exchange_create(device, world, rank, max_how_many, C.byref(self._h)))

    def handle(self) -> bytes:
        buf = C.create_string_buffer(self.HANDLE_BYTES)
        _check(lib().oryx_exchange_handle(self._h, buf))
        return buf.raw

    def connect(self, handles):
        """handles: list of `world` 64-byte handles in rank order (from the other processes)."""
        _check(lib().oryx_exchange_connect(self._h, b"".join(handles)))

    @staticmethod
    def connect_local(exchanges):
        arr = (C.c_void_p * len(exchanges))(*[x._h for x in exchanges])
        _check(lib().oryx_exchange_connect_local(arr, len(exchanges)))

    def topn(self, model: "Model", kind: int, d_query: int, how_many: int, row_stride: int, d_rows: int,
             d_scores: int, d_n: int, stream: int):
        _check(lib().oryx_topn_sharded_device(model._h, self._h, kind, d_query, how_many, row_stride, d_rows,
                                              d_scores, d_n, stream))

    def gram_allreduce_device(self, d_gram: int, features: int, d_status: int, stream: int):
        """Sum of the ranks' features x features float64 device matrices over peer memory, in place
        (oryx_gram_allreduce_device).  d_status: device int32, 0 = ok."""
        _check(lib().oryx_gram_allreduce_device(self._h, C.c_void_p(d_gram), features, C.c_void_p(d_status),
                                                C.c_void_p(stream)))

    def gram_sharded(self, model: "Model", which: int = Y):
        """The global Gram of a row-sharded store: this rank's oryx_gram, then the peer-memory all-reduce."""
        out = np.zeros((model.features, model.features), dtype=np.float64)
        used = C.c_int64()
        _check(lib().oryx_gram_sharded(model._h, self._h, which, _ptr(out, C.c_double), C.byref(used)))
        return out, used.value

    def close(self):
        if self._h:
            lib().oryx_exchange_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PeerStores:
    """This rank's view of every rank's X and Y stores when the ranks are separate processes
    (oryx_peer_stores_*): exchange `handle()`s (e.g. with torch.distributed.all_gather), `connect`, then
    `Solver.fold_in_rows_peers` gathers rows of any rank by global row id over NVLink."""

    HANDLE_BYTES = 512

    def __init__(self, model: "Model", world: int, rank: int):
        self._h = C.c_void_p()
        self.world, self.rank = world, rank
        _check(lib().oryx_peer_stores_create(model._h, world, rank, C.byref(self._h)))

    def handle(self) -> bytes:
        buf = C.create_string_buffer(self.HANDLE_BYTES)
        _check(lib().oryx_peer_stores_handle(self._h, buf))
        return buf.raw

    def connect(self, handles):
        """handles: `world` handles of HANDLE_BYTES in rank order (the own entry is ignored)."""
        _check(lib().oryx_peer_stores_connect(self._h, b"".join(handles)))

    def close(self):
        if self._h:
            lib().oryx_peer_stores_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def merge_topn_device(d_scores: int, d_rows: int, d_n: int, n_lists: int, how_many: int, row_stride: int,
                      d_out_rows: int, d_out_scores: int, d_out_n: int, stream: int):
    _check(lib().oryx_merge_topn_device(C.c_void_p(d_scores), C.c_void_p(d_rows), C.c_void_p(d_n), n_lists,
                                        how_many, row_stride, C.c_void_p(d_out_rows), C.c_void_p(d_out_scores),
                                        C.c_void_p(d_out_n), C.c_void_p(stream)))


def merge_topn_batch_device(d_scores: int, d_rows: int, d_n: int, n_lists: int, nq: int, how_many: int, row_stride: int,
                            d_out_rows: int, d_out_scores: int, d_out_n: int, stream: int):
    """Merge of every shard's answers to a page of queries, one launch (oryx_merge_topn_batch_device)."""
    _check(lib().oryx_merge_topn_batch_device(C.c_void_p(d_scores), C.c_void_p(d_rows), C.c_void_p(d_n), n_lists, nq, how_many,
                                              row_stride, C.c_void_p(d_out_rows), C.c_void_p(d_out_scores),
                                              C.c_void_p(d_out_n), C.c_void_p(stream)))


def packed_record_bytes(how_many: int) -> int:
    """Size of one packed per-shard result record (scores f64 | rows i32 | count i32), 8-byte aligned."""
    return (how_many * 12 + 4 + 7) // 8 * 8


def merge_topn_packed(d_records: int, n_lists: int, record_bytes: int, how_many: int, row_stride: int,
                      d_out_rows: int, d_out_scores: int, d_out_n: int, stream: int):
    _check(lib().oryx_merge_topn_packed(C.c_void_p(d_records), n_lists, record_bytes, how_many, row_stride,
                                        C.c_void_p(d_out_rows), C.c_void_p(d_out_scores), C.c_void_p(d_out_n),
                                        C.c_void_p(stream)))


class Solver:
    """Handle over oryx_solver (LinearSystemSolver.getSolver -> Solver)."""

    def __init__(self, gram, device: int = 0):
        g = np.ascontiguousarray(gram, dtype=np.float64)
        assert g.ndim == 2 and g.shape[0] == g.shape[1]
        self.n = g.shape[0]
        self._h = C.c_void_p()
        rank = C.c_int32()
        st = lib().oryx_solver_create(_ptr(g, C.c_double), self.n, device, C.byref(self._h), C.byref(rank))
        if st == E_SINGULAR:
            raise SingularMatrixSolverException(rank.value, _last_error())
        _check(st)

    def close(self):
        if getattr(self, "_h", None):
            lib().oryx_solver_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def solve_f2f(self, b) -> np.ndarray:
        b = _f32(b)
        x = np.zeros(self.n, dtype=np.float32)
        _check(lib().oryx_solve_f2f(self._h, _ptr(b, C.c_float), _ptr(x, C.c_float)))
        return x

    def fold_in_batch(self, implicit: bool, xu, has_xu, yi, values):
        yi = _f32(yi).reshape(-1, self.n)
        n = yi.shape[0]
        xu_a = None if xu is None else _f32(xu).reshape(n, self.n)
        has = None if has_xu is None else np.ascontiguousarray(has_xu, dtype=np.uint8)
        vals = np.ascontiguousarray(values, dtype=np.float64)
        out = np.zeros((n, self.n), dtype=np.float32)
        valid = np.zeros(n, dtype=np.uint8)
        _check(lib().oryx_fold_in_batch(self._h, int(implicit), _ptr(xu_a, C.c_float), _ptr(has, C.c_uint8),
                                        _ptr(yi, C.c_float), _ptr(vals, C.c_double), n, _ptr(out, C.c_float),
                                        _ptr(valid, C.c_uint8)))
        return out, valid

    def debug_fold_in_timings(self) -> dict:
        """Of this solver's last fold-in call: device ms of fold_in_kernel (all chunks), chunks, n, wall ms."""
        t = np.zeros(5, dtype=np.float64)
        _check(lib().oryx_debug_fold_in_timings(self._h, _ptr(t, C.c_double)))
        return {"kernel_ms": float(t[0]), "chunks": int(t[1]), "n": int(t[2]), "wall_ms": float(t[3]),
                "gemm_ctas_per_sm": int(t[4])}

    def fold_in_rows_sharded(self, models, xu_row_stride: int, yi_row_stride: int, implicit: bool, xu_which: int,
                             xu_rows, yi_rows, values):
        """fold_in_rows over stores sharded by rows across the GPUs of this process (global int64 rows)."""
        xr = np.ascontiguousarray(xu_rows, dtype=np.int64)
        yr = np.ascontiguousarray(yi_rows, dtype=np.int64)
        vals = np.ascontiguousarray(values, dtype=np.float64)
        n = xr.size
        out = np.empty((n, self.n), dtype=np.float32)
        valid = np.zeros(n, dtype=np.uint8)
        arr = (C.c_void_p * len(models))(*[m._h for m in models])
        _check(lib().oryx_fold_in_rows_sharded(self._h, arr, len(models), xu_row_stride, yi_row_stride, int(implicit),
                                               xu_which, _ptr(xr, C.c_int64), _ptr(yr, C.c_int64),
                                               _ptr(vals, C.c_double), n, _ptr(out, C.c_float), _ptr(valid, C.c_uint8)))
        return out, valid

    def fold_in_rows_peers(self, peers: "PeerStores", xu_row_stride: int, yi_row_stride: int, implicit: bool,
                           xu_which: int, xu_rows, yi_rows, values):
        """fold_in_rows over stores sharded by rows across PROCESSES (global int64 rows; oryx_fold_in_rows_peers)."""
        xr = np.ascontiguousarray(xu_rows, dtype=np.int64)
        yr = np.ascontiguousarray(yi_rows, dtype=np.int64)
        vals = np.ascontiguousarray(values, dtype=np.float64)
        n = xr.size
        out = np.empty((n, self.n), dtype=np.float32)
        valid = np.zeros(n, dtype=np.uint8)
        _check(lib().oryx_fold_in_rows_peers(self._h, peers._h, xu_row_stride, yi_row_stride, int(implicit), xu_which,
                                             _ptr(xr, C.c_int64), _ptr(yr, C.c_int64), _ptr(vals, C.c_double), n,
                                             _ptr(out, C.c_float), _ptr(valid, C.c_uint8)))
        return out, valid

    def fold_in_rows(self, model: "Model", implicit: bool, xu_which: int, xu_rows, yi_rows, values, out=None):
        """computeUpdatedXu for n interactions whose vectors are rows of the model's device stores.
        `out`: optional C-contiguous float32 [n x features] array to receive the rows; when it is page-locked
        (e.g. the numpy view of a torch pinned tensor) the library copies into it directly."""
        xr = np.ascontiguousarray(xu_rows, dtype=np.int32)
        yr = np.ascontiguousarray(yi_rows, dtype=np.int32)
        vals = np.ascontiguousarray(values, dtype=np.float64)
        n = xr.size
        if out is None:
            out = np.empty((n, self.n), dtype=np.float32)
        assert out.dtype == np.float32 and out.shape == (n, self.n) and out.flags["C_CONTIGUOUS"]
        valid = np.zeros(n, dtype=np.uint8)
        _check(lib().oryx_fold_in_rows(self._h, model._h, int(implicit), xu_which, _ptr(xr, C.c_int32),
                                       _ptr(yr, C.c_int32), _ptr(vals, C.c_double), n, _ptr(out, C.c_float),
                                       _ptr(valid, C.c_uint8)))
        return out, valid
