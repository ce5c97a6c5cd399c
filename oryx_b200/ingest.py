"""ctypes binding over include/oryx_ingest.h (liboryx_host.so): the update-stream text formats and
the two model managers.  No compute and no parsing happens in Python."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import native

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboryx_host.so")

_f32p, _i64p, _u8p = C.POINTER(C.c_float), C.POINTER(C.c_int64), C.POINTER(C.c_uint8)
SIGNATURES = {
    "oryx_parse_feature_messages": [C.c_char_p, _i64p, C.c_int64, C.c_int32, C.c_int32, _u8p, _i64p, _f32p, C.c_int64,
                                    _i64p, C.c_char_p, C.c_int64, _u8p, _i64p, _i64p, C.c_int64, C.c_char_p,
                                    C.c_int64, _i64p],
    "oryx_float_to_string": [C.c_float, C.c_char_p],
    "oryx_double_to_string": [C.c_double, C.c_char_p],
    "oryx_serving_manager_create": [C.c_double, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)],
    "oryx_serving_manager_consume": [C.c_void_p, C.c_char_p, _i64p, C.c_char_p, _i64p, C.c_int64],
    "oryx_serving_manager_consume_updates": [C.c_void_p, C.c_char_p, _i64p, C.c_int64],
    "oryx_serving_manager_load_features": [C.c_void_p, C.c_int32, C.c_char_p, _i64p, C.c_int64],
    "oryx_serving_manager_model": [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                   _i64p, _i64p, _f32p],
    "oryx_serving_manager_item_row": [C.c_void_p, C.c_char_p, C.POINTER(C.c_int32)],
    "oryx_speed_manager_create": [C.c_int32, C.c_double, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)],
    "oryx_speed_manager_consume": [C.c_void_p, C.c_char_p, _i64p, C.c_char_p, _i64p, C.c_int64],
    "oryx_speed_manager_build_updates": [C.c_void_p, C.c_char_p, _i64p, C.c_int64, C.c_char_p, C.c_int64, _i64p,
                                         C.c_int64, _i64p, _i64p],
    "oryx_ingest_last_error": [C.c_char_p, C.c_int32],
}
# entry points whose return type is not an int32 status
_OTHER = {
    "oryx_format_update": (C.c_int64, [C.c_int32, C.c_char_p, _f32p, C.c_int32, C.c_char_p, C.c_char_p, C.c_int64]),
    "oryx_serving_manager_destroy": (None, [C.c_void_p]),
    "oryx_speed_manager_destroy": (None, [C.c_void_p]),
}

_lib = None


def build() -> str:
    native.build()
    subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "host")])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise ImportError(f"{_SO} is not built; run oryx_b200.ingest.build()")
        native.lib()  # liboryx_b200.so first (same directory, found through the rpath as well)
        L = C.CDLL(_SO)
        for name, args in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = C.c_int32, args
        for name, (res, args) in _OTHER.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def _last_error() -> str:
    buf = C.create_string_buffer(2048)
    lib().oryx_ingest_last_error(buf, len(buf))
    return buf.value.decode("utf-8", "replace")


def _check(status: int):
    if status != native.OK:
        raise native.OryxError(status, _last_error())


def pack(strings):
    """list of str/bytes -> (bytes blob, int64 offsets[n+1]) -- the header's "message array"."""
    bs = [s.encode("utf-8") if isinstance(s, str) else s for s in strings]
    begin = np.zeros(len(bs) + 1, dtype=np.int64)
    if bs:
        np.cumsum([len(b) for b in bs], out=begin[1:])
    return b"".join(bs), begin


def _p(a, t):
    return a.ctypes.data_as(t)


def parse_feature_messages(messages, tagged: bool, threads: int = 0):
    """-> (n_parsed, error or None, list of (matrix, id, float32 vector, others or None))."""
    blob, begin = pack(messages)
    n = len(messages)
    cap = max(16, len(blob))
    matrix = np.zeros(n + 1, np.uint8)
    has = np.zeros(n + 1, np.uint8)
    vec_begin = np.zeros(n + 1, np.int64)
    id_begin = np.zeros(n + 1, np.int64)
    other_begin = np.zeros(n + 1, np.int64)
    other_id_begin = np.zeros(cap + 1, np.int64)
    values = np.zeros(cap, np.float32)
    ids = C.create_string_buffer(cap)
    other_ids = C.create_string_buffer(cap)
    n_parsed = C.c_int64(0)
    st = lib().oryx_parse_feature_messages(blob, _p(begin, _i64p), n, int(tagged), threads, _p(matrix, _u8p),
                                           _p(vec_begin, _i64p), _p(values, _f32p), cap, _p(id_begin, _i64p), ids, cap,
                                           _p(has, _u8p), _p(other_begin, _i64p), _p(other_id_begin, _i64p), cap,
                                           other_ids, cap, C.byref(n_parsed))
    err = None
    if st != native.OK:
        if st != native.E_INVALID:
            _check(st)
        err = _last_error()
    out = []
    idb, oib = ids.raw, other_ids.raw
    for i in range(n_parsed.value):
        others = None
        if has[i]:
            others = [oib[other_id_begin[k]:other_id_begin[k + 1]].decode("utf-8")
                      for k in range(other_begin[i], other_begin[i + 1])]
        out.append((chr(matrix[i]) if matrix[i] else None, idb[id_begin[i]:id_begin[i + 1]].decode("utf-8"),
                    values[vec_begin[i]:vec_begin[i + 1]].copy(), others))
    return n_parsed.value, err, out


def format_update(matrix, ident: str, vector, other=None) -> str:
    v = np.ascontiguousarray(vector, dtype=np.float32)
    cap = 64 + 8 * len(ident.encode()) + 24 * v.size + (8 * len(other.encode()) if other else 0)
    buf = C.create_string_buffer(cap)
    n = lib().oryx_format_update(ord(matrix) if matrix else 0, ident.encode(), _p(v, _f32p), v.size,
                                 other.encode() if other is not None else None, buf, cap)
    if n < 0:
        raise native.OryxError(native.E_NOMEM, "format buffer too small")
    return buf.raw[:n].decode("utf-8")


def float_to_string(v: float) -> str:
    buf = C.create_string_buffer(32)
    lib().oryx_float_to_string(v, buf)
    return buf.value.decode()


def double_to_string(v: float) -> str:
    buf = C.create_string_buffer(32)
    lib().oryx_double_to_string(v, buf)
    return buf.value.decode()


class ServingModelManager:
    """ALSServingModelManager (serving/als/model/ALSServingModelManager.java) over the C ABI."""

    def __init__(self, sample_rate: float = 1.0, num_cores: int = 0, device: int = 0, parse_threads: int = 0):
        self._h = C.c_void_p()
        _check(lib().oryx_serving_manager_create(sample_rate, num_cores, device, parse_threads, C.byref(self._h)))

    def close(self):
        if self._h:
            lib().oryx_serving_manager_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def consume(self, key_messages):
        kb, kbeg = pack([k for k, _ in key_messages])
        mb, mbeg = pack([m for _, m in key_messages])
        _check(lib().oryx_serving_manager_consume(self._h, kb, _p(kbeg, _i64p), mb, _p(mbeg, _i64p), len(key_messages)))

    def consume_updates(self, blob: bytes, begin: np.ndarray):
        _check(lib().oryx_serving_manager_consume_updates(self._h, blob, _p(begin, _i64p), len(begin) - 1))

    def load_features(self, which: int, blob: bytes, begin: np.ndarray):
        _check(lib().oryx_serving_manager_load_features(self._h, which, blob, _p(begin, _i64p), len(begin) - 1))

    def model_info(self):
        h, f, imp = C.c_void_p(), C.c_int32(), C.c_int32()
        nu, ni, frac = C.c_int64(), C.c_int64(), C.c_float()
        _check(lib().oryx_serving_manager_model(self._h, C.byref(h), C.byref(f), C.byref(imp), C.byref(nu),
                                                C.byref(ni), C.byref(frac)))
        return {"handle": h.value, "features": f.value, "implicit": bool(imp.value), "users": nu.value,
                "items": ni.value, "fraction_loaded": frac.value}

    def item_row(self, item: str) -> int:
        r = C.c_int32(-1)
        _check(lib().oryx_serving_manager_item_row(self._h, item.encode(), C.byref(r)))
        return r.value


class SpeedModelManager:
    """ALSSpeedModelManager (app/speed/als/ALSSpeedModelManager.java) over the C ABI."""

    def __init__(self, no_known_items: bool = False, min_model_load_fraction: float = 0.0, device: int = 0,
                 parse_threads: int = 0):
        self._h = C.c_void_p()
        _check(lib().oryx_speed_manager_create(int(no_known_items), min_model_load_fraction, device, parse_threads,
                                               C.byref(self._h)))

    def close(self):
        if self._h:
            lib().oryx_speed_manager_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def consume(self, key_messages):
        kb, kbeg = pack([k for k, _ in key_messages])
        mb, mbeg = pack([m for _, m in key_messages])
        _check(lib().oryx_speed_manager_consume(self._h, kb, _p(kbeg, _i64p), mb, _p(mbeg, _i64p), len(key_messages)))

    def build_updates(self, lines):
        blob, begin = pack(lines)
        n_out, out_len = C.c_int64(0), C.c_int64(0)
        cap, bcap = 1 << 16, 1 << 12
        for _ in range(2):
            out = C.create_string_buffer(cap)
            out_begin = np.zeros(bcap, np.int64)
            st = lib().oryx_speed_manager_build_updates(self._h, blob, _p(begin, _i64p), len(lines), out, cap,
                                                        _p(out_begin, _i64p), bcap, C.byref(n_out), C.byref(out_len))
            if st == native.E_NOMEM and out_len.value:
                cap, bcap = out_len.value + 16, n_out.value + 2
                continue
            _check(st)
            raw = out.raw
            return [raw[out_begin[i]:out_begin[i + 1]].decode("utf-8") for i in range(n_out.value)]
        raise native.OryxError(native.E_NOMEM, "build_updates output did not fit")
