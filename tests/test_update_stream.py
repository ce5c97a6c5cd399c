"""Update-stream text formats: the native parser/formatter in
oryx_b200/host/update_stream.cpp against the reference's own unit-test vectors (replayed in C++,
tests/cpp/test_update_text.cpp) and, through the C ABI of include/oryx_ingest.h, against the CPU
oracle oracle/update_text.py on seeded random messages.  CPU only."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

from oracle import update_text as oracle
from oryx_b200 import ingest, native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "oryx_b200", "host")


@pytest.fixture(scope="module", autouse=True)
def _built():
    ingest.build()


def test_ingest_library_exports_every_declared_symbol():
    text = open(os.path.join(ROOT, "include", "oryx_ingest.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    declared = sorted(set(re.findall(r"\b(?:int32_t|int64_t|void)\s+(oryx_[a-z0-9_]+)\s*\(", text)))
    assert len(declared) >= 15
    L = ingest.lib()
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/oryx_ingest.h but not exported"
    assert sorted(list(ingest.SIGNATURES) + list(ingest._OTHER)) == declared


def test_reference_text_unit_tests_replayed_in_cpp():
    exe = os.path.join(ROOT, "tests", "cpp", "test_update_text")
    src = exe + ".cpp"
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-I" + HOST, "-I" + os.path.join(ROOT, "include"), src, "-o", exe,
                           "-L" + os.path.join(ROOT, "oryx_b200"), "-loryx_host", "-loryx_b200",
                           "-Wl,-rpath," + os.path.join(ROOT, "oryx_b200"), "-pthread"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]


def _random_float32(rng, n):
    kind = rng.integers(0, 10, n)
    v = rng.standard_normal(n).astype(np.float32)
    v = np.where(kind == 0, (v * np.float32(1e-6)).astype(np.float32), v)
    v = np.where(kind == 1, (v * np.float32(1e9)).astype(np.float32), v)
    bits = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32).view(np.float32)
    v = np.where((kind == 2) & np.isfinite(bits), bits, v)  # any finite bit pattern, denormals included
    v = np.where(kind == 3, np.float32(0.0), v)
    return v.astype(np.float32)


_ALPHABET = ["a", "B", "7", "_", "-", " ", '"', "\\", "/", chr(10), chr(9), chr(0xE9), chr(0x6F22), chr(0x1F600), "", ","]


def _random_id(rng):
    n = int(rng.integers(1, 9))
    return "".join(_ALPHABET[int(i)] for i in rng.integers(0, len(_ALPHABET), n))


def _same_floats(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    if a.shape != b.shape:
        return False
    return bool(((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all())


def test_float_to_string_matches_oracle_and_round_trips():
    rng = np.random.default_rng(0xB2000010)
    for v in _random_float32(rng, 4000):
        s = ingest.float_to_string(float(v))
        assert s == oracle.java_number_to_string(v, single=True), (v, s)
        assert np.float32(s) == v
    for d in rng.standard_normal(2000) * 10.0 ** rng.integers(-30, 30, 2000):
        s = ingest.double_to_string(float(d))
        assert s == oracle.java_number_to_string(d, single=False)
        assert float(s) == d


@pytest.mark.parametrize("tagged", [True, False])
def test_writer_output_parses_back_bit_exactly(tagged):
    """toUpdateJSON / part-file lines -> parser: ids, float bits and known ids survive the trip, and
    the text equals the oracle's restatement of Jackson's output."""
    rng = np.random.default_rng(0xB2000011 + tagged)
    msgs, want = [], []
    for _ in range(600):
        f = int(rng.integers(0, 40))
        vec = _random_float32(rng, f)
        if rng.random() < 0.05 and f:
            vec[int(rng.integers(0, f))] = [np.nan, np.inf, -np.inf][int(rng.integers(0, 3))]
        ident = _random_id(rng)
        if tagged:
            matrix = "XY"[int(rng.integers(0, 2))]
            other = _random_id(rng) if rng.random() < 0.7 else None
            msg = ingest.format_update(matrix, ident, vec, other)
            assert msg == oracle.to_update_json(matrix, ident, vec, other)
            want.append((matrix, ident, vec, [other] if other is not None else None))
        else:
            msg = ingest.format_update(None, ident, vec)
            assert msg == oracle.to_feature_line(ident, vec)
            want.append((None, ident, vec, None))
        json.loads(msg)  # is JSON
        msgs.append(msg)
    for threads in (1, 4):
        n, err, got = ingest.parse_feature_messages(msgs, tagged, threads)
        assert err is None and n == len(msgs)
        for (m0, i0, v0, o0), (m1, i1, v1, o1) in zip(want, got):
            assert (m0, i0, o0) == (m1, i1, o1)
            assert _same_floats(v0, v1)


def _random_number_text(rng):
    k = int(rng.integers(0, 8))
    if k == 0:
        return str(int(rng.integers(-10**6, 10**6)))
    if k == 1:  # long decimal expansions (what a double-printing batch layer writes)
        return repr(float(rng.standard_normal() * 10.0 ** int(rng.integers(-8, 8))))
    if k == 2:  # more digits than a double holds
        return "%.30f" % float(rng.standard_normal())
    if k == 3:
        return "%de%d" % (int(rng.integers(-999, 999)), int(rng.integers(-60, 60)))
    if k == 4:  # halfway cases between adjacent floats: the double -> float narrowing must round once
        f = np.float32(rng.standard_normal())
        up = np.nextafter(f, np.float32(np.inf))
        return repr((float(f) + float(up)) / 2.0)
    if k == 5:
        return ["1e400", "-1e400", "1e-400", "0", "-0", "0.0", "-0.0e5", "1E+2",
                "123456789012345678901234567890"][int(rng.integers(0, 9))]
    if k == 6:
        return ['"NaN"', '"Infinity"', '"-INF"', '" 1.5 "', "null", '"2.5e3"'][int(rng.integers(0, 6))]
    return repr(float(np.float32(rng.standard_normal())))


def test_parser_matches_oracle_on_hostile_input():
    rng = np.random.default_rng(0xB2000012)
    msgs = []
    for _ in range(3000):
        nums = ",".join(_random_number_text(rng) for _ in range(int(rng.integers(0, 12))))
        ident = json.dumps(_random_id(rng), ensure_ascii=bool(rng.integers(0, 2)))
        if rng.random() < 0.2:
            ident = str(int(rng.integers(-5, 10**12)))  # numeric ids (ALSUpdate.java:421-424)
        if rng.random() < 0.05:
            ident = ["1.5", "1e3", "true", "-0", "1.0E10"][int(rng.integers(0, 5))]
        others = ""
        if rng.random() < 0.6:
            others = ",[" + ",".join(json.dumps(_random_id(rng)) if rng.random() < 0.9 else str(int(rng.integers(0, 99)))
                                     for _ in range(int(rng.integers(0, 5)))) + "]"
            if rng.random() < 0.1:
                others += ',{"ignored":[1,2,{"a":null}]}'
        ws = [" ", "", chr(9), chr(10) + " "][int(rng.integers(0, 4))]
        tag = "XY"[int(rng.integers(0, 2))]
        msgs.append(f'{ws}[{ws}"{tag}"{ws},{ws}{ident},{ws}[{nums}]{ws}{others}{ws}]{ws}')
    n, err, got = ingest.parse_feature_messages(msgs, True, 3)
    assert err is None and n == len(msgs), err
    for msg, (m1, i1, v1, o1) in zip(msgs, got):
        m0, i0, v0, o0 = oracle.parse_feature_message(msg, True)
        assert (m0, i0, o0) == (m1, i1, o1), msg
        assert _same_floats(v0, v1), (msg, v0, v1)


def test_malformed_messages_stop_the_stream_where_the_reference_throws():
    good = '["Y","I1",[1.0,2.0]]'
    for bad in ['["Z","I1",[1.0]]', '["X","I1"]', '["X","I1",[1.0,]]', '["X","I1",[1.0],"notalist"]', '{"a":1}',
                '["X",null,[1.0]]', '["X","I1",[true]]', '["X","I1",[1.0]', '["X","I1",[NaN]]', '["X","I1",[.5]]',
                '["X","I' + chr(1) + '",[1.0]]', '["X","I1",[1.0],[null]]']:
        with pytest.raises(Exception):
            oracle.parse_feature_message(bad, True)
        n, err, got = ingest.parse_feature_messages([good, good, bad, good], True, 1)
        assert n == 2 and err is not None and "message 2" in err, (bad, n, err)
        assert [g[1] for g in got] == ["I1", "I1"]


def test_empty_and_single():
    n, err, got = ingest.parse_feature_messages([], True)
    assert n == 0 and err is None and got == []
    n, err, got = ingest.parse_feature_messages(['["X","u",[]]'], True)
    assert n == 1 and got[0][0] == "X" and got[0][2].size == 0 and got[0][3] is None


def test_managers_fail_loudly_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    pmml = ('<PMML version="4.3"><Header/><Extension name="features" value="2"/><Extension name="implicit" '
            'value="true"/><Extension name="XIDs">a</Extension><Extension name="YIDs">b</Extension></PMML>')
    mgr = ingest.ServingModelManager()
    mgr.consume([("UP", '["Y","I1",[1.0,2.0]]')])  # no model yet: skipped (ALSServingModelManager.java:73-75)
    assert mgr.model_info()["handle"] is None
    with pytest.raises(native.OryxError) as ei:
        mgr.consume([("MODEL", pmml)])  # creating the device model needs a GPU: no CPU fallback
    assert "no CPU fallback" in str(ei.value)


def test_oracle_sum_with_nan_reference_vectors():
    """MLFunctionsTest.testSumWithNaN (app/oryx-app-common/src/test/.../fn/MLFunctionsTest.java:60-67)."""
    import math
    from oracle import update_text as ut
    nan = float("nan")
    assert ut.sum_with_nan([1.0]) == 1.0
    assert ut.sum_with_nan([1.0, 2.0, 3.0]) == 6.0
    assert ut.sum_with_nan([1.0, nan, 3.0]) == 3.0
    assert math.isnan(ut.sum_with_nan([1.0, 2.0, nan]))
    assert math.isnan(ut.sum_with_nan([nan]))
    agg = ut.aggregate_interactions([1, 1, 2, 1], [5, 5, 5, 5], [1.0, 2.5, nan, nan], [30, 10, 5, 20], True)
    assert agg == {(1, 5): 1.0}          # by time: 2.5, NaN, 1.0 -> the delete resets the tally; (2,5) is dropped
    agg = ut.aggregate_interactions([1, 1, 2, 1], [5, 5, 5, 5], [1.0, 2.5, nan, 4.0], [30, 10, 5, 20], False)
    assert agg == {(1, 5): 1.0}          # explicit: last by time wins
