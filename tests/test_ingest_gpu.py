"""The model managers of include/oryx_ingest.h through their C ABI on a GPU: the update topic of
MockALSModelUpdateGenerator.java:67-100 goes in as text, the serving model answers through the
oryx_model handle the manager hands back, and the speed manager's buildUpdates reproduces the fold-in
goldens of ALSSpeedIT.java:80-113."""
import ctypes as C
import json

import numpy as np
import pytest

import oracle
import oryx_b200 as ob
from als_harness import load_golden
from oracle import update_text
from oryx_b200 import ingest, native

pytestmark = pytest.mark.gpu


def _string_id(i: int) -> str:  # ALSUtilsTest.java:50-52
    return chr(ord("A") + i % 26) + str(i)


def _topic():
    s = load_golden("speed")
    x_ids, y_ids = list(s["X"]), list(s["Y"])
    pmml = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>\n<PMML version="4.3" '
            'xmlns="http://www.dmg.org/PMML-4_3">\n <Header><Application name="Oryx"/></Header>\n'
            ' <Extension name="features" value="2"/>\n <Extension name="implicit" value="true"/>\n'
            f' <Extension name="XIDs">{" ".join(x_ids)}</Extension>\n <Extension name="YIDs">{" ".join(y_ids)}</Extension>\n</PMML>')
    msgs = [("MODEL", pmml)]
    for k, v in s["Y"].items():
        msgs.append(("UP", update_text.to_update_json("Y", k, v, None)))
    for k, v in s["X"].items():
        msgs.append(("UP", update_text.to_update_json("X", k, v, None)[:-1] + ',["B1","E4"]]'))
    return s, msgs


def test_serving_manager_consumes_the_update_topic():
    s, msgs = _topic()
    ingest.build()
    mgr = ingest.ServingModelManager(sample_rate=1.0, num_cores=4)
    mgr.consume(msgs[3:4])  # an "UP" before any MODEL is skipped
    assert mgr.model_info()["handle"] is None
    mgr.consume(msgs)
    info = mgr.model_info()
    assert info["features"] == 2 and info["implicit"] and info["items"] == len(s["Y"]) and info["users"] == len(s["X"])
    assert info["fraction_loaded"] == 1.0
    # the handle is a plain oryx_model*: query it through the product ABI
    L = native.lib()
    h = C.c_void_p(info["handle"])
    y = np.array([s["Y"][k] for k in s["Y"]], dtype=np.float32)
    rows_of = [mgr.item_row(k) for k in s["Y"]]
    assert sorted(rows_of) == list(range(len(s["Y"]))) and mgr.item_row("nope") == -1
    q = np.array(list(s["X"].values())[0], dtype=np.float32)
    out_rows = np.zeros(3, np.int32); out_scores = np.zeros(3, np.float64); n = C.c_int32()
    assert L.oryx_topn(h, ob.DOT, native._ptr(q, C.c_float), 3, None, 0, native._ptr(out_rows, C.c_int32),
                       native._ptr(out_scores, C.c_double), C.byref(n)) == 0
    by_row = np.zeros_like(y)
    for k, r in zip(s["Y"], rows_of):
        by_row[r] = s["Y"][k]
    want_rows, want_scores = oracle.topn(ob.DOT, q, by_row, 3)
    assert out_rows[:n.value].tolist() == want_rows.tolist() and out_scores[:n.value].tobytes() == want_scores.tobytes()
    # part-file bulk load through the ABI (ALSUpdate.java:415-428 lines)
    lines = [update_text.to_feature_line("bulk%d" % i, np.array([i * 0.5, -1.0 / (i + 1)], np.float32)) for i in range(2000)]
    blob, begin = ingest.pack(lines)
    mgr.load_features(ob.Y, blob, begin)
    assert mgr.model_info()["items"] == len(s["Y"]) + 2000 and mgr.item_row("bulk1999") >= 0
    # malformed traffic surfaces as ORYX_E_INVALID with the reason
    with pytest.raises(native.OryxError) as ei:
        mgr.consume([("UP", '["Y","ok",[1.0,2.0]]'), ("UP", '["Y","bad",[1.0,]]')])
    assert ei.value.status == native.E_INVALID and mgr.item_row("ok") >= 0 and mgr.item_row("bad") == -1
    mgr.close()


def test_speed_manager_build_updates_goldens():
    s, msgs = _topic()
    mgr = ingest.SpeedModelManager(no_known_items=False)
    assert mgr.build_updates(["a,b,1,1"]) == []  # no model yet
    mgr.consume(msgs)
    lines = []
    for i, (user, item, value) in enumerate(s["interactions"]):
        lines.append(f"{user},{item},{value},{1000 + i}")
    ups = mgr.build_updates(lines)
    assert len(ups) == len(s["interactions"])  # one side of every interaction is new: one update each
    seen = 0
    for u in ups:
        matrix, ident, vec, other = update_text.parse_feature_message(u, True)
        expected = s["expected_X"].get(ident) if matrix == "X" else s["expected_Y"].get(ident)
        assert expected is not None and len(other) == 1
        assert np.max(np.abs(vec - np.asarray(expected, np.float32))) <= s["tol"]
        json.loads(u)
        seen += 1
    assert seen == len(ups)
    mgr.close()


@pytest.mark.parametrize("n,n_users,n_items", [(1, 1, 1), (1000, 7, 5), (200_000, 3000, 2000), (70_000, 1, 1)])
def test_aggregate_interactions_on_device_matches_oracle(n, n_users, n_items):
    """oryx_aggregate_interactions (sort + segmented reduce on the GPU) against the restatement of
    ALSSpeedModelManager.buildUpdates' aggregation step: implicit SUM_WITH_NAN with deletes, explicit last-wins,
    equal timestamps, one giant group."""
    import numpy as np
    import oryx_b200 as ob
    from oracle import update_text as ut
    rng = np.random.default_rng(n + n_users)
    users = rng.integers(0, n_users, n).astype(np.int32)
    items = rng.integers(0, n_items, n).astype(np.int32)
    strengths = rng.choice([1.0, 2.0, 0.5, -1.0, 3.25, float("nan")], n, p=[0.3, 0.2, 0.2, 0.1, 0.1, 0.1])
    ts = rng.integers(0, max(2, n // 3), n).astype(np.int64)        # many equal timestamps: input order breaks ties
    with ob.Model(4) as m:
        for implicit in (True, False):
            want = ut.aggregate_interactions(users, items, strengths, ts, implicit)
            gu, gi, gs = m.aggregate_interactions(users, items, strengths, ts, implicit)
            got = {(int(u), int(i)): float(s) for u, i, s in zip(gu, gi, gs)}
            assert len(gu) == len(got)                             # no duplicate groups
            assert got == want                                     # binary64 sums in the same order: exact
            keys = gu.astype(np.int64) * (1 << 32) + gi
            assert np.all(np.diff(keys) > 0)                       # ordered by (user, item)
