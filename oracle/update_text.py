"""CPU restatement of the text formats either side of the ALS path -- TEST INFRASTRUCTURE ONLY.

Nothing in the product imports this module; tests use it as the checker for
``oryx_b200/host/update_stream.cpp`` (C ABI ``include/oryx_ingest.h``).

Follows, in the reference:
  * "UP" payload reading, ``ALSServingModelManager.java:76-91`` / ``ALSSpeedModelManager.java:87-98``:
    ``TextUtils.readJSON(message, List.class)`` (``TextUtils.java:170-176``), ``update.get(1).toString()``,
    ``TextUtils.convertViaJSON(update.get(2), float[].class)`` (``:186-188``) -- every JSON number
    becomes a Java ``Double`` (correctly rounded) and is then narrowed to ``float``;
  * part-file lines, ``ALSUpdate.java:461-470``;
  * "UP" payload writing, ``ALSSpeedModelManager.java:207-215`` (``TextUtils.joinJSON``, ``:156-162``).
Third-party behaviour restated (absent from /root/reference): Jackson databind 2.x JSON reading and
writing, Java's ``Float.toString`` layout.  Parity status: PINNED for reading by the reference's own
fixtures (``TextUtilsTest.java:131-141``, ``MockALSModelUpdateGenerator.java:67-100``,
``src/site/markdown/docs/endusers.md:103-128``); Float.toString digits are the shortest that
round-trip (what JDK >= 19 prints; older JDKs occasionally print one digit more, which parses back to
the same float).
"""
import json
import math

import numpy as np


def java_number_to_string(v, single):
    """Float.toString (single=True) / Double.toString layout over shortest round-trip digits."""
    v = np.float32(v) if single else np.float64(v)
    if np.isnan(v):
        return "NaN"
    if np.isinf(v):
        return "Infinity" if v > 0 else "-Infinity"
    sign = "-" if np.signbit(v) else ""
    v = abs(v)
    if v == 0:
        return sign + "0.0"
    s = np.format_float_scientific(v, unique=True, trim="-", exp_digits=1)  # d.ddde-x
    mant, exp = s.split("e")
    e = int(exp)
    d = mant.replace(".", "")
    if len(d) == 1:
        # Float/Double.toString javadoc: when one digit suffices, the closest TWO-digit decimal that
        # still rounds to v is used (Float.MIN_VALUE prints as 1.4E-45, not 1.0E-45).
        s2 = np.format_float_scientific(v, precision=1, unique=False, trim="k", exp_digits=1)
        back = np.float32(s2) if single else np.float64(s2)
        if back == v:
            m2, e2 = s2.split("e")
            d, e = m2.replace(".", "").rstrip("0") or "0", int(e2)
    if -3 <= e < 7:
        if e >= 0:
            ip = (d + "0" * (e + 1))[: e + 1]
            fp = d[e + 1:] or "0"
            return sign + ip + "." + fp
        return sign + "0." + "0" * (-e - 1) + d
    return sign + d[0] + "." + (d[1:] or "0") + "E" + str(e)


def _scalar_to_string(x):
    if isinstance(x, str):
        return x
    if isinstance(x, bool):
        return "true" if x else "false"
    if isinstance(x, int):
        return str(x)
    if isinstance(x, float):
        return java_number_to_string(x, single=False)
    raise ValueError("id is not a scalar")


def _to_float(x):
    if x is None:
        return np.float32(0.0)
    if isinstance(x, bool):
        raise ValueError("boolean in float[]")
    if isinstance(x, (int, float)):
        with np.errstate(over="ignore"):
            return np.float32(float(x))  # Number -> Double -> (float)
    if isinstance(x, str):
        t = x.strip()
        special = {"NaN": math.nan, "Infinity": math.inf, "INF": math.inf, "-Infinity": -math.inf, "-INF": -math.inf}
        if t in special:
            return np.float32(special[t])
        return np.float32(t)  # Float.parseFloat: decimal -> float directly
    raise ValueError("not a number")


def _strict_constant(name):
    raise ValueError("non-standard JSON constant " + name)  # Jackson rejects bare NaN / Infinity


def parse_feature_message(message, tagged):
    """-> (matrix or None, id, float32 vector, other ids or None); raises ValueError when malformed."""
    dec = json.JSONDecoder(parse_constant=_strict_constant)
    update, _ = dec.raw_decode(message.lstrip(" \t\r\n"))  # trailing content is not looked at
    if not isinstance(update, list):
        raise ValueError("not a JSON array")
    k = 0
    matrix = None
    if tagged:
        matrix = _scalar_to_string(update[0])
        k = 1
    ident = _scalar_to_string(update[k])
    vec = update[k + 1]
    if not isinstance(vec, list):
        raise ValueError("vector is not an array")
    vector = np.array([_to_float(x) for x in vec], dtype=np.float32)
    if tagged and matrix not in ("X", "Y"):
        raise ValueError("bad matrix tag")
    others = None
    if tagged and len(update) > k + 2:
        if not isinstance(update[k + 2], list):
            raise ValueError("known ids are not an array")
        others = [_scalar_to_string(x) for x in update[k + 2]]
    return matrix, ident, vector, others


def _json_string(s):
    return json.dumps(s, ensure_ascii=False)


def format_floats(vector):
    out = []
    for v in np.asarray(vector, dtype=np.float32):
        t = java_number_to_string(v, single=True)
        out.append('"' + t + '"' if (np.isnan(v) or np.isinf(v)) else t)
    return "[" + ",".join(out) + "]"


def to_update_json(matrix, ident, vector, other):
    """ALSSpeedModelManager.toUpdateJSON; other=None <=> oryx.als.no-known-items."""
    parts = [_json_string(matrix), _json_string(ident), format_floats(vector)]
    if other is not None:
        parts.append("[" + _json_string(other) + "]")
    return "[" + ",".join(parts) + "]"


def to_feature_line(ident, vector):
    return "[" + _json_string(ident) + "," + format_floats(vector) + "]"


def sum_with_nan(ordered_strengths):
    """MLFunctions.SUM_WITH_NAN (app/oryx-app-common/.../common/fn/MLFunctions.java:67-78): a NaN (a delete)
    poisons the tally, the next value starts it afresh."""
    final = float("nan")
    for s in ordered_strengths:
        if final != final:
            final = s
        else:
            final += s
    return final


def aggregate_interactions(users, items, strengths, timestamps, implicit):
    """The aggregation step of ALSSpeedModelManager.buildUpdates (app/oryx-app/.../speed/als/
    ALSSpeedModelManager.java:141-171): stable sort by timestamp, group by (user, item), SUM_WITH_NAN (implicit)
    or last-wins (explicit, :164), drop NaN results (:167-168), strength narrowed to float (:171).
    Returns a dict {(user, item): strength}."""
    import numpy as np
    order = sorted(range(len(users)), key=lambda i: timestamps[i])   # sorted() is stable
    groups = {}
    for i in order:
        groups.setdefault((int(users[i]), int(items[i])), []).append(float(strengths[i]))
    out = {}
    for key, vals in groups.items():
        r = sum_with_nan(vals) if implicit else vals[-1]
        if r == r:
            out[key] = float(np.float32(r))
    return out
