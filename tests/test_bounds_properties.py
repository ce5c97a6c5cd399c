"""CPU restatements of two arguments the CUDA kernels rely on, checked against the oracle / exact arithmetic on
random and adversarial inputs (no GPU involved; the kernels themselves are checked by the -m gpu tests).

1. common.cuh warp_exact_dot: if every prefix of the integer-scaled terms fl32(a_j b_j) fits a binary64 significand,
   the reference's float-multiply / double-add chain never rounds, and its value is the exact sum.
2. batch_topn.cu: |bf16(q) . bf16(y) - q . y| <= ||bf16(q)|| ||y - bf16(y)|| + ||q - bf16(q)|| ||y||  (the pruning
   bound of the tensor-core batch path, before the accumulation term)."""
from fractions import Fraction

import numpy as np

import oracle


def _terms(a, b):
    return (a.astype(np.float32) * b.astype(np.float32)).astype(np.float32)      # IEEE binary32 multiply


def _fast_path(t):
    """The kernel's test: returns the exact sum as a Fraction when the chain provably never rounds, else None."""
    bits = t.view(np.uint32).astype(np.int64)
    ex = (bits >> 23) & 255
    if np.any(ex == 255):
        return None
    nz = (bits & 0x7FFFFFFF) != 0
    if not nz.any():
        return Fraction(0)
    e = np.where(ex == 0, 1, ex)
    emin, emax = int(e[nz].min()), int(e[nz].max())
    if emax - emin > 62 - 24 - (32 - (32 - int(len(t)).bit_length())):            # 32 - clz(n) = bit_length(n)
        return None
    p = 0
    for j in range(len(t)):
        if nz[j]:
            m = int(bits[j] & 0x7FFFFF) | (0 if ex[j] == 0 else 0x800000)
            v = m << (int(e[j]) - emin)
            p += -v if bits[j] & 0x80000000 else v
        if p != 0:
            mag = abs(p)
            if mag.bit_length() - ((mag & -mag).bit_length() - 1) > 53:
                return None
    return Fraction(p) * Fraction(2) ** (emin - 150)


def test_exact_chain_value_when_no_prefix_rounds():
    rng = np.random.default_rng(0xB2000A00)
    hits = 0
    for trial in range(600):
        n = int(rng.choice([1, 2, 7, 50, 64, 128, 250, 256]))
        a = rng.standard_normal(n).astype(np.float32)
        b = rng.standard_normal(n).astype(np.float32)
        mode = trial % 6
        if mode == 1:                                   # wide dynamic range: most of these chains round
            a *= np.exp2(rng.integers(-30, 31, n)).astype(np.float32)
        elif mode == 2:                                 # short significands: nothing ever rounds
            a, b = np.round(a * 16) / 16, np.round(b * 16) / 16
        elif mode == 3:                                 # exact cancellations and zeros
            a[: n // 2] = -a[n - n // 2:][::-1]
            b[: n // 2] = b[n - n // 2:][::-1]
            a[::5] = 0
        elif mode == 4:                                 # subnormal products
            a *= np.float32(2.0 ** -90)
            b *= np.float32(2.0 ** -50)
        elif mode == 5:                                 # moderate spread around the decision boundary
            a *= np.exp2(rng.integers(-12, 13, n)).astype(np.float32)
        a, b = a.astype(np.float32), b.astype(np.float32)
        with np.errstate(all="ignore"):
            t = _terms(a, b)
        got = _fast_path(t)
        if got is None:
            continue
        hits += 1
        want = oracle.dot(a, b)                         # VectorMath.dot: the chain itself
        assert Fraction(want) == got, (trial, n, want, float(got))
    assert hits > 250                                   # the fast path is the common case


def _bf16(x):
    """Round-to-nearest-even binary32 -> bfloat16 (returned as binary32)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return r.astype(np.uint32).view(np.float32)


def test_bf16_score_error_bound():
    rng = np.random.default_rng(0xB2000A01)
    worst = 0.0
    for trial in range(300):
        f = int(rng.choice([8, 50, 128, 256]))
        q = (rng.standard_normal(f) * np.exp2(rng.integers(-3, 4))).astype(np.float32)
        y = (rng.standard_normal((64, f)) * np.exp2(rng.integers(-3, 4, (64, 1)))).astype(np.float32)
        if trial % 4 == 1:
            y *= np.exp2(rng.integers(-6, 7, y.shape)).astype(np.float32)         # ragged magnitudes inside a row
        qb, yb = _bf16(q), _bf16(y)
        exact = y.astype(np.float64) @ q.astype(np.float64)
        approx = yb.astype(np.float64) @ qb.astype(np.float64)                     # exact products of the rounded operands
        r = np.linalg.norm((y - yb).astype(np.float64), axis=1).max()
        ymax = np.linalg.norm(y.astype(np.float64), axis=1).max()
        bound = np.linalg.norm(qb.astype(np.float64)) * r + np.linalg.norm((q - qb).astype(np.float64)) * ymax
        err = np.abs(approx - exact).max()
        assert err <= bound * (1 + 1e-12), (trial, err, bound)
        # and it is the tighter of the two bounds the kernel could use
        assert bound <= 2.0 ** -7 * 1.01 * np.linalg.norm(q.astype(np.float64)) * ymax
        worst = max(worst, err / bound)
    assert worst < 1.0
