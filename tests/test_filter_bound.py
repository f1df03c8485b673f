"""The float32 filter of score_fast.cu: its estimate stays within the stated bound of the float64 value.

CPU emulation (numpy) of `est = fma(B32, rcp_approx(z32), A32)` against `tmp = A + B / z` in float64, over
random constants of the admitted range (A, B >= 0, A + B <= 100.5) and random float32 terms t on both
branches, with the approximate reciprocal pushed to both ends of its documented 1-ulp error.  The kernel
only trusts floor(est) when est is farther than kFilterDelta = 2^-14 from an integer, so the maximum
deviation must stay below that; DESIGN.md 4.1 derives 7u * 100.5 = 4.2e-5."""
import numpy as np

DELTA = 2.0 ** -14
BOUND = 7 * 2.0 ** -24 * 100.5


def _estimate(A, B, t, neg, ulp_sign):
    a = (np.float32(1.0) + np.abs(t)).astype(np.float32)
    with np.errstate(over="ignore"):  # (1+|t|)^2 may exceed float32: z32 = inf, v = 0, as on the GPU
        zn = (a.astype(np.float64) * a.astype(np.float64) + 1.0).astype(np.float32)  # one rounding (fma)
    z32 = np.where(neg, zn, a)
    with np.errstate(over="ignore", divide="ignore"):
        v = (1.0 / z32.astype(np.float64)).astype(np.float32)
    v = (v.astype(np.float64) * (1.0 + ulp_sign * 2.0 ** -23)).astype(np.float32)   # rcp.approx: up to 1 ulp off
    v = np.where(np.abs(v) < np.float32(2.0 ** -126), np.float32(0), v)              # .ftz
    return (B.astype(np.float32).astype(np.float64) * v.astype(np.float64)
            + A.astype(np.float32).astype(np.float64)).astype(np.float32)           # fma: one rounding


def test_estimate_within_bound():
    rng = np.random.default_rng(20261107)
    n = 2_000_000
    worst = 0.0
    for mag in (1e-3, 1.0, 30.0, 1e4, 1e18, 3e37):
        s = rng.uniform(0.0, 100.5, n)
        f = rng.random(n) ** 2
        A, B = s * f, s * (1.0 - f)
        t = (rng.standard_normal(n) * mag).astype(np.float32)
        neg = rng.random(n) < 0.5
        w = np.abs(t.astype(np.float64))
        with np.errstate(over="ignore"):
            z = 1.0 + np.where(neg, (w + 1.0) ** 2, w)
        tmp = A + B / z
        for sign in (-1.0, 1.0):
            est = _estimate(A, B, t, neg, sign).astype(np.float64)
            worst = max(worst, float(np.max(np.abs(est - tmp))))
    assert worst <= BOUND < DELTA, (worst, BOUND, DELTA)
