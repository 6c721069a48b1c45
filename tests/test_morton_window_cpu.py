"""CPU check of the Morton-window argument the kNN walks start from (pcnn_search.cuh: walk_start_level;
pcnn_common.cuh: quant21 / spread21 / morton63).

The walk of a query q whose seed bound is w does not start at the root: with r >= sqrt(w) it starts at the
lowest ancestor of q's home group whose code range strictly contains [code(q - r), code(q + r)].  That is
only allowed if EVERY point within the bound has its Morton code inside that window, i.e. if
  (a) the quantisation is monotone in each coordinate (float32 arithmetic included), and
  (b) bit interleaving preserves componentwise order: a <= b componentwise  =>  morton(a) <= morton(b),
  (c) r, q - r, q + r are computed conservatively (directed rounding).
The functions are restated in numpy / fractions and the claim is checked on random clouds with points on and
inside the ball.  No GPU, no library."""
from fractions import Fraction

import numpy as np

F32 = np.float32


def quant21(v, mn, inv_cell):  # pcnn_common.cuh: float32 arithmetic, clamped, truncated
    f = F32(F32(v - mn) * inv_cell)
    f = np.where(f > 0, f, F32(0))
    f = np.minimum(f, F32(2097151.0))
    return f.astype(np.uint64)


def spread21(v):
    v = v & np.uint64(0x1fffff)
    v = (v | (v << np.uint64(32))) & np.uint64(0x1f00000000ffff)
    v = (v | (v << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
    v = (v | (v << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
    v = (v | (v << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
    v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
    return v


def morton63(p, mn, inv_cell):
    p = np.atleast_2d(p).astype(F32)
    a, b, c = (quant21(p[:, i], mn[i], inv_cell) for i in range(3))
    return spread21(a) | (spread21(b) << np.uint64(1)) | (spread21(c) << np.uint64(2))


def _down(x):
    f = F32(float(x))
    while Fraction(float(f)) > x:
        f = np.nextafter(f, F32(-np.inf))
    while True:
        g = np.nextafter(f, F32(np.inf))
        if np.isfinite(g) and Fraction(float(g)) <= x:
            f = g
        else:
            return f


def _up(x):
    return -_down(-x)


def _sqrt_up(v):
    x = Fraction(float(v))
    f = F32(np.sqrt(np.float64(v)))
    while Fraction(float(f)) ** 2 < x:
        f = np.nextafter(f, F32(np.inf))
    return f


def window(q, w, mn, inv_cell):
    """[code(q - r), code(q + r)] as walk_start_level computes it"""
    r = _up(Fraction(float(_sqrt_up(w))) * Fraction(float(F32(1.000001))))
    lo = np.array([_down(Fraction(float(q[i])) - Fraction(float(r))) for i in range(3)], F32)
    hi = np.array([_up(Fraction(float(q[i])) + Fraction(float(r))) for i in range(3)], F32)
    return morton63(lo, mn, inv_cell)[0], morton63(hi, mn, inv_cell)[0]


def test_interleaving_preserves_componentwise_order():
    rng = np.random.default_rng(1)
    a = rng.integers(0, 1 << 21, (20000, 3), dtype=np.uint64)
    d = rng.integers(0, 1 << 10, (20000, 3), dtype=np.uint64) * rng.integers(0, 2, (20000, 3), dtype=np.uint64)
    b = np.minimum(a + d, np.uint64((1 << 21) - 1))
    ca = spread21(a[:, 0]) | (spread21(a[:, 1]) << np.uint64(1)) | (spread21(a[:, 2]) << np.uint64(2))
    cb = spread21(b[:, 0]) | (spread21(b[:, 1]) << np.uint64(1)) | (spread21(b[:, 2]) << np.uint64(2))
    assert np.all(ca <= cb)


def test_quantisation_is_monotone_in_float32():
    rng = np.random.default_rng(2)
    mn, inv = F32(-37.25), F32(2097152.0) / F32(181.7)
    v = np.sort((rng.normal(0, 60, 200000)).astype(F32))
    v = np.concatenate([v, np.nextafter(v, F32(np.inf))])
    v.sort()
    qv = quant21(v, mn, inv)
    assert np.all(np.diff(qv.astype(np.int64)) >= 0)


def test_every_point_within_the_bound_lies_in_the_window():
    rng = np.random.default_rng(3)
    for it in range(60):
        scale = F32(10.0 ** rng.integers(-1, 3))
        pts = (rng.normal(0, 1, (4000, 3)) * scale).astype(F32)
        mn = pts.min(0)
        side = F32((pts.max(0) - mn).max())
        inv = F32(2097152.0) / side
        codes = morton63(pts, mn, inv)
        for _ in range(6):
            q = (pts[rng.integers(0, len(pts))] + rng.normal(0, 0.05, 3) * scale).astype(F32)
            if it % 3 == 0:
                q = (q + rng.normal(0, 2, 3) * scale).astype(F32)  # also queries outside the cloud's cube
            d = pts.astype(np.float64) - q.astype(np.float64)
            key = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
            kth = np.sort(key)[rng.integers(0, 64)]
            w = _up(Fraction(float(kth)))           # the float32 bound the walk prunes with (ru32)
            wlo, whi = window(q, w, mn, inv)
            inside = key <= np.float64(w)
            assert inside.sum() >= 1
            assert np.all(codes[inside] >= wlo) and np.all(codes[inside] <= whi)
