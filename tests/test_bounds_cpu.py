"""CPU check of the arithmetic argument behind the kNN kernels' float32 filters (pcnn_common.cuh, lbsq32x;
DESIGN 4.1 "Bounds without a margin"), in exact rational arithmetic:

  s   = fma_rd(az, az, fma_rd(ay, ay, mul_rd(ax, ax)))  with ax = |a - b| rounded toward zero, ...
  key = ((dx*dx) + dy*dy) + dz*dz in double, round to nearest, dx = (double)p.x - (double)q.x, ...
        (the reference's key: Homework2/hw2/include/kdtree.hpp:329-364 via resultSet.hpp)

Claims: (1) s <= the real squared distance; (2) key >= real * (1 - 2^-50); hence (3) whenever s > w for a
float32 w that is >= some other key k2, key > k2 -- a candidate the filter rejects can never belong in the
list, ties included.  (3) is also checked directly on near-equal pairs.  Directed float32 rounding is
emulated with fractions; nothing here touches the GPU or the library."""
from fractions import Fraction

import numpy as np
import pytest


def _f32_down(x):
    """largest float32 <= x (x a Fraction)"""
    f = np.float32(float(x))
    while Fraction(float(f)) > x:
        f = np.nextafter(f, np.float32(-np.inf))
    while True:
        g = np.nextafter(f, np.float32(np.inf))
        if np.isfinite(g) and Fraction(float(g)) <= x:
            f = g
        else:
            return f


def _f32_toward_zero(x):
    return _f32_down(x) if x >= 0 else -_f32_down(-x)


def _f32_up(x):
    return -_f32_down(-x)


def _absdiff_rz(a, b):
    return abs(_f32_toward_zero(Fraction(float(a)) - Fraction(float(b))))


def _lbsq32x(ax, ay, az):
    ax, ay, az = (Fraction(float(v)) for v in (ax, ay, az))
    t = _f32_down(ax * ax)
    t = _f32_down(ay * ay + Fraction(float(t)))
    return _f32_down(az * az + Fraction(float(t)))


def _key64(p, q):
    d = [np.float64(p[i]) - np.float64(q[i]) for i in range(3)]
    return (d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]


def _real(p, q):
    return sum((Fraction(float(p[i])) - Fraction(float(q[i]))) ** 2 for i in range(3))


def _cases(n, seed):
    rng = np.random.default_rng(seed)
    for i in range(n):
        scale = np.float32(10.0 ** rng.integers(-3, 5))
        q = (rng.normal(0, 1, 3) * scale).astype(np.float32)
        kind = i % 4
        if kind == 0:      # a near neighbour
            p = (q + rng.normal(0, 1e-3, 3).astype(np.float32) * scale).astype(np.float32)
        elif kind == 1:    # far, exponents of p and q far apart in one coordinate
            p = (rng.normal(0, 1, 3) * scale).astype(np.float32)
            p[0] = np.float32(p[0] * 1e-9)
        elif kind == 2:    # one ulp away in every coordinate
            p = np.nextafter(q, np.float32(np.inf))
        else:
            p = (rng.normal(0, 1, 3) * scale).astype(np.float32)
        yield p, q


def test_filter_is_a_lower_bound_of_the_real_distance_and_the_key_is_close_to_it():
    for p, q in _cases(1200, 1):
        s = _lbsq32x(_absdiff_rz(p[0], q[0]), _absdiff_rz(p[1], q[1]), _absdiff_rz(p[2], q[2]))
        real = _real(p, q)
        key = Fraction(float(_key64(p, q)))
        assert Fraction(float(s)) <= real
        assert key >= real * (1 - Fraction(1, 2 ** 50))
        assert key <= real * (1 + Fraction(1, 2 ** 50))


def test_a_rejected_candidate_never_belongs_in_the_list():
    # w = ru32(k2) for the key k2 of another point at almost the same distance: if the filter says s > w,
    # the exact key must exceed k2 -- strictly, so that not even a tie with a lower index is lost
    rng = np.random.default_rng(2)
    rejected = 0
    for p, q in _cases(1200, 3):
        p2 = p.copy()
        j = int(rng.integers(0, 3))
        steps = int(rng.integers(-3, 4))
        for _ in range(abs(steps)):
            p2[j] = np.nextafter(p2[j], np.float32(np.inf if steps > 0 else -np.inf))
        k2 = _key64(p2, q)
        w = _f32_up(Fraction(float(k2)))
        s = _lbsq32x(_absdiff_rz(p[0], q[0]), _absdiff_rz(p[1], q[1]), _absdiff_rz(p[2], q[2]))
        if s > w:
            rejected += 1
            assert _key64(p, q) > k2
    assert rejected > 20  # the near-equal pairs do exercise the rejection


def test_ru32_is_monotone_so_float_order_implies_double_order():
    # the k-best lists compare f = ru32(key): f_a < f_b must imply key_a < key_b
    rng = np.random.default_rng(4)
    keys = np.sort(np.abs(rng.normal(0, 1, 4000)) * 10.0 ** rng.integers(-6, 6, 4000))
    keys = np.concatenate([keys, np.nextafter(keys, np.inf), np.nextafter(keys, 0)])
    keys.sort()
    f = np.array([_f32_up(Fraction(float(k))) for k in keys[::7]], np.float32)
    k = keys[::7]
    assert np.all(np.diff(f.astype(np.float64)) >= 0)
    lt = f[:-1] < f[1:]
    assert np.all(k[:-1][lt] < k[1:][lt])
    assert np.all(f.astype(np.float64) >= k)  # rounded UP: also the pruning bound
