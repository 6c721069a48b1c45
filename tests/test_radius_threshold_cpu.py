"""CPU check of the radius search's float32 pre-tests (pcnn_radius.cu: radius_threshold / make_thr / sq3).

The reference includes a point when `sqrt(key) <= radius` with key = ((dx*dx)+dy*dy)+dz*dz in double
(Homework2/hw2/include/resultSet.hpp:130-140 behind kdtree.hpp / octree.hpp radius searches).  The kernels
decide most candidates in float32: s = fma(dx, dx, fma(dy, dy, dz*dz)) with dx = p - q in float32, and
    s <= t_in  => hit,    s > t_out => miss,    otherwise the double test  key <= thr
where thr is the largest double whose IEEE square root is <= radius.  Here the host functions are restated,
float32 round-to-nearest fused multiply-adds are emulated exactly (fractions), and the three-way decision is
compared with the reference's test on candidates placed on and around the sphere.  No GPU, no library."""
from fractions import Fraction

import numpy as np

F32 = np.float32


def radius_threshold(r):  # pcnn_radius.cu: largest double t with sqrt(t) <= r
    t = np.float64(r) * np.float64(r)
    for _ in range(8):
        if np.sqrt(t) <= r:
            t = np.nextafter(t, np.inf)
    for _ in range(16):
        if np.sqrt(t) > r:
            t = np.nextafter(t, -np.inf)
    return t


def make_thr(radius):  # pcnn_radius.cu: make_thr
    thr = radius_threshold(np.float64(radius))
    lo, hi = thr * (1.0 - 9.5367431640625e-07), thr * (1.0 + 9.5367431640625e-07)
    a = F32(lo)
    if np.float64(a) > lo:
        a = np.nextafter(a, F32(0))
    b = F32(hi)
    if np.float64(b) < hi:
        b = np.nextafter(b, F32(np.inf))
    return thr, a, b


def _rn32(x):
    """exact Fraction -> float32, round to nearest even"""
    f = F32(float(x))  # may be off by one ulp through the double: repair by exact comparison
    best = f
    for g in (np.nextafter(f, F32(-np.inf)), f, np.nextafter(f, F32(np.inf))):
        if abs(Fraction(float(g)) - x) < abs(Fraction(float(best)) - x):
            best = g
        elif abs(Fraction(float(g)) - x) == abs(Fraction(float(best)) - x) and g != best:
            if (np.frombuffer(F32(g).tobytes(), np.uint32)[0] & 1) == 0:
                best = g
    return best


def sq3(ax, ay, az):  # fmaf(ax, ax, fmaf(ay, ay, az * az)), every step rounded once
    ax, ay, az = (Fraction(float(v)) for v in (ax, ay, az))
    t = _rn32(az * az)
    t = _rn32(ay * ay + Fraction(float(t)))
    return _rn32(ax * ax + Fraction(float(t)))


def key64(p, q):
    d = [np.float64(p[i]) - np.float64(q[i]) for i in range(3)]
    return (d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]


def test_threshold_is_the_largest_double_whose_root_fits():
    rng = np.random.default_rng(1)
    for r in np.concatenate([[0.5, 1.0, 0.2, 1.66, 3.0], np.abs(rng.normal(0, 1, 200)) * 10.0 ** rng.integers(-3, 3, 200)]):
        t = radius_threshold(np.float64(r))
        assert np.sqrt(t) <= r and np.sqrt(np.nextafter(t, np.inf)) > r


def test_float_pretests_agree_with_the_reference_test():
    rng = np.random.default_rng(2)
    decided = sliver = 0
    for it in range(1500):
        radius = np.float64([0.5, 1.0, 0.2, 1.66, 0.03, 7.5][it % 6])
        thr, t_in, t_out = make_thr(radius)
        scale = F32(10.0 ** rng.integers(-1, 3))
        q = (rng.normal(0, 1, 3) * scale).astype(F32)
        u = rng.normal(0, 1, 3)
        u /= np.linalg.norm(u)
        # on the sphere to within a few float32 ulps of the coordinates, or clearly inside / outside
        eps = [0.0, 1e-7, -1e-7, 3e-7, -3e-7, 1e-6, -1e-6, 1e-3, -1e-3, 0.3, -0.3][it % 11]
        p = (q.astype(np.float64) + u * radius * (1.0 + eps)).astype(F32)
        s = sq3(p[0] - q[0], p[1] - q[1], p[2] - q[2])
        reference = np.sqrt(key64(p, q)) <= radius          # resultSet.hpp:130-140
        assert (key64(p, q) <= thr) == reference           # the double test the kernels fall back to
        if s <= t_in:
            decided += 1
            assert reference, "float pre-test accepted a point the reference rejects"
        elif s > t_out:
            decided += 1
            assert not reference, "float pre-test rejected a point the reference accepts"
        else:
            sliver += 1
    assert decided > 800 and sliver > 50  # both the float decisions and the sliver are exercised
