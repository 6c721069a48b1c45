"""CPU check of the ICP correspondence certificate (pcnn_icp.cu: cert_radius / cert_holds).

A search at position c leaves the nearest target point x1 and a floor under the float32 keys of all other
points; the certificate stores c and r = cert_radius(floor).  After the source point has moved to p,
cert_holds(p, (c, r), K(p, x1)) claims that x1 is STILL what the float32 1-NN search of the reference
(nanoflann's key, Homework9/hw9/include/nanoflann.hpp:383-408: dx = q - p, accumulated x -> y -> z, float32)
would return -- strictly smaller key than every other point, so no tie question arises -- and the search is
skipped.  Here the two device functions are restated with their directed roundings emulated exactly
(fractions), and the claim is checked by brute force over random and adversarial configurations, with the
tightest floor a search can produce (the true second-smallest key).  Nothing here touches the GPU."""
from fractions import Fraction

import numpy as np

F32 = np.float32


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


def _sqrt_dir(v, up):
    """float32 square root of the float32 v >= 0, rounded down / up, decided by exact comparison of squares"""
    x = Fraction(float(v))
    f = F32(np.sqrt(np.float64(v)))
    if up:
        while Fraction(float(f)) ** 2 < x:
            f = np.nextafter(f, F32(np.inf))
        while f > 0 and Fraction(float(np.nextafter(f, F32(-np.inf)))) ** 2 >= x:
            f = np.nextafter(f, F32(-np.inf))
    else:
        while Fraction(float(f)) ** 2 > x:
            f = np.nextafter(f, F32(-np.inf))
        while Fraction(float(np.nextafter(f, F32(np.inf)))) ** 2 <= x:
            f = np.nextafter(f, F32(np.inf))
    return f


def cert_radius(floor_key):  # pcnn_icp.cu: __fmul_rd(__fsqrt_rd(floor_key), 0.99999f)
    return _down(Fraction(float(_sqrt_dir(floor_key, False))) * Fraction(float(F32(0.99999))))


def cert_holds(p, c, r, k1):  # pcnn_icp.cu: cert_holds
    if not r > 0:
        return False
    d = [abs(F32(p[i]) - F32(c[i])) for i in range(3)]
    sq = [_up(Fraction(float(v)) ** 2) for v in d]
    m2 = _up(Fraction(float(_up(Fraction(float(sq[0])) + Fraction(float(sq[1]))))) + Fraction(float(sq[2])))
    s = _up(Fraction(float(_up(Fraction(float(_sqrt_dir(k1, True))) + Fraction(float(_sqrt_dir(m2, True)))))) *
            Fraction(float(F32(1.00001))))
    return s < r


def key32(x, q):  # nanoflann's float32 key, no fused multiply-add
    d = (q.astype(F32) - x.astype(F32)).astype(F32)
    return F32(F32(F32(d[..., 0] * d[..., 0]) + F32(d[..., 1] * d[..., 1])) + F32(d[..., 2] * d[..., 2]))


def _scenario(rng, scale):
    n = 48
    X = (rng.normal(0, 1, (n, 3)) * scale).astype(F32)
    c = (X[rng.integers(0, n)] + rng.normal(0, 0.3, 3) * scale).astype(F32)
    if rng.random() < 0.3:  # an almost equidistant runner-up
        j = int(rng.integers(0, n))
        X[(j + 1) % n] = (2 * c - X[j] + rng.normal(0, 1e-4, 3) * scale).astype(F32)
    keys = key32(X, c)
    order = np.lexsort((np.arange(n), keys))
    i1 = int(order[0])
    floor_key = keys[order[1]]
    return X, c, i1, floor_key


def test_certificate_is_sound():
    rng = np.random.default_rng(9)
    held = refused = 0
    for it in range(800):
        scale = F32(10.0 ** rng.integers(-2, 3))
        X, c, i1, floor_key = _scenario(rng, scale)
        r = cert_radius(floor_key)
        k1c = key32(X[i1], c)
        room = max(float(np.sqrt(np.float64(floor_key)) - np.sqrt(np.float64(k1c))), 0.0)
        # motions from far inside the certified band to beyond it; the direction that hurts most is away
        # from x1 towards the runner-up, random directions probe the rest
        frac = [0.05, 0.3, 0.45, 0.499, 0.51, 0.8, 1.1, 1.6][it % 8]
        u = rng.normal(0, 1, 3)
        if it % 2 and np.any(c != X[i1]):
            u = c.astype(np.float64) - X[i1].astype(np.float64)
        u /= np.linalg.norm(u)
        p = (c.astype(np.float64) + u * room * frac).astype(F32)
        k1 = key32(X[i1], p)
        if cert_holds(p, c, r, k1):
            held += 1
            kp = key32(X, p)
            others = np.delete(np.arange(len(X)), i1)
            assert np.all(kp[others] > k1), "a certified correspondence is not the strict float32 nearest neighbour"
        else:
            refused += 1
    assert held > 100 and refused > 100  # both outcomes are exercised


def test_certificate_refuses_without_a_radius():
    assert not cert_holds(np.zeros(3, F32), np.zeros(3, F32), F32(-1.0), F32(0.0))
    assert not cert_holds(np.zeros(3, F32), np.zeros(3, F32), F32(0.0), F32(0.0))
