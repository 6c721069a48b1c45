"""Synthetic KITTI-like clouds: a 64-beam spinning LiDAR ray-cast against a
ground plane, 40 building boxes and 30 car boxes, with 2 cm range noise.  Host-side numpy,
float32 output, deterministic per seed.  The same buffers feed the CUDA path, the oracle and
the CPU baseline.  Real frames have the on-disk layout of
Homework2/hw2/000000.bin (x, y, z, intensity float32; test.hpp:24-28)."""
import numpy as np

GROUND_Z = -1.73
MAX_RANGE = 80.0
N_BEAMS = 64
ELEV_TOP_DEG, ELEV_BOTTOM_DEG = 2.0, -24.8
RANGE_SIGMA = 0.02
BASE_SEED = 20200607


class World:
    """Ground plane + axis-aligned boxes (buildings and cars), fixed by ``seed``."""

    def __init__(self, seed=BASE_SEED, n_buildings=40, n_cars=30, path_len=200.0):
        rng = np.random.default_rng(seed)
        lo, hi = [], []
        while len(lo) < n_buildings:
            w, d, h = rng.uniform(4, 20), rng.uniform(4, 20), rng.uniform(3, 10)
            cx, cy = rng.uniform(-70, 70 + path_len), rng.uniform(-70, 70)
            if abs(cy) - d / 2 < 3.0:  # keep the driving corridor (the x axis) clear
                continue
            lo.append([cx - w / 2, cy - d / 2, GROUND_Z])
            hi.append([cx + w / 2, cy + d / 2, GROUND_Z + h])
        while len(lo) < n_buildings + n_cars:
            cx, cy = rng.uniform(-60, 60 + path_len), rng.uniform(-30, 30)
            if abs(cy) - 0.9 < 2.0:
                continue
            lo.append([cx - 2.0, cy - 0.9, GROUND_Z])
            hi.append([cx + 2.0, cy + 0.9, GROUND_Z + 1.5])
        self.lo = np.asarray(lo, np.float64)
        self.hi = np.asarray(hi, np.float64)


def _ray_dirs(n_az, az_offset=0.0):
    elev = np.deg2rad(np.linspace(ELEV_TOP_DEG, ELEV_BOTTOM_DEG, N_BEAMS))
    az = az_offset + np.arange(n_az) * (2 * np.pi / n_az)
    ce, se = np.cos(elev)[None, :], np.sin(elev)[None, :]
    d = np.stack([np.cos(az)[:, None] * ce, np.sin(az)[:, None] * ce, np.broadcast_to(se, (n_az, N_BEAMS))], -1)
    return d.reshape(-1, 3)  # azimuth-major scan order


def _cast(world, origin, dirs):
    """Nearest hit of every ray: (t, normal); t = inf for a miss."""
    m = dirs.shape[0]
    t_best = np.full(m, np.inf)
    normal = np.zeros((m, 3))
    dz = dirs[:, 2]
    with np.errstate(divide="ignore", invalid="ignore"):
        tg = np.where(dz < 0, (GROUND_Z - origin[2]) / dz, np.inf)
    ok = tg < t_best
    t_best = np.where(ok, tg, t_best)
    normal[ok] = (0, 0, 1)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = [1.0 / dirs[:, a] for a in range(3)]
    sgn = [-np.sign(dirs[:, a]) for a in range(3)]
    # broad phase: a box can only be hit by rays whose azimuth falls inside its angular span
    for b in range(world.lo.shape[0]):
        tn, tf = [], []
        for a in range(3):
            t0 = (world.lo[b, a] - origin[a]) * inv[a]
            t1 = (world.hi[b, a] - origin[a]) * inv[a]
            tn.append(np.minimum(t0, t1))
            tf.append(np.maximum(t0, t1))
        tnear = np.maximum(np.maximum(tn[0], tn[1]), tn[2])
        tfar = np.minimum(np.minimum(tf[0], tf[1]), tf[2])
        hit = np.flatnonzero((tnear <= tfar) & (tnear > 0) & (tnear < t_best))
        if hit.size == 0:
            continue
        th = tnear[hit]
        t_best[hit] = th
        nb = np.zeros((hit.size, 3))
        done = np.zeros(hit.size, bool)
        for a in range(3):  # first axis attaining the entry time, as argmax would pick
            m_a = (tn[a][hit] == th) & ~done
            nb[m_a, a] = sgn[a][hit][m_a]
            done |= m_a
        normal[hit] = nb
    return t_best, normal


def _unique_rows(pts):
    """Indices (ascending) of the first occurrence of every distinct xyz row."""
    b = np.ascontiguousarray(pts, np.float32).view(np.uint32).astype(np.uint64)
    h = (b[:, 0] * np.uint64(0x9E3779B97F4A7C15)) ^ (b[:, 1] * np.uint64(0xC2B2AE3D27D4EB4F)) ^ \
        (b[:, 2] * np.uint64(0x165667B19E3779F9))
    if np.unique(h).size == h.size:  # no hash collision => all rows distinct
        return np.arange(pts.shape[0])
    _, first = np.unique(pts, axis=0, return_index=True)
    first.sort()
    return first


def lidar_frame(n, seed=BASE_SEED, origin=(0.0, 0.0, 0.0), world=None, noise_seed=None):
    """One frame of exactly ``n`` unique points: (xyz float32 [n,3], normals float32 [n,3]).
    ``noise_seed`` draws an independent range-noise realisation of the same ray hits."""
    world = world or World(seed)
    origin = np.asarray(origin, np.float64)
    n_az = int(np.ceil(n / N_BEAMS / 0.55))
    while True:
        dirs = _ray_dirs(n_az)
        t, nrm = _cast(world, origin, dirs)
        keep = t <= MAX_RANGE
        if keep.sum() >= n:
            break
        n_az = int(n_az * 1.25) + 1
    rng_sel = np.random.default_rng(seed + 17)
    rng_noise = np.random.default_rng((seed if noise_seed is None else noise_seed) + 101)
    idx = np.flatnonzero(keep)
    noise = rng_noise.normal(0.0, RANGE_SIGMA, size=t.shape[0])
    pts = (origin + dirs[idx] * (t[idx] + noise[idx])[:, None]).astype(np.float32)
    nrm = nrm[idx].astype(np.float32)
    first = _unique_rows(pts)
    if first.size < n:
        raise RuntimeError("synthetic frame: not enough unique points (%d < %d)" % (first.size, n))
    sel = np.sort(rng_sel.choice(first.size, size=n, replace=False))
    sel = first[sel]
    return np.ascontiguousarray(pts[sel]), np.ascontiguousarray(nrm[sel])


def aggregate(n, frames, seed=BASE_SEED, spacing=2.0, lateral=0.0, world=None, noise_seed=None):
    """``frames`` frames ``spacing`` metres apart along +x (offset ``lateral`` in y), expressed
    in the first frame, truncated to exactly ``n`` points: (xyz [n,3], normals [n,3])."""
    world = world or World(seed)
    per = int(np.ceil(n / frames)) + 64
    P, N = [], []
    for f in range(frames):
        p, nr = lidar_frame(per, seed=seed + f, origin=(f * spacing, lateral, 0.0), world=world,
                            noise_seed=None if noise_seed is None else noise_seed + f)
        P.append(p)
        N.append(nr)
    pts, nrm = np.concatenate(P), np.concatenate(N)
    first = _unique_rows(pts)
    if first.size < n:
        raise RuntimeError("synthetic aggregate: not enough unique points")
    first = first[:n]
    return np.ascontiguousarray(pts[first]), np.ascontiguousarray(nrm[first])


def rot_zx(deg_z, deg_x):
    cz, sz = np.cos(np.deg2rad(deg_z)), np.sin(np.deg2rad(deg_z))
    cx, sx = np.cos(np.deg2rad(deg_x)), np.sin(np.deg2rad(deg_x))
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1.0]])
    Rx = np.array([[1.0, 0, 0], [0, cx, -sx], [0, sx, cx]])
    return Rz @ Rx


KNOWN_R = rot_zx(2.0, 0.5)
KNOWN_T = np.array([0.30, -0.20, 0.05])


def icp_pair(n, frames=1, seed=BASE_SEED):
    """(src, tar, tar_normals, T_true): ``tar`` is a cloud; ``src`` is an independent noise
    realisation of the same surface hits moved by the INVERSE of the known transform, so that
    ICP from the identity should recover T_true = [KNOWN_R | KNOWN_T] (src -> tar)."""
    world = World(seed)
    if frames == 1:
        tar, nrm = lidar_frame(n, seed=seed, world=world)
        src0, _ = lidar_frame(n, seed=seed, world=world, noise_seed=seed + 5000)
    else:
        tar, nrm = aggregate(n, frames, seed=seed, world=world)
        src0, _ = aggregate(n, frames, seed=seed, world=world, noise_seed=seed + 5000)
    # src = R^T (src0 - t)
    src = ((src0.astype(np.float64) - KNOWN_T) @ KNOWN_R).astype(np.float32)
    T = np.eye(4)
    T[:3, :3] = KNOWN_R
    T[:3, 3] = KNOWN_T
    return np.ascontiguousarray(src), tar, nrm, T


def uniform_cloud(n, seed=0, scale=100.0):
    rng = np.random.default_rng(seed)
    return (rng.random((n, 3)) * scale).astype(np.float32)
