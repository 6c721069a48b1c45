"""ctypes loader for the parity checkers under oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT.

Two interchangeable back ends with the same methods:

* ``Oracle("port")`` -- oracle/libpcnn_oracle.so, the CPU restatement (port_nn.cpp,
  port_icp.cpp); always buildable.
* ``Oracle("ref")``  -- oracle/_ref/libpcnn_ref.so (+ libpcnn_ref_icp.so): the reference's own
  headers compiled unmodified (ref_nn.cpp drives Homework2/hw2/include/*.hpp; ref_icp.cpp
  drives the vendored nanoflann.hpp).  Built only where /root/reference is mounted; the built
  files travel to the GPU box.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
KDTREE, OCTREE = 0, 1

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)


def _ptr(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def build_oracle(ref=True):
    """Compile the checkers (building the checker is not using it)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    if ref:
        subprocess.check_call(["make", "-s", "-C", _HERE, "_ref"])


def have(which):
    if which == "port":
        return os.path.exists(os.path.join(_HERE, "libpcnn_oracle.so"))
    return os.path.exists(os.path.join(_HERE, "_ref", "libpcnn_ref.so"))


def _xyz(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    assert a.ndim == 2 and a.shape[1] in (3, 4), a.shape
    return a


class Oracle:
    def __init__(self, which="port"):
        self.which = which
        if which == "port":
            self.nn = C.CDLL(os.path.join(_HERE, "libpcnn_oracle.so"))
            self.icp_lib = self.nn
            self.pfx = "pcport_"
        elif which == "ref":
            self.nn = C.CDLL(os.path.join(_HERE, "_ref", "libpcnn_ref.so"))
            self.icp_lib = C.CDLL(os.path.join(_HERE, "_ref", "libpcnn_ref_icp.so"))
            self.pfx = "pcref_"
        else:
            raise ValueError(which)
        f = self._fn("build")
        f.restype = C.c_void_p
        f.argtypes = [C.c_int, _f32p, C.c_int64, C.c_int, C.c_int, C.c_double]
        f = self._fn("destroy")
        f.restype = None
        f.argtypes = [C.c_void_p]
        f = self._fn("knn")
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, _f32p, C.c_int64, C.c_int, C.c_int, _i32p, _f64p, _i32p, C.c_int]
        f = self._fn("radius")
        f.restype = C.c_int64
        f.argtypes = [C.c_void_p, _f32p, C.c_int64, C.c_int, C.c_double, _i64p, C.c_int]
        f = self._fn("radius_fetch")
        f.restype = None
        f.argtypes = [C.c_void_p, _i32p, _f64p]
        for name in ("knn_checksum",):
            f = self._fn(name)
            f.restype = C.c_int64
            f.argtypes = [C.c_void_p, _f32p, C.c_int64, C.c_int, C.c_int, C.c_int]
        f = self._fn("radius_checksum")
        f.restype = C.c_int64
        f.argtypes = [C.c_void_p, _f32p, C.c_int64, C.c_int, C.c_double, C.c_int]
        f = self._fn("hardware_threads")
        f.restype = C.c_int
        f.argtypes = []
        f = getattr(self.icp_lib, self.pfx + "icp")
        f.restype = C.c_int
        f.argtypes = [_f32p, C.c_int64, _f32p, C.c_int64, _f32p, _f32p, C.c_int, C.c_float,
                      C.c_float, C.c_int, C.c_int, C.c_int, _f32p, _i64p, _f32p, _i32p]
        f = getattr(self.icp_lib, self.pfx + "nn1")
        f.restype = C.c_int
        f.argtypes = [_f32p, C.c_int64, _f32p, C.c_int64, _i32p, _f32p, C.c_int]
        if which == "port":
            f = self.nn.pcport_brute_knn
            f.restype = C.c_int
            f.argtypes = [_f32p, C.c_int64, C.c_int, _f32p, C.c_int64, C.c_int, C.c_int, _i32p,
                          _f64p, C.c_int]

    def _fn(self, name):
        return getattr(self.nn, self.pfx + name)

    # ---- ISS keypoints (port only: oracle/port_iss.cpp)
    def iss(self, xyz, local_radius, non_max_radius, gamma21=0.9, gamma32=0.9, min_neighbors=5, weighted=True,
            nthreads=8):
        """(keypoint indices ascending, lambda3[n]) as ISSKeypoint::compute (iss_detector.cpp:38-111)."""
        a = np.ascontiguousarray(_xyz(xyz)[:, :3])
        n = a.shape[0]
        lam = np.empty(max(n, 1), np.float32)
        flag = np.zeros(max(n, 1), np.uint8)
        f = self.nn.pco_iss
        f.restype = C.c_int
        f.argtypes = [_f32p, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, _f32p,
                      C.POINTER(C.c_uint8), C.c_int]
        f(_ptr(a, _f32p), n, local_radius, non_max_radius, gamma21, gamma32, min_neighbors, 1 if weighted else 0,
          _ptr(lam, _f32p), flag.ctypes.data_as(C.POINTER(C.c_uint8)), nthreads)
        return np.flatnonzero(flag[:n]).astype(np.int32), lam[:n]

    # ---- descriptor matching + RANSAC (port only: oracle/port_ransac.cpp)
    def desc_nn1(self, a, b, nthreads=8):
        a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
        idx = np.full(max(len(a), 1), -1, np.int32); d2 = np.zeros(max(len(a), 1), np.float32)
        f = self.nn.pco_desc_nn1
        f.restype = C.c_int
        f.argtypes = [_f32p, C.c_int64, _f32p, C.c_int64, C.c_int, _i32p, _f32p, C.c_int]
        f(_ptr(a, _f32p), len(a), _ptr(b, _f32p), len(b), a.shape[1], _ptr(idx, _i32p), _ptr(d2, _f32p), nthreads)
        return idx[:len(a)], d2[:len(a)]

    def match(self, src_desc, tar_desc, mode=0, rejection_rate=0.0, nthreads=8):
        sd = np.ascontiguousarray(src_desc, np.float32); td = np.ascontiguousarray(tar_desc, np.float32)
        out = np.zeros((max(len(sd) + len(td), 1), 2), np.int32)
        f = self.nn.pco_match
        f.restype = C.c_int64
        f.argtypes = [_f32p, C.c_int64, _f32p, C.c_int64, C.c_int, C.c_int, C.c_float, _i32p, C.c_int]
        n = f(_ptr(sd, _f32p), len(sd), _ptr(td, _f32p), len(td), sd.shape[1], mode, rejection_rate, _ptr(out, _i32p), nthreads)
        return out[:n].copy()

    def ransac(self, src_kp, tar_kp, pairs, max_iter, dist_threshold, coplanar_min=0.15, seed=0, nthreads=8):
        s = np.ascontiguousarray(src_kp, np.float32)[:, :3].copy(); t = np.ascontiguousarray(tar_kp, np.float32)[:, :3].copy()
        p = np.ascontiguousarray(pairs, np.int32)
        R = np.zeros(9, np.float32); tt = np.zeros(3, np.float32); it = C.c_int64(-1)
        f = self.nn.pco_ransac
        f.restype = C.c_int64
        f.argtypes = [_f32p, _f32p, _i32p, C.c_int64, C.c_int, C.c_float, C.c_float, C.c_uint64, _f32p, _f32p,
                      C.POINTER(C.c_int64), C.c_int]
        cons = f(_ptr(s, _f32p), _ptr(t, _f32p), _ptr(p, _i32p), len(p), max_iter, dist_threshold, coplanar_min, seed,
                 _ptr(R, _f32p), _ptr(tt, _f32p), C.byref(it), nthreads)
        return {"R": R.reshape(3, 3), "t": tt, "consensus": int(cons), "best_iter": int(it.value)}

    def iss_counts(self, xyz, radius, nthreads=8):
        a = np.ascontiguousarray(_xyz(xyz)[:, :3])
        out = np.zeros(max(a.shape[0], 1), np.int32)
        f = self.nn.pco_iss_counts
        f.restype = C.c_int
        f.argtypes = [_f32p, C.c_int64, C.c_float, _i32p, C.c_int]
        f(_ptr(a, _f32p), a.shape[0], radius, _ptr(out, _i32p), nthreads)
        return out[:a.shape[0]]

    def hardware_threads(self):
        return int(self._fn("hardware_threads")())

    # ---- index
    def build(self, kind, xyz, leaf_size=32, min_extent=1e-4):
        xyz = _xyz(xyz)
        h = self._fn("build")(kind, _ptr(xyz, _f32p), xyz.shape[0], xyz.shape[1], leaf_size,
                              float(min_extent))
        return _Tree(self, h, xyz)

    # ---- ICP (restated loop); returns dict
    def icp(self, src, tar, tar_normals=None, init_T=None, max_iter=30, max_corres_dist=1.0,
            loss_epsilon=0.0, mode=0, accum=1, nthreads=1, want_corr=False):
        src = np.ascontiguousarray(src, dtype=np.float32).reshape(-1, 3)
        tar = np.ascontiguousarray(tar, dtype=np.float32).reshape(-1, 3)
        nrm = None if tar_normals is None else np.ascontiguousarray(tar_normals, np.float32).reshape(-1, 3)
        T0 = np.eye(4, dtype=np.float32) if init_T is None else np.ascontiguousarray(init_T, np.float32)
        T = np.zeros(16, np.float32)
        info = np.zeros(4, np.int64)
        loss = np.zeros(1, np.float32)
        corr = np.zeros(src.shape[0], np.int32) if want_corr else None
        getattr(self.icp_lib, self.pfx + "icp")(
            _ptr(src, _f32p), src.shape[0], _ptr(tar, _f32p), tar.shape[0], _ptr(nrm, _f32p),
            _ptr(T0, _f32p), max_iter, max_corres_dist, loss_epsilon, mode, accum, nthreads,
            _ptr(T, _f32p), _ptr(info, _i64p), _ptr(loss, _f32p), _ptr(corr, _i32p))
        return {"T": T.reshape(4, 4), "iters": int(info[0]), "converged": int(info[1]),
                "status": int(info[2]), "pairs": int(info[3]), "loss": float(loss[0]), "corr": corr}

    def nn1(self, tar, q, nthreads=1):
        tar = np.ascontiguousarray(tar, dtype=np.float32).reshape(-1, 3)
        q = np.ascontiguousarray(q, dtype=np.float32).reshape(-1, 3)
        idx = np.zeros(q.shape[0], np.int32)
        d2 = np.zeros(q.shape[0], np.float32)
        getattr(self.icp_lib, self.pfx + "nn1")(_ptr(tar, _f32p), tar.shape[0], _ptr(q, _f32p),
                                                q.shape[0], _ptr(idx, _i32p), _ptr(d2, _f32p), nthreads)
        return idx, d2

    def brute_knn(self, xyz, q, k, nthreads=1):
        """Canonical answer: k smallest under (d2 double, index); returns (idx, d2)."""
        assert self.which == "port"
        xyz, q = _xyz(xyz), _xyz(q)
        idx = np.zeros((q.shape[0], k), np.int32)
        d2 = np.zeros((q.shape[0], k), np.float64)
        self.nn.pcport_brute_knn(_ptr(xyz, _f32p), xyz.shape[0], xyz.shape[1], _ptr(q, _f32p),
                                 q.shape[0], q.shape[1], k, _ptr(idx, _i32p), _ptr(d2, _f64p), nthreads)
        return idx, d2


def pca_normals(xyz, max_nn=10, radius=5.0, oracle=None, nthreads=8):
    """CPU restatement of the reference's normal estimation loop
    (Homework1/YuF_KIT-.../pca_normal.py:85-101, PCA() :17-33): hybrid search (<= max_nn nearest
    within radius, the point itself included), PCA of the found points, eigenvector of the
    smallest eigenvalue; zero when fewer than 3 are found.  The neighbour search is this oracle's
    exact brute-force kNN (open3d is not installed here: PARITY UNPINNED); the PCA is numpy's
    eigh on the double scatter matrix (the reference calls np.linalg.eig on the same matrix).
    Returns (normals [n,3] float64, eigenvalues [n,3] ascending) -- sign of a normal arbitrary."""
    orc = oracle or Oracle("port")
    p = np.ascontiguousarray(xyz, np.float32)[:, :3]
    n = len(p)
    k = min(max_nn, n)
    idx, d2 = orc.brute_knn(p, p, k, nthreads=nthreads)
    dist = np.sqrt(d2)
    ok = (idx >= 0) & (dist <= radius)
    m = ok.sum(1)
    P = p.astype(np.float64)[np.where(ok, idx, 0)]                     # [n, k, 3]
    w = ok[:, :, None].astype(np.float64)
    c = (P * w).sum(1) / np.maximum(m, 1)[:, None]                      # :18
    D = (P - c[:, None, :]) * w                                          # :19
    XTX = np.einsum("nki,nkj->nij", D, D)                                # :20
    ev, evec = np.linalg.eigh(XTX)                                       # ascending
    normals = evec[:, :, 0].copy()
    normals[m < 3] = 0.0                                                 # :94
    return normals, ev


def voxel_grid(xyz, normals, leaf, sequential_float=False):
    """CPU restatement of pcl::VoxelGrid<pcl::PointNormal>::applyFilter as the reference calls it
    (Homework9/hw9/src/registration.cpp:37-41, :688-691), after PCL 1.8-1.10
    pcl/filters/impl/voxel_grid.hpp and pcl/common/impl/accumulators.hpp (PCL is not vendored in
    the reference and not installed here: PARITY UNPINNED, no golden vector exists).
    Returns (xyz, normals, counts): one row per occupied voxel in ascending voxel index.
    Means in float64 rounded once by default; ``sequential_float=True`` accumulates in float32 in
    input order, which is what PCL does when std::sort happens to keep that order (small cases)."""
    p = np.ascontiguousarray(xyz, np.float32)[:, :3]
    n = len(p)
    leaf = np.broadcast_to(np.asarray(leaf, np.float32), (3,))
    inv = (np.float32(1.0) / leaf).astype(np.float32)
    if n == 0:
        return p.copy(), (None if normals is None else np.zeros((0, 3), np.float32)), np.zeros(0, np.int64)
    lo, hi = p.min(0), p.max(0)
    d = ((hi - lo) * inv).astype(np.int64) + 1
    min_b = np.floor(lo * inv).astype(np.int64)
    max_b = np.floor(hi * inv).astype(np.int64)
    div = max_b - min_b + 1
    if int(d[0]) * int(d[1]) * int(d[2]) > 2**31 - 1 or int(div[0]) * int(div[1]) * int(div[2]) > 2**31 - 1:
        return p.copy(), (None if normals is None else np.ascontiguousarray(normals, np.float32)[:, :3].copy()), np.ones(n, np.int64)
    ijk = (np.floor(p * inv) - min_b.astype(np.float32)).astype(np.int64)   # float32 arithmetic, then the int cast
    idx = ijk[:, 0] + ijk[:, 1] * div[0] + ijk[:, 2] * div[0] * div[1]
    order = np.argsort(idx, kind="stable")
    sidx = idx[order]
    heads = np.flatnonzero(np.append(True, sidx[1:] != sidx[:-1]))
    counts = np.diff(np.append(heads, n))
    nr = None if normals is None else np.ascontiguousarray(normals, np.float32)[:, :3]
    if sequential_float:
        ox = np.zeros((len(heads), 3), np.float32)
        on = None if nr is None else np.zeros((len(heads), 3), np.float32)
        for v, (b, c) in enumerate(zip(heads, counts)):
            acc = np.zeros(3, np.float32)
            accn = np.zeros(3, np.float32)
            for i in order[b:b + c]:
                acc = (acc + p[i]).astype(np.float32)
                if nr is not None:
                    accn = (accn + nr[i]).astype(np.float32)
            ox[v] = acc / np.float32(c)
            if nr is not None:
                with np.errstate(invalid="ignore", divide="ignore"):
                    on[v] = accn / np.sqrt((accn * accn).sum(dtype=np.float32))
        return ox, on, counts
    s = np.add.reduceat(p[order].astype(np.float64), heads, axis=0)
    ox = (s / counts[:, None]).astype(np.float32)
    on = None
    if nr is not None:
        sn = np.add.reduceat(nr[order].astype(np.float64), heads, axis=0)
        with np.errstate(invalid="ignore", divide="ignore"):
            on = (sn / np.sqrt((sn * sn).sum(1))[:, None]).astype(np.float32)
    return ox, on, counts


def ref_hw7_radius_all(xyz, radius):
    """Neighbour lists of every point by the UNMODIFIED Homework7 float KD-tree
    (oracle/_ref/libpcnn_ref_hw7.so, oracle/ref_iss_nn.cpp): CSR (offsets, idx in traversal order)."""
    L = C.CDLL(os.path.join(_HERE, "_ref", "libpcnn_ref_hw7.so"))
    f = L.ref7_radius_all
    f.restype = C.c_int
    f.argtypes = [_f32p, C.c_int64, C.c_float, _i64p, _i32p]
    a = np.ascontiguousarray(_xyz(xyz)[:, :3])
    n = a.shape[0]
    off = np.zeros(n + 1, np.int64)
    f(_ptr(a, _f32p), n, radius, _ptr(off, _i64p), None)
    idx = np.empty(max(int(off[n]), 1), np.int32)
    f(_ptr(a, _f32p), n, radius, _ptr(off, _i64p), _ptr(idx, _i32p))
    return off, idx[:int(off[n])]


class _Tree:
    def __init__(self, lib, h, xyz):
        self.lib, self.h, self.xyz = lib, h, xyz

    def __del__(self):
        if getattr(self, "h", None):
            self.lib._fn("destroy")(self.h)
            self.h = None

    def knn(self, q, k, nthreads=1, want_cmp=False):
        q = _xyz(q)
        idx = np.zeros((q.shape[0], k), np.int32)
        dist = np.zeros((q.shape[0], k), np.float64)
        cmp_ = np.zeros(q.shape[0], np.int32) if want_cmp else None
        self.lib._fn("knn")(self.h, _ptr(q, _f32p), q.shape[0], q.shape[1], k, _ptr(idx, _i32p),
                            _ptr(dist, _f64p), _ptr(cmp_, _i32p), nthreads)
        return (idx, dist, cmp_) if want_cmp else (idx, dist)

    def radius(self, q, r, nthreads=1):
        q = _xyz(q)
        off = np.zeros(q.shape[0] + 1, np.int64)
        total = self.lib._fn("radius")(self.h, _ptr(q, _f32p), q.shape[0], q.shape[1], float(r),
                                       _ptr(off, _i64p), nthreads)
        idx = np.zeros(total, np.int32)
        dist = np.zeros(total, np.float64)
        self.lib._fn("radius_fetch")(self.h, _ptr(idx, _i32p), _ptr(dist, _f64p))
        return off, idx, dist

    def knn_checksum(self, q, k, nthreads=1):
        q = _xyz(q)
        return int(self.lib._fn("knn_checksum")(self.h, _ptr(q, _f32p), q.shape[0], q.shape[1], k, nthreads))

    def radius_checksum(self, q, r, nthreads=1):
        q = _xyz(q)
        return int(self.lib._fn("radius_checksum")(self.h, _ptr(q, _f32p), q.shape[0], q.shape[1],
                                                   float(r), nthreads))
