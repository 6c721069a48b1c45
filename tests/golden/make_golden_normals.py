"""Regenerates tests/golden/golden_normals.npz.  Needs the reference mount; run from the repo root:
    python tests/golden/make_golden_normals.py

The reference's normal estimation (Homework1/YuF_KIT-.../pca_normal.py) cannot be imported: the
module pulls in open3d, pyntcloud and a pybind11 extension at import time, none of which are
installed.  Its PCA() (pca_normal.py:17-33) is pure numpy, though: this script cuts that
function's SOURCE TEXT out of the reference file where it lies (nothing is copied into the repo),
executes it, and runs it exactly as the per-point loop of :90-101 does -- PCA(points[idx_list, :])
and normals[i] = real(eig_vectors[:, 2]) when num >= 3.  The neighbour lists stand in for
open3d's search_hybrid_vector_3d(radius, max_nn): the (at most) max_nn nearest within radius, the
point itself included, from the exact brute-force kNN of oracle/port_nn.cpp.

golden_normals.npz  (cloud = golden_iss.npz["pts"] = Homework7/hw7/test_data/chair_0001.ply, 10 000 points)
  radius, max_nn     0.02, 10: on this cloud 5 886 points get fewer than 10 neighbours and 67 fewer
                     than 3 (zero normal), so the radius bound and the num >= 3 rule are both
                     exercised; and a second set at the reference's own 5.0 / 10 (:93), where the
                     radius never binds (the lists are the 10 nearest: normals_r5 only)
  nbr_idx            [n, max_nn] int32, -1 padded: the neighbour lists fed to PCA() at radius 0.02
  normals[_r5]       [n, 3] float64: what the reference's PCA() returned (zero rows: fewer than 3)
  evals              [n, 3] float64 eigenvalues, descending as PCA(sort=True) orders them
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pcoracle as po  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/Homework1/YuF_KIT-第1章作业/pca_normal.py"


def reference_pca():
    """The reference's PCA() compiled from its own source text (pca_normal.py:17-33)."""
    lines = open(REF, encoding="utf-8").read().split("\n")
    a = next(i for i, ln in enumerate(lines) if ln.startswith("def PCA("))
    b = next(i for i, ln in enumerate(lines) if i > a and ln.startswith("def "))
    ns = {"np": np}
    exec(compile("\n".join(lines[a:b]), REF, "exec"), ns)
    return ns["PCA"]


def run(PCA, pts, radius, max_nn, P):
    n = len(pts)
    idx, d2 = P.brute_knn(pts, pts, max_nn, nthreads=8)
    keep = np.sqrt(d2) <= radius
    nbr = np.where(keep, idx, -1).astype(np.int32)
    points = pts.astype(np.float64)
    normals = np.zeros((n, 3))
    evals = np.zeros((n, 3))
    for i in range(n):
        lst = nbr[i][nbr[i] >= 0]
        if len(lst) >= 3:                                   # :94
            w, v = PCA(points[lst, :])                      # :100
            normals[i, :] = np.real(v[:, 2])                # :101
            evals[i, :] = np.real(w)
    return nbr, normals, evals


def main():
    pts = np.load(os.path.join(HERE, "golden_iss.npz"))["pts"]
    PCA = reference_pca()
    P = po.Oracle("port")
    out = {"radius": np.float64(0.02), "max_nn": np.int32(10), "radius_r5": np.float64(5.0)}
    out["nbr_idx"], out["normals"], out["evals"] = run(PCA, pts, 0.02, 10, P)
    _, out["normals_r5"], _ = run(PCA, pts, 5.0, 10, P)
    np.savez_compressed(os.path.join(HERE, "golden_normals.npz"), **out)
    m = (out["nbr_idx"] >= 0).sum(1)
    print("golden_normals.npz: %d points, neighbours per point min %d mean %.2f, %d zero normals" %
          (len(pts), m.min(), m.mean(), int((np.abs(out["normals"]).sum(1) == 0).sum())))


if __name__ == "__main__":
    main()
