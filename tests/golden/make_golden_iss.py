"""Regenerates tests/golden/golden_iss.npz.  Needs the reference mount (/root/reference) and
oracle/_ref built (`make -C oracle all`); run from the repo root:
    python tests/golden/make_golden_iss.py

golden_iss.npz
  pts            Homework7/hw7/test_data/chair_0001.ply, xyz float32 (10,000 points) -- one of
                 the clouds the reference's own main.cpp runs its detector on
  local_radius, non_max_radius, gamma21, gamma32, min_neighbors
                 the parameters of Homework7/hw7/main.cpp:93-101 (test = 0.02: 6*test, 4*test, 0.9, 0.9, 5)
  ref_counts_local / ref_counts_nms
                 |rnn_idx[i]| for every point at both radii, from the UNMODIFIED Homework7 float
                 KD-tree (oracle/_ref/libpcnn_ref_hw7.so): pins the neighbourhoods
  ref_nbr_hash_local
                 per point, sum of (index * 2654435761 mod 2^32) over its neighbour list: pins the sets
  port_keypoints, port_lambda3 (weighted) / port_keypoints_unweighted, port_lambda3_unweighted
                 the restated detector (oracle/port_iss.cpp) on those neighbourhoods.  The
                 reference's own eigenvalue step needs Eigen, which is not available: unpinned.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pcoracle as po  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
PLY = "/root/reference/Homework7/hw7/test_data/chair_0001.ply"


def read_ply_xyz(path):
    b = open(path, "rb").read()
    i = b.index(b"end_header\n") + len(b"end_header\n")
    hdr = b[:i].decode()
    n = int([ln for ln in hdr.split("\n") if ln.startswith("element vertex")][0].split()[-1])
    nprop = sum(1 for ln in hdr.split("\n") if ln.startswith("property float"))
    return np.frombuffer(b[i:i + n * 4 * nprop], dtype="<f4").reshape(n, nprop)[:, :3].copy()


def nbr_hash(off, idx):
    h = (idx.astype(np.uint64) * np.uint64(2654435761)) % np.uint64(1 << 32)
    return np.add.reduceat(np.append(h, np.uint64(0)), off[:-1]).astype(np.uint64) * (np.diff(off) > 0)


def main():
    pts = read_ply_xyz(PLY)
    test = np.float32(0.02)
    rl, rn = np.float32(6 * test), np.float32(4 * test)
    out = {"pts": pts, "local_radius": rl, "non_max_radius": rn, "gamma21": np.float32(0.9),
           "gamma32": np.float32(0.9), "min_neighbors": np.int32(5)}
    off, idx = po.ref_hw7_radius_all(pts, float(rl))
    out["ref_counts_local"] = np.diff(off).astype(np.int32)
    out["ref_nbr_hash_local"] = nbr_hash(off, idx)
    off2, _ = po.ref_hw7_radius_all(pts, float(rn))
    out["ref_counts_nms"] = np.diff(off2).astype(np.int32)
    P = po.Oracle("port")
    kp, lam = P.iss(pts, float(rl), float(rn), 0.9, 0.9, 5, True)
    out["port_keypoints"], out["port_lambda3"] = kp, lam
    kp, lam = P.iss(pts, float(rl), float(rn), 0.9, 0.9, 5, False)
    out["port_keypoints_unweighted"], out["port_lambda3_unweighted"] = kp, lam
    np.savez_compressed(os.path.join(HERE, "golden_iss.npz"), **out)
    print("chair_0001: %d points, mean |N| = %.1f, %d / %d keypoints (weighted / unweighted)" % (
        len(pts), out["ref_counts_local"].mean(), len(out["port_keypoints"]), len(out["port_keypoints_unweighted"])))


if __name__ == "__main__":
    main()
