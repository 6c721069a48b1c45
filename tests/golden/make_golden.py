"""Regenerates the fixtures in tests/golden/ by running the UNMODIFIED reference
(oracle/_ref, built from /root/reference by `make -C oracle _ref`).  Needs the reference
mount; run from the repo root:  python tests/golden/make_golden.py

Fixtures
  kitti_100k.npy      first 100,000 points of Homework2/hw2/000000.bin (xyz float32) -- the
                      input of the reference's own recorded test (Homework2/hw2/README:13)
  golden_nn.npz       reference outputs on that cloud: the query-5 known answer of
                      result_cpp.txt:11-33 / :44-66 plus 2,000 kNN rows and 200 radius rows per
                      tree kind, exactly as the reference's result sets hold them
  golden_icp.npz      restated ICP loop driven by the reference's nanoflann on a synthetic pair
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pcoracle as po  # noqa: E402
import pcnn_b200 as pc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF_BIN = "/root/reference/Homework2/hw2/000000.bin"


def main():
    raw = np.fromfile(REF_BIN, dtype=np.float32).reshape(-1, 4)
    db = np.ascontiguousarray(raw[:100000, :3])
    np.save(os.path.join(HERE, "kitti_100k.npy"), db)
    R = po.Oracle("ref")
    out = {}
    qsel = np.arange(0, 100000, 50, dtype=np.int64)  # 2,000 in-cloud queries
    rsel = qsel[::10]                                # 200 radius queries
    out["q_index"], out["r_index"] = qsel, rsel
    # perturbed (off-cloud) queries
    rng = np.random.default_rng(7)
    qoff = (db[rng.integers(0, len(db), 500)] + rng.normal(0, 1.5, (500, 3))).astype(np.float32)
    out["q_off"] = qoff
    for kind, name, r in ((po.KDTREE, "kd", 1.0), (po.OCTREE, "oc", 0.5)):
        t = R.build(kind, db, 32, 1e-4)
        i5, d5, c5 = t.knn(db[5:6], 8, want_cmp=True)
        o5, ri5, rd5 = t.radius(db[5:6], 1.66)
        out[name + "_q5_knn_idx"], out[name + "_q5_knn_dist"] = i5[0], d5[0]
        out[name + "_q5_radius_idx"], out[name + "_q5_radius_dist"] = ri5, rd5
        out[name + "_q5_cmp"] = c5
        idx, dist = t.knn(db[qsel], 8, nthreads=8)
        out[name + "_knn_idx"], out[name + "_knn_dist"] = idx, dist
        idx, dist = t.knn(qoff, 8, nthreads=8)
        out[name + "_knn_off_idx"], out[name + "_knn_off_dist"] = idx, dist
        idx, dist = t.knn(db[rsel], 32, nthreads=8)
        out[name + "_knn32_idx"], out[name + "_knn32_dist"] = idx, dist
        off, ridx, rdist = t.radius(db[rsel], r, nthreads=8)
        out[name + "_radius"] = np.float64(r)
        out[name + "_radius_off"], out[name + "_radius_idx"], out[name + "_radius_dist"] = off, ridx, rdist
    np.savez_compressed(os.path.join(HERE, "golden_nn.npz"), **out)

    src, tar, nrm, T = pc.synth.icp_pair(5000, seed=pc.synth.BASE_SEED + 3)
    icp = {"n": np.int64(5000), "seed": np.int64(pc.synth.BASE_SEED + 3), "T_true": T,
           "src_sum": np.float64(src.astype(np.float64).sum()), "tar_sum": np.float64(tar.astype(np.float64).sum())}
    for mode in (0, 1):
        for accum in (0, 1):
            r = R.icp(src, tar, nrm, max_iter=30, max_corres_dist=1.0, loss_epsilon=0.0, mode=mode, accum=accum)
            icp["T_m%d_a%d" % (mode, accum)] = r["T"]
            icp["pairs_m%d_a%d" % (mode, accum)] = np.int64(r["pairs"])
    ni, nd = R.nn1(tar, src)
    icp["nn1_idx"], icp["nn1_d2"] = ni, nd
    np.savez_compressed(os.path.join(HERE, "golden_icp.npz"), **icp)
    print("wrote fixtures to", HERE)


if __name__ == "__main__":
    main()
