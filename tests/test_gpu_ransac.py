"""GPU parity tests for descriptor matching and RANSAC, through the C ABI
(pcnn_desc_nn1, pcnn_match_descriptors, pcnn_ransac) against the restatement
(oracle/port_ransac.cpp).  The reference's own run is not reproducible (mt19937 seeded from
std::random_device, registration.cpp:298-299) and needs Eigen + PCL descriptors: unpinned.

Bars: 1-NN indices and float32 squared distances bit-exact (same arithmetic, no FMA on either
side); correspondence lists identical; RANSAC: same winning iteration and consensus, pose within 1e-5."""
import numpy as np
import pytest

import pcnn_b200 as pc

pytestmark = pytest.mark.gpu


def _descs(ns, nt, dim=33, seed=0):
    rng = np.random.default_rng(seed)
    sd = rng.random((ns, dim)).astype(np.float32) * 100.0  # FPFH bins are percentages
    td = (sd[rng.integers(0, ns, nt)] + rng.normal(0, 0.5, (nt, dim))).astype(np.float32)
    return sd, td


@pytest.mark.parametrize("ns,nt,dim", [(3000, 2500, 33), (257, 1, 33), (1, 300, 33), (500, 500, 3), (100, 130, 64)])
def test_desc_nn1_bit_exact(oracle_port, ns, nt, dim):
    sd, td = _descs(ns, nt, dim, seed=ns + dim)
    idx, d2 = pc.desc_nn1(td, sd)
    oidx, od2 = oracle_port.desc_nn1(td, sd)
    assert np.array_equal(idx.cpu().numpy(), oidx)
    assert np.array_equal(d2.cpu().numpy(), od2)


def test_desc_nn1_ties_lowest_index(oracle_port):
    b = np.zeros((10, 33), np.float32)
    b[3:] = 1.0
    a = np.zeros((4, 33), np.float32)
    idx, d2 = pc.desc_nn1(a, b)
    assert (idx.cpu().numpy() == 0).all() and (d2.cpu().numpy() == 0).all()
    a[:] = 1.0
    idx, _ = pc.desc_nn1(a, b)
    assert (idx.cpu().numpy() == 3).all()


@pytest.mark.parametrize("mode,omode", [("inter", 0), ("union", 1)])
@pytest.mark.parametrize("rate", [0.0, 0.2, 0.95, 1.0])
def test_match_descriptors_equals_restatement(oracle_port, mode, omode, rate):
    sd, td = _descs(2000, 1700, 33, seed=5)
    got = pc.match_descriptors(sd, td, mode, rate).cpu().numpy()
    ref = oracle_port.match(sd, td, omode, rate)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)
    if mode == "inter" and rate == 0.0:
        assert len(got) > 500 and len(np.unique(got[:, 0])) == len(got) and len(np.unique(got[:, 1])) == len(got)


def _scene(n=400, outliers=0.4, seed=3):
    rng = np.random.default_rng(seed)
    src = rng.uniform(-10, 10, (n, 3)).astype(np.float32)
    R = pc.synth.rot_zx(25.0, 7.0).astype(np.float32)
    t = np.array([1.5, -2.0, 0.7], np.float32)
    tar = (src @ R.T + t + rng.normal(0, 0.01, (n, 3))).astype(np.float32)
    pairs = np.stack([np.arange(n), np.arange(n)], 1).astype(np.int32)
    bad = rng.random(n) < outliers
    pairs[bad, 1] = rng.integers(0, n, bad.sum())
    return src, tar, pairs, R, t


@pytest.mark.parametrize("seed", [0, 1, 12345678901234])
def test_ransac_equals_restatement_and_recovers_pose(oracle_port, seed):
    src, tar, pairs, R, t = _scene()
    got = pc.ransac(src, tar, pairs, max_iter=5000, dist_threshold=0.1, coplanar_min=0.15, seed=seed)
    ref = oracle_port.ransac(src, tar, pairs, 5000, 0.1, 0.15, seed)
    assert got["best_iter"] == ref["best_iter"] and got["consensus"] == ref["consensus"]
    assert np.abs(got["R"] - ref["R"]).max() <= 1e-5 and np.abs(got["t"] - ref["t"]).max() <= 1e-5
    assert got["consensus"] > 0.5 * len(pairs)
    assert np.abs(got["R"] - R).max() < 0.01 and np.abs(got["t"] - t).max() < 0.1


def test_ransac_full_reference_iteration_count(oracle_port):
    # main.cpp:86 runs 80,000 hypotheses; a larger correspondence set
    src, tar, pairs, R, t = _scene(n=3000, outliers=0.7, seed=9)
    got = pc.ransac(src, tar, pairs, max_iter=80000, dist_threshold=0.1, seed=42)
    ref = oracle_port.ransac(src, tar, pairs, 80000, 0.1, 0.15, 42, nthreads=16)
    assert got["best_iter"] == ref["best_iter"] and got["consensus"] == ref["consensus"]
    assert np.abs(got["R"] - R).max() < 0.01


def test_ransac_and_match_edges():
    import ctypes as C
    import torch
    from pcnn_b200 import api
    # fewer than 4 correspondences: an argument error (the reference would spin forever)
    src = torch.zeros((3, 3), device="cuda"); pairs = torch.zeros((3, 2), dtype=torch.int32, device="cuda")
    prm = api.RansacParams(10, 0.1, 0.15, 0, 0)
    res = api.RansacResult()
    rc = api.lib().pcnn_ransac(api._t_ptr(src), api._t_ptr(src), api._t_ptr(pairs), 3, C.byref(prm), C.byref(res), 0, None)
    assert rc == api.ERR_INVALID
    # all points coplanar: no hypothesis passes the check, identity comes back with consensus 0
    gx, gy = np.meshgrid(np.arange(5, dtype=np.float32), np.arange(5, dtype=np.float32))
    flat = np.stack([gx.ravel(), gy.ravel(), np.zeros(25, np.float32)], 1)
    p = np.stack([np.arange(25), np.arange(25)], 1).astype(np.int32)
    got = pc.ransac(flat, flat, p, max_iter=64, dist_threshold=0.1)
    assert got["consensus"] == 0 and got["best_iter"] == -1 and np.array_equal(got["R"], np.eye(3, dtype=np.float32))
    # empty descriptor sets
    e = pc.match_descriptors(np.zeros((0, 33), np.float32), np.ones((5, 33), np.float32))
    assert e.shape == (0, 2)
    # dimension beyond the supported width
    a = torch.zeros((4, 65), device="cuda")
    idx = torch.zeros(4, dtype=torch.int32, device="cuda"); d2 = torch.zeros(4, device="cuda")
    rc = api.lib().pcnn_desc_nn1(api._t_ptr(a), 4, api._t_ptr(a), 4, 65, api._t_ptr(idx), api._t_ptr(d2), 0, None)
    assert rc == api.ERR_UNSUPPORTED
