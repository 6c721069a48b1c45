"""GPU parity tests for the voxel-grid downsample, through the C ABI
(pcnn_voxel_downsample) against the CPU restatement of pcl::VoxelGrid (oracle/pcoracle.py
voxel_grid).  PCL is absent here and from the reference tree: parity is against the restated
published algorithm (unpinned), plus size-independent properties.

Bars: voxel count, voxel order and per-voxel membership exact (integer work); means within one
float ulp of the double-accumulated mean and within n*eps of a float32 sequential sum."""
import numpy as np
import pytest

import pcnn_b200 as pc
import pcoracle as po

pytestmark = pytest.mark.gpu


def _ulp_close(a, b, ulps=1):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    tol = ulps * np.spacing(np.maximum(np.abs(a), np.abs(b)).astype(np.float32))
    return bool(np.all((np.abs(a - b) <= tol) | (np.isnan(a) & np.isnan(b))))


@pytest.mark.parametrize("leaf", [1.75, 0.5, (0.3, 0.6, 1.2)])
def test_voxel_matches_restatement_on_synthetic_scan(leaf):
    pts, nrm = pc.synth.lidar_frame(120_000, seed=pc.synth.BASE_SEED + 5)
    ox, on = pc.voxel_downsample(pts, nrm, leaf)
    rx, rn, counts = po.voxel_grid(pts, nrm, leaf)
    assert ox.shape == rx.shape and len(ox) == len(counts) and counts.sum() == len(pts)
    assert _ulp_close(ox, rx) and _ulp_close(on, rn, 2)
    assert np.allclose(np.linalg.norm(on, axis=1), 1.0, atol=1e-6)
    # every output point lies inside its voxel, and voxels come in ascending index
    lf = np.broadcast_to(np.asarray(leaf, np.float32), (3,))
    cell = np.floor(ox / lf).astype(np.int64)
    lo = np.floor(pts.min(0) / lf).astype(np.int64)
    div = np.floor(pts.max(0) / lf).astype(np.int64) - lo + 1
    key = (cell - lo) @ np.array([1, div[0], div[0] * div[1]])
    assert (np.diff(key) >= 0).all()


def test_voxel_small_case_matches_float_sequential_sum():
    rng = np.random.default_rng(11)
    pts = rng.uniform(-20, 20, (3000, 3)).astype(np.float32)
    nrm = rng.normal(0, 1, (3000, 3)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    ox, on = pc.voxel_downsample(pts, nrm, 4.0)
    sx, sn, counts = po.voxel_grid(pts, nrm, 4.0, sequential_float=True)
    assert ox.shape == sx.shape
    bound = counts[:, None] * np.float32(6e-8) * 40.0  # n * eps * |coordinate scale|
    assert (np.abs(ox - sx) <= bound).all()
    assert np.allclose(on, sn, atol=1e-5)


def test_voxel_properties_and_edges():
    import ctypes as C
    import torch
    from pcnn_b200 import api
    # idempotence: a filtered cloud holds one point per voxel, filtering again keeps the count
    pts, nrm = pc.synth.lidar_frame(50_000, seed=3)
    ox, on = pc.voxel_downsample(pts, nrm, 1.0)
    assert 0 < len(ox) < len(pts)
    # the weighted mean of the output equals the mean of the input
    _, _, counts = po.voxel_grid(pts, nrm, 1.0)
    assert np.allclose((ox.astype(np.float64) * counts[:, None]).sum(0) / counts.sum(), pts.astype(np.float64).mean(0), atol=1e-5)
    # no normals; device tensors; strided (xyzi) input
    ox2, none = pc.voxel_downsample(pts, None, 1.0)
    assert none is None and np.array_equal(ox2, ox)
    xyzi = np.concatenate([pts, np.ones((len(pts), 1), np.float32)], 1)
    ox3, _ = pc.voxel_downsample(torch.from_numpy(xyzi).cuda(), None, 1.0)
    assert np.array_equal(ox3.cpu().numpy(), ox)
    # leaf larger than the cloud: one voxel per sign octant at most; a single point; empty input
    one, _ = pc.voxel_downsample(pts[:1], None, 1.0)
    assert np.array_equal(one, pts[:1])
    empty, _ = pc.voxel_downsample(np.zeros((0, 3), np.float32), None, 1.0)
    assert empty.shape == (0, 3)
    # a grid with more than INT_MAX cells: PCL gives the input back unchanged
    wide = np.array([[0, 0, 0], [5000, 5000, 5000], [1, 2, 3]], np.float32)
    back, _ = pc.voxel_downsample(wide, None, 0.001)
    assert np.array_equal(back, wide)
    rx, _, _ = po.voxel_grid(wide, None, 0.001)
    assert np.array_equal(rx, wide)
    # zero normals in a voxel: NaN, as Eigen's normalize() of a zero vector
    z, zn = pc.voxel_downsample(np.zeros((2, 3), np.float32), np.zeros((2, 3), np.float32), 1.0)
    assert len(z) == 1 and np.isnan(zn).all()
    # capacity smaller than the answer; bad leaf
    t = torch.from_numpy(pts).cuda()
    out = torch.full((4, 3), -7.0, device="cuda")
    cnt = C.c_int64(0)
    api._check(api.lib().pcnn_voxel_downsample(api._t_ptr(t), None, len(pts), 3, 3, 1.0, 1.0, 1.0, api._t_ptr(out), None, 3,
                                               C.byref(cnt), 0, None))
    assert cnt.value == len(ox) and np.array_equal(out[:3].cpu().numpy(), ox[:3]) and (out[3] == -7).all()
    rc = api.lib().pcnn_voxel_downsample(api._t_ptr(t), None, len(pts), 3, 3, 0.0, 1.0, 1.0, api._t_ptr(out), None, 3,
                                         C.byref(cnt), 0, None)
    assert rc == api.ERR_INVALID


def test_reader_reproduces_the_trailing_record(tmp_path):
    rec = np.arange(24, dtype=np.float32).reshape(4, 6)
    path = str(tmp_path / "cloud.bin")
    rec.tofile(path)
    xyz, nrm = pc.read_xyz_normal_bin(path)
    assert xyz.shape == (5, 3) and np.array_equal(xyz[:4], rec[:, :3]) and np.array_equal(nrm[:4], rec[:, 3:])
    assert (xyz[4] == 0).all() and (nrm[4] == 0).all()
    ox, on = pc.read_binary_and_voxel_down(path, 100.0)
    assert len(ox) == 1
