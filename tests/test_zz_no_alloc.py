"""GPU test of pcnn_radius_count_noalloc (include/pcnn.h), the allocation-free count pass added:
through the C ABI, against pcnn_radius_count, on two calls in a
row (the scratch is one file-scope device array that keeps the last call's contents).  Collected
last on purpose."""
import ctypes as C

import numpy as np
import pytest

import pcnn_b200 as pc
from pcnn_b200 import api

pytestmark = pytest.mark.gpu

MAX_Q = 262144  # PCNN_NOALLOC_MAX_QUERIES


def _count(L, ix, dq, r, noalloc, torch):
    nq = dq.shape[0]
    off = torch.full((nq + 1,), -7, dtype=torch.int64, device=dq.device)
    if noalloc:
        api._check(L.pcnn_radius_count_noalloc(ix._h, api._t_ptr(dq), nq, dq.shape[1], float(r), api._t_ptr(off), api._stream()))
    else:
        api._check(L.pcnn_radius_count(ix._h, api._t_ptr(dq), nq, dq.shape[1], float(r), api._t_ptr(off), None, api._stream()))
    torch.cuda.synchronize()
    return off.cpu().numpy()


def test_radius_count_noalloc_matches_radius_count_twice_in_a_row():
    import torch
    L = pc.lib()
    pts, _ = pc.synth.lidar_frame(60_000, seed=pc.synth.BASE_SEED + 71)
    rng = np.random.default_rng(5)
    ix = pc.Index(pts)
    try:
        # two calls in a row with DIFFERENT batches and radii: the second must not see the first's leftovers
        qa = (pts[rng.integers(0, len(pts), 50_000)] + rng.normal(0, 0.2, (50_000, 3))).astype(np.float32)
        qb = (pts[rng.integers(0, len(pts), 9_001)] + rng.normal(0, 1.0, (9_001, 3))).astype(np.float32)
        da, db = torch.from_numpy(qa).cuda(), torch.from_numpy(qb).cuda()
        want_a = _count(L, ix, da, 0.8, False, torch)
        want_b = _count(L, ix, db, 0.3, False, torch)
        assert want_a[-1] > 10 * len(qa) and want_b[-1] > 0          # not a vacuous comparison
        free0, _ = torch.cuda.mem_get_info()
        got_a = _count(L, ix, da, 0.8, True, torch)
        got_b = _count(L, ix, db, 0.3, True, torch)
        got_a2 = _count(L, ix, da, 0.8, True, torch)
        free1, _ = torch.cuda.mem_get_info()
        # integer output: bit-exact, the bar of every index / count comparison in this repository
        assert np.array_equal(got_a, want_a) and np.array_equal(got_b, want_b) and np.array_equal(got_a2, want_a)
        assert free1 >= free0 - (4 << 20)   # the three calls (and the test's own 400 kB tensors, cached by torch) took no device memory
        # float4-strided queries, an empty batch, and the stated maximum
        q4 = torch.cat([db, torch.ones((db.shape[0], 1), device="cuda")], 1).contiguous()
        assert np.array_equal(_count(L, ix, q4, 0.3, True, torch), want_b)
        assert _count(L, ix, da[:0], 0.8, True, torch).tolist() == [0]
        big = torch.from_numpy(pts[rng.integers(0, len(pts), MAX_Q)]).cuda()
        assert np.array_equal(_count(L, ix, big, 0.25, True, torch), _count(L, ix, big, 0.25, False, torch))
        # beyond the maximum: refused with a clear error, nothing enqueued
        over = torch.zeros((MAX_Q + 1, 3), dtype=torch.float32, device="cuda")
        off = torch.zeros(MAX_Q + 2, dtype=torch.int64, device="cuda")
        rc = L.pcnn_radius_count_noalloc(ix._h, api._t_ptr(over), MAX_Q + 1, 3, 0.5, api._t_ptr(off), None)
        assert rc == api.ERR_UNSUPPORTED and b"PCNN_NOALLOC_MAX_QUERIES" in L.pcnn_last_error()
        assert L.pcnn_radius_count_noalloc(ix._h, api._t_ptr(da), 10, 3, float("nan"), api._t_ptr(off), None) == api.ERR_INVALID
        # the regular entry point still agrees after the shared scratch has been used
        assert np.array_equal(_count(L, ix, da, 0.8, False, torch), want_a)
    finally:
        ix.close()
