"""Parity helpers shared by the CPU and GPU tests.

Contract (include/pcnn.h): kNN rows are the k smallest under (d2 in double,
original index).  The reference breaks an exact-distance tie by visit order
(resultSet.hpp:69-90), so against the reference's own rows we require
  - the distance rows to be bit-identical, and
  - the index rows to agree wherever the reference's distance at that slot is unique in the
    row and differs from the k-th (boundary) distance; inside a tie group only membership of
    an equally distant point is required.
Against the canonical brute-force answer (pcport_brute_knn) everything must match exactly.
"""
import numpy as np


def ref_d2(db, q, idx):
    """d2 = ((dx*dx)+dy*dy)+dz*dz in double, per (query row, neighbour) pair."""
    p = db[idx].astype(np.float64)
    qq = q.astype(np.float64)[:, None, :] if idx.ndim == 2 else q.astype(np.float64)
    d = p - qq
    s = d[..., 0] * d[..., 0]
    s = s + d[..., 1] * d[..., 1]
    s = s + d[..., 2] * d[..., 2]
    return s


def assert_knn_matches_reference(idx, dist, ref_idx, ref_dist, db, q):
    """idx/dist: ours (canonical ties); ref_*: rows exactly as the reference's KNNResultSet."""
    idx, ref_idx = np.asarray(idx), np.asarray(ref_idx)
    dist, ref_dist = np.asarray(dist, np.float64), np.asarray(ref_dist, np.float64)
    assert idx.shape == ref_idx.shape
    assert np.array_equal(dist, ref_dist), "distance rows differ from the reference"
    diff = idx != ref_idx
    n_tie_rows = 0
    if diff.any():
        rows = np.flatnonzero(diff.any(axis=1))
        n_tie_rows = len(rows)
        for r in rows:
            for j in np.flatnonzero(diff[r]):
                d = dist[r, j]
                same = np.isclose(dist[r], d, rtol=0, atol=0)
                assert same.sum() >= 2 or j == idx.shape[1] - 1, \
                    "row %d slot %d: index differs without a distance tie" % (r, j)
                # our pick must really be at that distance from the query
                ours = np.sqrt(ref_d2(db, q[r:r + 1], idx[r:r + 1, j:j + 1]))[0, 0]
                assert ours == d, "row %d slot %d: our index is not at the tied distance" % (r, j)
    return n_tie_rows


def assert_knn_equals_canonical(idx, dist, can_idx, can_d2):
    assert np.array_equal(np.asarray(idx), can_idx), "index rows differ from the canonical answer"
    want = np.sqrt(can_d2)
    assert np.array_equal(np.asarray(dist, np.float64), want), "distances differ from sqrt(d2)"


def csr_rows_sorted(off, idx, dist=None):
    """Sort every CSR row by index; returns (idx_sorted, dist_sorted).  One argsort of a combined
    (row, index) key: rows hold distinct indices, so the key is unique."""
    off = np.asarray(off)
    idx = np.asarray(idx)
    if idx.size == 0:
        return idx, (None if dist is None else np.asarray(dist))
    key = np.repeat(np.arange(len(off) - 1, dtype=np.int64), np.diff(off))
    key *= int(idx.max()) + 1
    key += idx
    if dist is None:
        key.sort()
        key %= int(idx.max()) + 1
        return key.astype(idx.dtype), None
    order = np.argsort(key)
    return idx[order], np.asarray(dist)[order]


def assert_radius_matches_reference(off, idx, dist, ref_off, ref_idx, ref_dist):
    assert np.array_equal(np.asarray(off), np.asarray(ref_off)), "hit counts differ"
    a, ad = csr_rows_sorted(off, idx, dist)
    b, bd = csr_rows_sorted(ref_off, ref_idx, ref_dist)
    assert np.array_equal(a, b), "hit sets differ"
    if dist is not None:
        assert np.array_equal(np.asarray(ad, np.float64), np.asarray(bd, np.float64)), "hit distances differ"


def rot_err(Ta, Tb):
    return float(np.abs(np.asarray(Ta)[:3, :3] - np.asarray(Tb)[:3, :3]).max())


def trans_err(Ta, Tb):
    return float(np.abs(np.asarray(Ta)[:3, 3] - np.asarray(Tb)[:3, 3]).max())


def try_option(ix, name, value):
    """set_option that reports False for kernel variants the shipped library does not hold (they
    exist only in the -DPCNN_STUDY build, `make -C .../csrc STUDY=1`, run with PCNN_LIB=...)."""
    import pcnn_b200 as pc
    try:
        ix.set_option(name, value)
        return True
    except pc.PcnnError as exc:
        if exc.code == pc.api.ERR_UNSUPPORTED:
            return False
        raise
