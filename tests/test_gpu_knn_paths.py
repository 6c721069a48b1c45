"""GPU parity tests for the code paths knn_dfs_kernel / knn_dfs_heap_kernel gained in the second half of
round 2: lists keyed by float32 with exact comparison only where floats tie, the empty boxes behind
odd tree levels, the interleaved launch order (24 strata, padded grids), the cooperative finish of a
warp's last walks.  Everything against the canonical brute force of the oracle (exact double keys, ties
by original index: Homework2/hw2/include/resultSet.hpp:65-91), rows bit-exact."""
import numpy as np
import pytest

import parity
import pcnn_b200 as pc

pytestmark = pytest.mark.gpu


def _cloud(n, seed, scale=20.0):
    rng = np.random.default_rng(seed)
    a = rng.normal(0, 1, (n, 3)).astype(np.float32) * np.float32(scale)
    a[: n // 3, 2] *= np.float32(0.01)  # a dense sheet: long walks for queries off the sheet
    return a


def _check(db, q, k, oracle_port):
    ix = pc.Index(db)
    idx, dist = ix.knn(q, k, dist="f64")
    cidx, cd2 = oracle_port.brute_knn(db, q, k, nthreads=8)
    parity.assert_knn_equals_canonical(idx, dist, cidx, cd2)
    ix.close()


@pytest.mark.parametrize("n", [8 * 5 + 3, 8 * 9, 8 * 21 + 1, 8 * 1365, 8 * 1365 + 5, 8 * 43691 + 7])
@pytest.mark.parametrize("k", [1, 8, 20])
def test_odd_levels(oracle_port, n, k):
    # group counts 6, 9, 22, 1365, 1366, 43692: odd node counts at several levels (1365 = 0b10101010101
    # is odd at every other level), i.e. pairs whose second box is the empty pad
    db = _cloud(n, n)
    rng = np.random.default_rng(k)
    q = np.concatenate([db[:: max(1, n // 300)], (db[rng.integers(0, n, 200)] + rng.normal(0, 3, (200, 3))).astype(np.float32)])
    _check(db, q, k, oracle_port)


@pytest.mark.parametrize("nq", [33, 127, 129, 128 * 24 - 1, 128 * 24, 128 * 24 + 1, 128 * 24 * 3 + 77, 64 * 24 + 1, 20000])
@pytest.mark.parametrize("k", [8, 32])
def test_batch_sizes_around_the_strata(oracle_port, nq, k):
    # the ordered batch is cut into 24 strata and the grid padded to a multiple of 24 blocks
    # (128 queries per block at k <= 16, 64 at 16 < k <= 32): every row must still be written once
    db = _cloud(30000, 7)
    rng = np.random.default_rng(nq)
    q = (db[rng.integers(0, len(db), nq)] + rng.normal(0, 0.5, (nq, 3))).astype(np.float32)
    _check(db, q, k, oracle_port)


@pytest.mark.parametrize("k", [1, 8, 16])
def test_far_out_queries_among_near_ones(oracle_port, k):
    # one or two queries per warp far from the cloud, the rest on it: their walks outlast the others by far,
    # which is what the cooperative finish takes over (<= 4 lanes left after 8 rounds)
    db = _cloud(120000, 11)
    rng = np.random.default_rng(k)
    q = (db[rng.integers(0, len(db), 4096)] + rng.normal(0, 0.05, (4096, 3))).astype(np.float32)
    far = rng.choice(4096, 160, replace=False)
    q[far] = (rng.normal(0, 1, (160, 3)) * 400.0 + 900.0).astype(np.float32)
    q[far[:20], 2] = np.float32(1e4)
    _check(db, q, k, oracle_port)


@pytest.mark.parametrize("k", [4, 8, 12, 32])
def test_lattice_ties(oracle_port, k):
    # an integer lattice: many candidates at exactly the same distance, so the float32 keys of the list tie
    # constantly and the order falls to the exact (double key, original index) comparison
    g = np.arange(-12, 13, dtype=np.float32)
    db = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    rng = np.random.default_rng(k)
    db = db[rng.permutation(len(db))]
    q = np.concatenate([db[::37], db[::53] + np.float32(0.5), (rng.integers(-12, 12, (300, 3)) + np.array([0.5, 0.0, 0.0])).astype(np.float32)])
    _check(db, q, k, oracle_port)


def test_keys_within_one_float32_ulp(oracle_port):
    # candidates whose double keys differ by less than a float32 ulp of the key: equal float keys, different
    # double keys -- the cold path has to order them by the doubles, not by the index
    rng = np.random.default_rng(5)
    base = rng.normal(0, 1, (4000, 3)).astype(np.float32) * np.float32(50.0)
    jitter = base.copy()
    jitter[:, 0] = np.nextafter(jitter[:, 0], np.float32(np.inf))  # one ulp away in x
    db = np.concatenate([base, jitter])
    db = db[rng.permutation(len(db))]
    q = (base[::5] * np.float32(1.0001) + np.float32(0.25)).astype(np.float32)
    for k in (2, 8, 24):
        _check(db, q, k, oracle_port)
