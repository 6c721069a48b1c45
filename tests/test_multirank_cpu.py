"""World-size-2 tests of the N > 1 host logic on CPU (gloo): query sharding tiles the batch,
and the ICP all-reduce (31 sums + 1 int64 max) gives every rank the totals of the whole
source cloud -- checked against the oracle's single-process sums of the same pairs."""
import os
import socket

import numpy as np
import pytest

import pcnn_b200 as pc


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _partial_sums(src, tar, nn_idx, d2, lo, hi, gate):
    """What one rank's pcnn_icp_accumulate contributes for point-to-point (layout of
    pcnn_icp.cu: [0..2] sum p, [3..5] sum q, [6..14] sum q_r * p_c, [28] count, [31] packed)."""
    s = np.zeros(32, np.float64)
    keep = np.flatnonzero(d2[lo:hi] < gate) + lo
    p = src[keep].astype(np.float64)
    q = tar[nn_idx[keep]].astype(np.float64)
    s[0:3] = p.sum(0)
    s[3:6] = q.sum(0)
    s[6:15] = (q[:, :, None] * p[:, None, :]).sum(0).reshape(9)
    s[28] = len(keep)
    packed = -1
    if len(keep):
        last = int(keep.max())
        packed = (last << 32) | int(np.float32(d2[last]).view(np.uint32))
    s[31:32] = np.array([packed], np.int64).view(np.float64)
    return s


def _worker(rank, world, port, src, tar, nn_idx, d2, gate, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = pc.shard_range(len(src), world, rank)
    sums = torch.from_numpy(_partial_sums(src, tar, nn_idx, d2, lo, hi, gate))
    pc.allreduce_icp_sums(sums, dist)
    out[rank] = (lo, hi, sums.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_tiles_the_batch():
    for n in (0, 1, 7, 120000, 1000003):
        for world in (1, 2, 3, 8):
            blocks = [pc.shard_range(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        pc.shard_range(10, 2, 2)


def test_icp_allreduce_world2_gloo(oracle_port):
    import torch.multiprocessing as mp
    src, tar, nrm, T = pc.synth.icp_pair(6000, seed=pc.synth.BASE_SEED + 5)
    nn_idx, d2 = oracle_port.nn1(tar, src)
    gate = 1.0
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), src, tar, nn_idx, d2, gate, out), nprocs=world, join=True)
    whole = _partial_sums(src, tar, nn_idx, d2, 0, len(src), gate)
    assert out[0][0] == 0 and out[0][1] == out[1][0] and out[1][1] == len(src)
    for r in range(world):
        got = out[r][2]
        assert np.allclose(got[:31], whole[:31], rtol=1e-12, atol=1e-9)
        assert got[31:].view(np.int64)[0] == whole[31:].view(np.int64)[0]  # the LAST kept pair wins
    assert np.array_equal(out[0][2], out[1][2])  # replicated update sees identical sums
