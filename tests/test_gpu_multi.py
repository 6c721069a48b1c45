"""2-GPU tests (skipped on a one-GPU box): sharded ICP with the all-reduce done INSIDE the
iteration kernel through peer mailboxes mapped over CUDA IPC (pcnn_icp_step_peer), one process
per GPU, against the NCCL split-phase loop and the one-GPU result."""
import os
import socket

import numpy as np
import pytest

import parity
import pcnn_b200 as pc

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, src, tar, nrm, mode, iters, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ix = pc.Index(torch.from_numpy(tar).to(dev))
    dn = torch.from_numpy(nrm).to(dev)
    lo, hi = pc.shard_range(len(src), world, rank)
    d_src = torch.from_numpy(src[lo:hi]).to(dev)
    # fused: one kernel per iteration, no collective call
    st = pc.IcpState(ix, d_src, tar_normals=dn, max_iter=iters, mode=mode, src_index_base=lo)
    st.attach_peers(dist)
    dist.barrier()
    for _ in range(iters):
        st.step_peer()
    peer = st.finish()
    st.close()
    # split phase with NCCL all-reduce
    st = pc.IcpState(ix, d_src, tar_normals=dn, max_iter=iters, mode=mode, src_index_base=lo)
    for _ in range(iters):
        st.accumulate()
        pc.allreduce_icp_sums(st.sums, dist)
        st.update()
    nccl = st.finish()
    st.close()
    out[rank] = (peer, nccl)
    ix.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", [0, 1])
def test_icp_peer_exchange_two_gpus(mode):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    src, tar, nrm, T = pc.synth.icp_pair(20000, seed=pc.synth.BASE_SEED + 3)
    ix = pc.Index(tar)
    one = ix.icp(src, tar_normals=nrm, max_iter=10, mode=mode)
    ix.close()
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), src, tar, nrm, mode, 10, out), nprocs=world, join=True)
    p0, n0 = out[0]
    p1, n1 = out[1]
    assert p0["status"] == 0 and p0["iters"] == 10 and p0["pairs"] == one["pairs"]
    assert np.array_equal(p0["T"], p1["T"])          # rank-ordered sums: same bits on both ranks
    assert np.array_equal(n0["T"], n1["T"])
    for got in (p0, n0):
        assert parity.rot_err(got["T"], one["T"]) < 1e-6 and parity.trans_err(got["T"], one["T"]) < 1e-6
