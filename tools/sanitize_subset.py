"""Developer tool (GPU box): a reduced pass over the hot path for compute-sanitizer
(memcheck / racecheck / synccheck): small clouds, every shipped kernel family once --
index build, kNN k = 8 (lock-step walk), k = 32 (shared-memory heap), k = 64 (warp per query),
brute-force tiles, radius count + fill (stack walk per query and per run of queries), the
allocation-free count, 1-NN, ICP both modes with certificates (fused loop), the split phase, and
the three-rank peer exchange on one GPU (mailboxes, system-scope flags)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C
import numpy as np, torch
import pcnn_b200 as pc
from pcnn_b200 import api

N = int(os.environ.get("SAN_N", "6000"))
src, tar, nrm, T = pc.synth.icp_pair(N, seed=pc.synth.BASE_SEED + 3)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(tar).to(dev))
dq = torch.from_numpy(src).to(dev)
ix.set_option("nvtx", 1)  # the ranges are exercised too
for k in (8, 32, 64):
    ix.knn(dq, k, dist="f64")
ix.set_option("method", 3); ix.knn(dq[:512], 8); ix.set_option("method", 0)
for batch in (4, 5):
    ix.set_option("radius_batch", batch)
    off, idx, d = ix.radius(dq, 0.5, dist="f32")
ix.set_option("radius_batch", 1)
off2 = torch.empty(len(src) + 1, dtype=torch.int64, device=dev)
api._check(pc.lib().pcnn_radius_count_noalloc(ix._h, api._t_ptr(dq), len(src), 3, 0.5, api._t_ptr(off2), api._stream()))
assert torch.equal(off, off2)
ix.nn1(dq)
dn = torch.from_numpy(nrm).to(dev)
for mode in (0, 1):
    r = ix.icp(dq, tar_normals=dn, max_iter=12, mode=mode)
    assert r["status"] == 0
    cuts = [0, N // 3, N // 3, N]
    states = [pc.IcpState(ix, dq[cuts[i]:cuts[i + 1]].contiguous(), tar_normals=dn, max_iter=6, mode=mode, src_index_base=cuts[i]) for i in range(3)]
    ptrs = [s.export_mailbox()[1] for s in states]
    streams = [torch.cuda.Stream() for _ in states]
    torch.cuda.synchronize()
    for i, s in enumerate(states):
        s.attach_local_peers(i, ptrs)
    for _ in range(6):
        for s, stm in zip(states, streams):
            with torch.cuda.stream(stm):
                s.step_peer()
    res = []
    for s, stm in zip(states, streams):
        with torch.cuda.stream(stm):
            res.append(s.finish())
    assert all(np.array_equal(x["T"], res[0]["T"]) and x["status"] == 0 for x in res)
    for s in states:
        s.detach_peers()
    for s in states:
        s.close()
    st = pc.IcpState(ix, dq, tar_normals=dn, max_iter=4, mode=mode)
    for _ in range(4):
        st.accumulate(); st.update()
    st.finish(); st.close()
kp, lam = ix.iss_keypoints(0.5, 0.3, device_out=True)
ix.estimate_normals(10, 1.0, device_out=True)
pc.voxel_downsample(torch.from_numpy(tar).to(dev), dn, 0.5)
ix.close()
torch.cuda.synchronize()
print("sanitize_subset: done")
