"""Developer tool (GPU box): run radius r=0.5 (200k queries) and ICP (100k) once each, for ncu."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
dq = torch.from_numpy(qry[:200000]).to(dev)
for it in range(2):
    off, idx, d = ix.radius(dq, 0.5, dist="f32")
torch.cuda.synchronize()
print("hits", idx.shape[0])
src, tar, tnrm, T = pc.synth.icp_pair(100_000, seed=pc.synth.BASE_SEED + 3)
tix = pc.Index(tar)
r = tix.icp(torch.from_numpy(src).to(dev), tar_normals=torch.from_numpy(tnrm).to(dev), max_iter=6, mode=0)
print(r["loop_ms"])
