"""Developer tool (GPU box): 12 point-to-plane ICP iterations, 1M source x 1M target, for ncu
(-k regex:icp_accumulate -s 8 -c 1 captures a converged iteration)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench
dev = torch.device("cuda", 0)
pts, nrm, qry = bench.make_clouds(0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
src1m = ((qry.astype(np.float64) - pc.synth.KNOWN_T) @ pc.synth.KNOWN_R).astype(np.float32)
r = ix.icp(torch.from_numpy(src1m).to(dev), tar_normals=torch.from_numpy(nrm).to(dev), max_iter=int(os.environ.get("ITERS", "12")), mode=int(os.environ.get("MODE", "1")))
print(r["loop_ms"], r["pairs"])
