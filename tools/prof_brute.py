"""Developer tool (GPU box): the brute-force tile path (method 3: cloud streamed through shared
memory by bulk async copies, exact double key on every candidate) on a 32k-point cloud x 200k
queries, k = 8, for ncu (-k regex:knn_brute).  Prints the kernel time and candidates/s."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
dev = torch.device("cuda", 0)
pts, _ = pc.synth.lidar_frame(32768, seed=pc.synth.BASE_SEED + 1)
qry, _ = pc.synth.lidar_frame(120000, seed=pc.synth.BASE_SEED + 1, noise_seed=5)
qry = np.concatenate([qry, qry[:80000] + np.float32(0.01)])
ix = pc.Index(torch.from_numpy(pts).to(dev))
ix.set_option("profile", 1)
dq = torch.from_numpy(qry).to(dev)
ix.set_option("method", 7)
ref, _ = ix.knn(dq, 8, dist="f64")
ix.set_option("method", 3)
for _ in range(3):
    idx, d = ix.knn(dq, 8, dist="f64")
torch.cuda.synchronize()
ms = ix.last_kernel_ms()
print("brute kernel %.3f ms, %.3g candidate evaluations/s (3 DMUL + 2 DADD + 3 DADD(sub) each), equal to method 7: %s" % (
    ms, len(qry) * len(pts) / (ms * 1e-3), bool(torch.equal(idx, ref))))
