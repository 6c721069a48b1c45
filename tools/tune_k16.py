"""Developer tool (GPU box): kernel time of kNN at 8 < k <= 16 (the K = 16 instance of
knn_dfs_kernel) on the 1M x 1M bench clouds, and of PCA normals (k = 10 on the cloud itself)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench

pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
ix.set_option("profile", 1)
dq = torch.from_numpy(qry).to(dev)
for k in (10, 16):
    ms = []
    for it in range(5):
        idx, d = ix.knn(dq, k, dist="f32")
        torch.cuda.synchronize()
        ms.append(ix.last_kernel_ms())
    print("k=%d kernel %.3f ms  checksum %d" % (k, min(ms), int(idx.sum().item())), flush=True)
ms = []
for it in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = ix.estimate_normals(10, 1.0, device_out=True)
    e1.record()
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
print("normals k=10: %.3f ms" % min(ms), flush=True)
