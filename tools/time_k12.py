"""Developer tool (GPU box): step time of kNN at k = 10 / 12 / 16 (the K = 12 instance serves 9..12) and of the PCA normals."""
import sys; sys.path.insert(0, "/root/repo")
import torch, numpy as np, pcnn_b200 as pc, bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
dq = torch.from_numpy(qry).to(dev)
for k in (10, 12, 16):
    ix.knn(dq, k, dist="f32"); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ix.knn(dq, k, dist="f32"); e1.record(); torch.cuda.synchronize()
    print("k", k, "%.3f ms" % e0.elapsed_time(e1))
ix.estimate_normals(10, 1.0, device_out=True); torch.cuda.synchronize()
e0.record(); ix.estimate_normals(10, 1.0, device_out=True); e1.record(); torch.cuda.synchronize()
print("normals %.3f ms" % e0.elapsed_time(e1))
