"""Developer tool (GPU box): the standalone float32 1-NN (pcnn_nn1) on the bench clouds, for comparison
with what an all-points ICP iteration spends on its searches."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
dq = torch.from_numpy(qry).to(dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, f in (("nn1 (float32 1-NN, ordering + search)", lambda: ix.nn1(dq)), ("knn k=1 (double keys, ordering + search)", lambda: ix.knn(dq, 1))):
    best = 1e9
    for _ in range(6):
        torch.cuda.synchronize(); e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("%s: %.3f ms per 1M queries" % (name, best), flush=True)
