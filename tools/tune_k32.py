"""Developer tool (GPU box): k = 32 kNN, 200k queries of the bench clouds, methods 2 and 9."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
ix.set_option("profile", 1)
nq = int(os.environ.get("NQ", "200000"))
dq = torch.from_numpy(qry[:nq]).to(dev)
ref = None
for k in (32, 20):
    for m in (2, 9):
        ix.set_option("method", m)
        ms = []
        for it in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); idx, d = ix.knn(dq, k, dist="f64"); e1.record(); torch.cuda.synchronize()
            ms.append((e0.elapsed_time(e1), ix.last_kernel_ms()))
        a = (idx.cpu().numpy(), d.cpu().numpy())
        same = True if m == 2 else bool(np.array_equal(a[0], ref[0]) and np.array_equal(a[1], ref[1]))
        if m == 2: ref = a
        print("k %d method %d: step %.3f ms kernel %.3f ms -> %.1f M q/s  equal_to_method2=%s" % (
            k, m, min(x[0] for x in ms), min(x[1] for x in ms), nq / min(x[0] for x in ms) / 1e3, same), flush=True)
