"""Developer tool (GPU box): what query ordering buys on the bench clouds.  Device-resident kNN k = 8,
1M x 1M: whole call and search kernel with sort_queries = 1 (default), 0 on the generator's scan
order, and 0 on a random shuffle of the same queries."""
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
rng = np.random.default_rng(0)
sets = {"scan order": qry, "shuffled": qry[rng.permutation(len(qry))]}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, qq in sets.items():
    dq = torch.from_numpy(np.ascontiguousarray(qq)).to(dev)
    for sort in (1, 0):
        ix.set_option("sort_queries", sort)
        best = (1e9, 1e9)
        for _ in range(5):
            torch.cuda.synchronize(); e0.record()
            ix.knn(dq, 8, dist="f32")
            e1.record(); torch.cuda.synchronize()
            best = min(best, (e0.elapsed_time(e1), ix.last_kernel_ms()))
        print("%-10s sort_queries=%d: call %.3f ms, search kernel %.3f ms" % (name, sort, best[0], best[1]), flush=True)
ix.set_option("sort_queries", 1)
dq = torch.from_numpy(np.ascontiguousarray(sets["shuffled"])).to(dev)
dq2 = torch.from_numpy(np.ascontiguousarray(sets["scan order"])).to(dev)
for bits in (30, 27, 24, 21, 18, 15):
    ix.set_option("sort_bits", bits)
    row = []
    for d in (dq2, dq):
        best = (1e9, 1e9)
        for _ in range(5):
            torch.cuda.synchronize(); e0.record()
            ix.knn(d, 8, dist="f32")
            e1.record(); torch.cuda.synchronize()
            best = min(best, (e0.elapsed_time(e1), ix.last_kernel_ms()))
        row.append("call %.3f ms (kernel %.3f)" % best)
    print("sort_bits=%d: scan order %s | shuffled %s" % (bits, row[0], row[1]), flush=True)
