"""Developer tool (GPU box): kernel time of batched kNN (k=8, 1M x 1M bench clouds) for each
traversal method / leaf size, with a bit-exact cross-check between methods."""
import sys, os, time
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
K = 8
ref = None
cfgs = [("thread", 1, {"leaf_log": 1})]
cfgs.append(("lockstep b60", 6, {"coherent_budget": 60}))
cfgs.append(("dfs", 7, {}))
cfgs.append(("refill", 10, {}))
for kv in os.environ.get("OPTS", "").split():
    a, b = kv.split("=")
    ix.set_option(a, int(b))
only = os.environ.get("TUNE_ONLY")
if only:
    cfgs = [c for c in cfgs if c[0].startswith(only)]
for name, method, opts in cfgs:
    ix.set_option("method", method)
    for k_, v in opts.items():
        ix.set_option(k_, v)
    ms = []
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        idx, d = ix.knn(dq, K, dist="f32")
        e1.record()
        torch.cuda.synchronize()
        ms.append((e0.elapsed_time(e1), ix.last_kernel_ms()))
    idx = idx.cpu().numpy()
    same = True if ref is None else bool(np.array_equal(ref, idx))
    if ref is None:
        ref = idx
    print("%-14s step %.3f ms  kernel %.3f ms  equal_to_first=%s" % (name, min(m[0] for m in ms), min(m[1] for m in ms), same), flush=True)

