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
for bits in (30, 24, 21, 18, 15):
    cfgs.append(("dfs bits%d" % bits, 7, {"sort_bits": bits}))
cfgs.append(("lockstep b60 bits24", 6, {"coherent_budget": 60, "sort_bits": 24}))
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

# e2e through the host entry for a few pipeline stage sizes (PCNN_HOST_CHUNK is read per call)
import time
ix.set_option("method", 0); ix.set_option("sort_bits", 30)
h_q = torch.from_numpy(qry).pin_memory().numpy()
h_idx = torch.empty((len(qry), K), dtype=torch.int32).pin_memory().numpy()
h_d = torch.empty((len(qry), K), dtype=torch.float32).pin_memory().numpy()
for chunk in (0, 32768, 65536, 131072, 262144, 524288, 1000000):
    if chunk: os.environ["PCNN_HOST_CHUNK"] = str(chunk)
    ix.knn(h_q, K, out=(h_idx, h_d))
    t0 = time.perf_counter()
    for _ in range(10): ix.knn(h_q, K, out=(h_idx, h_d))
    print("e2e chunk %7d: %.3f ms/step" % (chunk, (time.perf_counter() - t0) * 100), flush=True)
