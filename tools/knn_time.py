"""Developer tool (GPU box): kernel time of the shipped batched kNN on the bench clouds (1M x 1M),
K from the environment (default 8), best of 5; prints a checksum of the rows so that two builds can be
compared bit for bit.  Small enough to run under ncu (-k regex:knn_dfs)."""
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
if os.environ.get("SORT_BITS"):
    ix.set_option("sort_bits", int(os.environ["SORT_BITS"]))
if os.environ.get("METHOD"):  # a study variant (PCNN_LIB=.../libpcnn_study.so)
    ix.set_option("method", int(os.environ["METHOD"]))
dq = torch.from_numpy(qry).to(dev)
for K in [int(x) for x in os.environ.get("KS", "8").split(",")]:
    best = 1e9
    for _ in range(int(os.environ.get("REPS", "5"))):
        idx, dist = ix.knn(dq, K, dist="f64")
        best = min(best, ix.last_kernel_ms())
    chk = int((idx.to(torch.int64) * torch.arange(1, K + 1, device=dev)).sum().item()) ^ int(dist.view(torch.int64).sum().item())
    print("k=%d kernel %.4f ms  %.1f M q/s  checksum %x" % (K, best, len(qry) / best / 1e3, chk & 0xffffffffffff), flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for pdl in ((0, 1, 0, 1) if not os.environ.get("METHOD") else ()):
    ix.set_option("pdl", pdl)
    best = 1e9
    for _ in range(8):
        torch.cuda.synchronize(); e0.record()
        ix.knn(dq, 8, dist="f32")
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("pdl=%d: whole device-resident call (ordering + search) %.4f ms, search kernel %.4f ms" % (pdl, best, ix.last_kernel_ms()), flush=True)
