"""Developer tool (GPU box): stage timeline of pcnn_knn_host on the bench workload
(run with PCNN_PIPE_TRACE=1; PCNN_HOST_CHUNK=<queries per stage> to vary the stage size)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench

pts, nrm, qry = bench.make_clouds(0)
ix = pc.Index(torch.from_numpy(pts).to("cuda:0"))
K = 8
h_q = torch.from_numpy(qry).pin_memory().numpy()
h_idx = torch.empty((len(qry), K), dtype=torch.int32).pin_memory().numpy()
h_dist = torch.empty((len(qry), K), dtype=torch.float32).pin_memory().numpy()
for it in range(int(os.environ.get("ITERS", "4"))):
    t0 = time.perf_counter()
    ix.knn(h_q, K, out=(h_idx, h_dist))
    print("call %d: %.3f ms" % (it, 1e3 * (time.perf_counter() - t0)), file=sys.stderr, flush=True)
chk = int((torch.from_numpy(h_idx).to(torch.int64) * torch.arange(1, K + 1)).sum().item()) ^ int(torch.from_numpy(h_dist).view(torch.int32).to(torch.int64).sum().item())
print("checksum %x" % (chk & 0xffffffffffff), file=sys.stderr, flush=True)
