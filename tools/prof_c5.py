"""Developer tool (GPU box): config C5's per-GPU work for ncu -- kNN k = 32 of 8 frames of 120k queries
against the 10M-point synthetic map (ncu -k regex:knn_dfs_heap -s 1 -c 1 captures the second launch)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
dev = torch.device("cuda", 0)
w = pc.synth.World(pc.synth.BASE_SEED + 5)
big, _ = pc.synth.aggregate(10_000_000, 84, seed=pc.synth.BASE_SEED + 5, world=w)
frames = [pc.synth.lidar_frame(120_000, seed=pc.synth.BASE_SEED + 5 + 1000 + f, origin=(2.6 * f + 1.0, 0.5, 0.0), world=w)[0] for f in range(8)]
ix = pc.Index(torch.from_numpy(big).to(dev))
ix.set_option("profile", 1)
dq = torch.from_numpy(np.concatenate(frames)).to(dev)
for k in [int(x) for x in os.environ.get("KS", "32").split(",")]:
    best = 1e9
    for _ in range(3):
        ix.knn(dq, k, dist="f32")
        best = min(best, ix.last_kernel_ms())
    print("k=%d: %d queries vs 10M points: kernel %.3f ms, %.1f M q/s" % (k, len(dq), best, len(dq) / best / 1e3), flush=True)
