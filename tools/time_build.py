"""Developer tool (GPU box): warm index build time at 1M and 10M points (uniform slab)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, pcnn_b200 as pc
rng = np.random.default_rng(1)
for n in (1_000_000, 10_000_000):
    pts = torch.from_numpy((rng.random((n, 3), dtype=np.float32) * np.array([400, 400, 8], np.float32))).cuda()
    pc.Index(pts).close()
    ix = pc.Index(pts)
    print(n, "build %.3f ms" % ix.info()["build_ms"], flush=True)
    ix.close()
