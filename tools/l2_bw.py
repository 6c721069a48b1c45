"""Developer tool (GPU box): L2-resident copy bandwidth (read + write bytes), the denominator of
the candidate-traffic diagnostic in bench.py's roofline.  torch copy of a buffer that
fits L2 (2 x 16 MB), best of 20, CUDA events."""
import json, sys, torch
dev = torch.device("cuda", 0)
out = {}
for mb in (8, 16, 32):
    n = mb * 1024 * 1024 // 4
    a = torch.empty(n, dtype=torch.float32, device=dev).normal_()
    b = torch.empty_like(a)
    for _ in range(5):
        b.copy_(a)
    best = 1e9
    for _ in range(20):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            b.copy_(a)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 10)
    out["copy_%dMB" % mb] = {"ms": best, "GBps_read_plus_write": 2 * mb * 1.048576e-3 / (best * 1e-3)}
print(json.dumps(out))
