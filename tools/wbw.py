"""Developer tool (GPU box): pure-write and read+write HBM bandwidth with plain kernels (torch fill_ /
copy_), so the radius fill pass has a measured write roof next to MEASURED_PEAKS.json's copy figure."""
import torch
dev = torch.device("cuda", 0)
n = 1 << 30
a = torch.empty(n, dtype=torch.float32, device=dev)
b = torch.empty(n, dtype=torch.float32, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, fn, nbytes in (("fill 4 GiB (write only)", lambda: a.fill_(1.0), 4 * n),
                         ("copy 4 GiB -> 4 GiB (read + write)", lambda: b.copy_(a), 8 * n),
                         ("sum 4 GiB (read only)", lambda: a.sum(), 4 * n)):
    best = 1e9
    for _ in range(6):
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("%-40s %.3f ms  %.0f GB/s" % (name, best, nbytes / best / 1e6))
