"""Developer tool (GPU box): radius r=0.5, 200k queries of the bench clouds; kernel times of the
count and fill passes (profile events) and the whole call, batch on/off."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C
import numpy as np, torch
import pcnn_b200 as pc
from pcnn_b200 import api
import bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
ix.set_option("profile", 1)
nq = int(os.environ.get("NQ", "200000"))
dq = torch.from_numpy(qry[:nq]).to(dev).contiguous()
L = pc.lib()
ref = None
cfgs = ((0, 4, 45), (2, 4, 45), (3, 4, 45), (4, 4, 45))
if os.environ.get("CFG"):  # e.g. CFG="1,4,45;0,4,45"
    cfgs = tuple(tuple(int(v) for v in c.split(",")) for c in os.environ["CFG"].split(";"))
iters = int(os.environ.get("ITERS", "4"))
for batch, div, bits in cfgs:
    ix.set_option("radius_batch", batch)
    ix.set_option("radius_run_div", div)
    ix.set_option("radius_sort_bits", bits)
    if os.environ.get("MULT"):  # "count,fill" percentages of the mean group extent (used when div = 0)
        mc, mf = (int(v) for v in os.environ["MULT"].split(","))
        ix.set_option("radius_run_mult_count", mc)
        ix.set_option("radius_run_mult_fill", mf)
    for r in [float(x) for x in os.environ.get('RADII', '0.5,0.2,1.0').split(',')]:
        best = None
        for it in range(iters):
            off = torch.empty(nq + 1, dtype=torch.int64, device=dev)
            total = C.c_int64(0)
            torch.cuda.synchronize(); t0 = time.perf_counter()
            api._check(L.pcnn_radius_count(ix._h, api._t_ptr(dq), nq, 3, float(r), api._t_ptr(off), C.byref(total), api._stream()))
            t1 = time.perf_counter(); kc = ix.last_kernel_ms()
            n = int(total.value)
            idx = torch.empty(n, dtype=torch.int32, device=dev); d = torch.empty(n, dtype=torch.float32, device=dev)
            torch.cuda.synchronize(); t2 = time.perf_counter()
            api._check(L.pcnn_radius_fill(ix._h, api._t_ptr(dq), nq, 3, float(r), api._t_ptr(off), api._t_ptr(idx), api._t_ptr(d), None, api._stream()))
            torch.cuda.synchronize(); t3 = time.perf_counter(); kf = ix.last_kernel_ms()
            cur = (1e3 * (t1 - t0), kc, 1e3 * (t3 - t2), kf)
            best = cur if best is None or cur[0] + cur[2] < best[0] + best[2] else best
        key = (r,)
        srt = None
        if r == 0.5:
            o = off.cpu().numpy(); i_ = idx.cpu().numpy()
            chk = (int(o[-1]), int(np.bitwise_xor.reduce(i_.astype(np.int64) * 2654435761 % (1 << 31))))
            if ref is None: ref = chk
            srt = "same_as_first=%s" % (chk == ref)
        alg = nq * 24 + 8 * n + 16 * len(pts)
        print("bits %d div %d " % (bits, div) + "batch %d r %.1f hits/q %.0f: count call %.3f ms (kernel %.3f)  fill call %.3f ms (kernel %.3f)  alg GB/s on kernels %.0f  %s" % (
            batch, r, n / nq, best[0], best[1], best[2], best[3], alg / ((best[1] + best[3]) * 1e-3) / 1e9, srt or ""), flush=True)
