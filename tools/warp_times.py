"""Developer tool (GPU box): per-warp start/duration of the warp-coherent kNN kernel on the
bench workload -- where does the tail come from?"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ctypes as C
import numpy as np, torch
import pcnn_b200 as pc
import bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
ix.set_option("profile", 1)
ix.set_option("method", 4)
for a in sys.argv[1:]:
    k_, v_ = a.split("=")
    ix.set_option(k_, int(v_))
dq = torch.from_numpy(qry).to(dev)
nw = (len(qry) + 31) // 32
buf = torch.zeros(10 * nw, dtype=torch.int64, device=dev)
for it in range(3):
    ix.knn(dq, 8, dist="f32")
pc.lib().pcnn_debug_warp_times(ix._h, C.c_void_p(buf.data_ptr()))
ix.knn(dq, 8, dist="f32")
torch.cuda.synchronize()
print("kernel ms", ix.last_kernel_ms())
pc.lib().pcnn_debug_warp_times(ix._h, None)
allb = buf.cpu().numpy()
b = allb[:2 * nw].reshape(-1, 2)
cn = allb[2 * nw:].reshape(-1, 8)
t0 = b[:, 0] - b[:, 0].min()
dur = b[:, 1]
end_ns = t0 + dur / 1.965  # cycles -> ns at 1965 MHz (approx)
print("warps", len(dur), "dur cycles mean %.0f p50 %.0f p90 %.0f p99 %.0f p99.9 %.0f max %.0f" % (
    dur.mean(), *np.percentile(dur, [50, 90, 99, 99.9]), dur.max()))
print("kernel span ns (first start to last end): %.0f" % end_ns.max())
for frac in (0.5, 0.8, 0.9, 0.95, 0.99, 0.999):
    print("  %5.1f%% of warps ended by %.0f ns" % (100 * frac, np.quantile(end_ns, frac)))
top = np.argsort(-dur)[:15]
for w in top:
    print("  warp %6d start %8.0f ns dur %8.0f cyc (%.0f ns)  box %d scan %d ins %d rbox %d drain %d window %d" % ((w, t0[w], dur[w], dur[w] / 1.965) + tuple(cn[w, :6])))
print("mean counters: box %.1f scan %.1f ins %.1f rbox %.1f drain %.1f window %.0f" % tuple(cn[:, :6].mean(0)))
late = np.argsort(-end_ns)[:10]
print("last to finish:")
for w in late:
    print("  warp %6d start %8.0f end %8.0f dur_ns %.0f" % (w, t0[w], end_ns[w], dur[w] / 1.965))
np.save(os.path.join(ROOT, "gpurun_out", "warp_times.npy"), b)
# lock-step walks (method 7): c[0] = longest walk of the warp (steps), c[3] = steps summed over lanes,
# c[1] = group scans (warp-wide drains that ran), c[2] = insertions, c[4] = rounds
mx, sm = cn[:, 0].astype(np.float64), cn[:, 3].astype(np.float64)
ok = mx > 0
print("walk utilisation (sum of lane steps / 32 x longest lane): mean %.3f  p10 %.3f p50 %.3f p90 %.3f; mean longest %.1f mean lane %.1f" % (
    (sm[ok] / (32 * mx[ok])).mean(), *np.percentile(sm[ok] / (32 * mx[ok]), [10, 50, 90]), mx[ok].mean(), (sm[ok] / 32).mean()))
print("work-weighted utilisation: %.3f" % (sm[ok].sum() / (32 * mx[ok].sum())))
