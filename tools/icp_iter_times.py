"""Developer tool (GPU box): where an ICP iteration's time goes.  Runs the split-phase API
(pcnn_icp_accumulate = search kernel + sums/fold kernel, pcnn_icp_update = solve) with CUDA events
between the calls and prints, per iteration, searches / search+sums us / update us, for the 100k pair
(both modes) and the 1M x 1M point-to-plane pair; then the fused loop's average for comparison."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench
dev = torch.device("cuda", 0)


def study(name, ix, d_src, d_nrm, mode, iters=30):
    for pdl in (0, 1):
        ix.set_option("icp_pdl", pdl)
        ix.icp(d_src, tar_normals=d_nrm, max_iter=3, mode=mode)
        best = min(ix.icp(d_src, tar_normals=d_nrm, max_iter=iters, mode=mode)["loop_ms"] for _ in range(3))
        print("%s: icp_pdl=%d fused loop %.1f us/iter" % (name, pdl, 1e3 * best / iters), flush=True)
    ix.icp(d_src, tar_normals=d_nrm, max_iter=3, mode=mode)
    st = pc.IcpState(ix, d_src, tar_normals=d_nrm, max_iter=iters, mode=mode)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * iters + 1)]
    torch.cuda.synchronize()
    ev[0].record()
    for i in range(iters):
        st.accumulate(); ev[2 * i + 1].record()
        st.update(); ev[2 * i + 2].record()
    torch.cuda.synchronize()
    acc = [ev[2 * i].elapsed_time(ev[2 * i + 1]) * 1e3 for i in range(iters)]
    upd = [ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) * 1e3 for i in range(iters)]
    st.finish(); st.close()
    r = ix.icp(d_src, tar_normals=d_nrm, max_iter=iters, mode=mode)
    print("%s: fused loop %.1f us/iter | split: search+sums %s | update %s" % (
        name, 1e3 * r["loop_ms"] / r["iters"], " ".join("%.0f" % a for a in acc), " ".join("%.0f" % u for u in upd[:6])), flush=True)


src, tar, tnrm, T = pc.synth.icp_pair(100_000, seed=pc.synth.BASE_SEED + 3)
tix = pc.Index(tar)
d_src, d_nrm = torch.from_numpy(src).to(dev), torch.from_numpy(tnrm).to(dev)
for mode in (1, 0):
    study("100k mode %d" % mode, tix, d_src, d_nrm, mode)
tix.close()
if os.environ.get("BIG", "1") == "1":
    pts, nrm, qry = bench.make_clouds(0)
    ix = pc.Index(torch.from_numpy(pts).to(dev))
    src1m = ((qry.astype(np.float64) - pc.synth.KNOWN_T) @ pc.synth.KNOWN_R).astype(np.float32)
    study("1M p2plane", ix, torch.from_numpy(src1m).to(dev), torch.from_numpy(nrm).to(dev), 1)
