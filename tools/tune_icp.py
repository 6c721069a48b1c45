"""Developer tool (GPU box): ICP iteration time (device loop, CUDA events inside pcnn_icp) on the
bench clouds, for the library build named by PCNN_LIB.  Prints the pose bits so builds can be
compared for identical results."""
import sys, os, hashlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench

dev = torch.device("cuda", 0)
pts, nrm, qry = bench.make_clouds(0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
src1m = ((qry.astype(np.float64) - pc.synth.KNOWN_T) @ pc.synth.KNOWN_R).astype(np.float32)
d_src, d_nrm = torch.from_numpy(src1m).to(dev), torch.from_numpy(nrm).to(dev)
src, tar, tnrm, T = pc.synth.icp_pair(100_000, seed=pc.synth.BASE_SEED + 3)
tix = pc.Index(tar)
d_s100, d_n100 = torch.from_numpy(src).to(dev), torch.from_numpy(tnrm).to(dev)
out = [os.environ.get("PCNN_LIB", "default").split("/")[-1], "bps=" + os.environ.get("PCNN_ICP_BLOCKS_PER_SM", "auto")]
for name, index, s, n, mode in (("p2plane_1M", ix, d_src, d_nrm, 1), ("p2p_1M", ix, d_src, d_nrm, 0),
                                ("p2plane_100k", tix, d_s100, d_n100, 1), ("p2p_100k", tix, d_s100, d_n100, 0)):
    index.icp(s, tar_normals=n, max_iter=3, mode=mode)
    best = 1e9
    for _ in range(3):
        r = index.icp(s, tar_normals=n, max_iter=30, mode=mode)
        best = min(best, 1e3 * r["loop_ms"] / r["iters"])
    out.append("%s %.1f us/it [%s]" % (name, best, hashlib.md5(r["T"].tobytes()).hexdigest()[:6]))
print("  ".join(out), flush=True)

if os.environ.get("PER_ITER"):
    for name, index, s, n, mode in (("p2plane_1M", ix, d_src, d_nrm, 1), ("p2p_100k", tix, d_s100, d_n100, 0)):
        st = pc.IcpState(index, s, tar_normals=n, max_iter=30, mode=mode)
        w = pc.IcpState(index, s, tar_normals=n, max_iter=2, mode=mode)  # loads the split-phase kernels
        w.accumulate(); w.update(); w.finish(); w.close()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(31)]
        evs[0].record()
        for i in range(30):
            st.accumulate(); st.update()
            evs[i + 1].record()
        torch.cuda.synchronize()
        st.finish(); st.close()
        print(name, "per-iteration us:", " ".join("%.0f" % (1e3 * evs[i].elapsed_time(evs[i + 1])) for i in range(30)), flush=True)

if os.environ.get("SEARCHES"):
    for name, index, s, n, mode in (("p2plane_1M", ix, d_src, d_nrm, 1), ("p2p_100k", tix, d_s100, d_n100, 0)):
        st = pc.IcpState(index, s, tar_normals=n, max_iter=30, mode=mode)
        cnt = []
        for i in range(30):
            st.accumulate(); st.update()
            cnt.append(st.finish()["searches"])
        st.close()
        print(name, "searches per iteration:", " ".join(str(c) for c in cnt), flush=True)
