"""Developer tool (GPU box): one call of each SURVEY 8(f) operation on the bench clouds, for ncu.
usage: python tools/prof_next.py   (under `ncu --set full -k regex:iss_pass|vox_|normals_|ransac_|desc_nn1 ...`)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import pcnn_b200 as pc
import bench
pts, nrm, qry = bench.make_clouds(0)
dev = torch.device("cuda", 0)
ix = pc.Index(torch.from_numpy(pts).to(dev))
for rep in range(2):
    torch.cuda.synchronize(); t = time.perf_counter()
    kp, lam = ix.iss_keypoints(0.5, 0.3, device_out=True); torch.cuda.synchronize()
    print("iss ms", 1e3 * (time.perf_counter() - t), int(kp.shape[0]))
    t = time.perf_counter(); n_ = ix.estimate_normals(10, 1.0, device_out=True); torch.cuda.synchronize()
    print("normals ms", 1e3 * (time.perf_counter() - t))
    d_p, d_n = torch.from_numpy(pts).to(dev), torch.from_numpy(nrm).to(dev)
    t = time.perf_counter(); vx, vn = pc.voxel_downsample(d_p, d_n, 0.5); torch.cuda.synchronize()
    print("voxel ms", 1e3 * (time.perf_counter() - t), int(vx.shape[0]))
    rng = np.random.default_rng(5)
    sd = torch.from_numpy((rng.random((20000, 33)) * 100).astype(np.float32)).to(dev)
    td = sd[torch.randperm(20000, device=dev)] + 0.5 * torch.randn((20000, 33), device=dev)
    t = time.perf_counter(); m = pc.match_descriptors(sd, td, "inter", 0.2); torch.cuda.synchronize()
    print("match 20k ms", 1e3 * (time.perf_counter() - t))
    ks = rng.uniform(-10, 10, (3000, 3)).astype(np.float32)
    kt = (ks @ pc.synth.rot_zx(25.0, 7.0).astype(np.float32).T + np.float32(1.0)).astype(np.float32)
    kpairs = np.stack([np.arange(3000), np.arange(3000)], 1).astype(np.int32)
    bad = rng.random(3000) < 0.7
    kpairs[bad, 1] = rng.integers(0, 3000, bad.sum())
    t = time.perf_counter(); r = pc.ransac(ks, kt, kpairs, 80000, 0.1, seed=1)
    print("ransac ms (incl. uploads)", 1e3 * (time.perf_counter() - t), r["consensus"])
