#!/usr/bin/env python
"""bench.py -- headline benchmark of the nearest-neighbour hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Metric (BASELINE.json): kNN queries/s at k = 8.  One step = one pass of batched kNN (k=8)
over one batch of 1M synthetic KITTI-like queries against a 1M-point reference cloud
(the C2 cloud shape).  With N GPUs every rank holds the full index (cloud
broadcast from rank 0 over NCCL) and its own batch of queries: weak scaling, no data-path
collective.  `value` counts queries already resident in HBM; `e2e` goes through the C-ABI
host entry (pcnn_knn_host) with pinned host buffers, H2D and D2H inside the timed region.
Extra keys: radius r=0.5 and ICP iterations/s on the same clouds (bounded, after the
headline), roofline of the dominant kernel, CPU baseline on the box's host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

N_POINTS = 1_000_000
N_QUERIES = 1_000_000
K = 8
FRAMES = 9
WORKLOAD = ("kNN k=8, 1M-point synthetic KITTI-like aggregate (9 frames) x 1M queries per GPU "
            "(second aggregate, path offset 1 m)")


ISS_LOCAL_R, ISS_NMS_R = 0.5, 0.3  # metres, for the KITTI-like aggregate


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def make_clouds(rank):
    import pcnn_b200 as pc
    t = time.time()
    pts, nrm = pc.synth.aggregate(N_POINTS, FRAMES, seed=pc.synth.BASE_SEED + 2)
    qry, _ = pc.synth.aggregate(N_QUERIES, FRAMES, seed=pc.synth.BASE_SEED + 2, lateral=1.0,
                                noise_seed=1000 + 37 * rank)
    log("[rank %d] synthetic clouds in %.1fs" % (rank, time.time() - t))
    return pts, nrm, qry


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region: NVML polled every 5 ms from a
    thread (nvidia-smi -lms cannot go below ~100 ms, longer than a short timed region)."""
    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
           "sw_power_cap": 0x4, "hw_power_brake": 0x80}

    def __init__(self, gpu_index):
        self.sm, self.reasons, self.mx, self.ok = [], set(), None, False
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            # NVML enumerates physical devices: honour CUDA_VISIBLE_DEVICES if it is a list of ints
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            try:
                phys = [int(x) for x in vis.split(",") if x.strip() != ""]
                gpu_index = phys[gpu_index] if phys else gpu_index
            except Exception:
                pass
            self.h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
        except Exception as exc:  # pragma: no cover
            self.err = repr(exc)

    def _poll(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for name, bit in self.BAD.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def stop(self):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "err", "?")]}
        self._stop.set()
        self.th.join(timeout=1)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                "samples": len(self.sm), "reasons": sorted(self.reasons)}


# ---------------------------------------------------------------------------- CPU side
def cpu_oracle():
    """The reference's own headers compiled unmodified (oracle/_ref, kind "reference") when that
    library is present, else the restated port (kind "port").  Used ONLY as the timed CPU
    baseline / reference arm and as a checker, never on the product path."""
    import pcoracle as po
    if po.have("ref"):
        return po.Oracle("ref"), "reference"
    if not po.have("port"):
        po.build_oracle(ref=False)
    return po.Oracle("port"), "port"


def pc_synth():
    import pcnn_b200 as pc
    return pc.synth


def cpu_knn_rate(tree, qry, cores, budget_s, chunk):
    """queries/s of the reference KD-tree loop (benchmark.hpp:59-66) split over `cores` host
    threads, on as many `chunk`-query slices of the workload as fit in ~budget_s."""
    done, t0 = 0, time.perf_counter()
    while done < len(qry):
        tree.knn_checksum(qry[done:done + chunk], K, nthreads=cores)
        done += min(chunk, len(qry) - done)
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return done / dt, done, dt


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # rank 0 alone runs the CPU arm
    import pcoracle as po
    orc, kind = cpu_oracle()
    cores = orc.hardware_threads()
    pts, _, qry = make_clouds(0)
    t = time.perf_counter()
    tree = orc.build(po.KDTREE, pts, 32, 1e-4)
    build_s = time.perf_counter() - t
    sample = min(N_QUERIES, max(20_000, 25_000 * cores))  # per step; bounded
    q = qry[:sample]
    for _ in range(args.warmup):
        tree.knn_checksum(q[: max(1000, sample // 10)], K, nthreads=cores)
    times = []
    for _ in range(args.steps):
        t = time.perf_counter()
        tree.knn_checksum(q, K, nthreads=cores)
        times.append(time.perf_counter() - t)
    total = sum(times)
    value = sample * args.steps / total
    desc = ("%s KD-tree (leaf 32), first %d of the 1M queries per step, %d host threads; "
            "build %.2fs not timed" % (kind, sample, cores, build_s))
    line = {
        "impl": "reference", "metric": "kNN queries/s (k=8)", "value": value, "unit": "queries/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_points": N_POINTS, "n_queries_per_step": sample, "k": K},
        "cpu_baseline": {"value": value, "unit": "queries/s", "cores": cores, "kind": kind, "sample": desc},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline(pts, qry):
    """Reported baseline, rank 0 at N=1 only: the reference KD-tree loop on the box's host cores,
    on a bounded sample (~12 s) of the same workload."""
    import pcoracle as po
    orc, kind = cpu_oracle()
    cores = orc.hardware_threads()
    t = time.perf_counter()
    tree = orc.build(po.KDTREE, pts, 32, 1e-4)
    build_s = time.perf_counter() - t
    tree.knn_checksum(qry[:2000], K, nthreads=cores)
    rate, done, dt = cpu_knn_rate(tree, qry, cores, 12.0, max(20_000, 25_000 * cores))
    rate1, done1, dt1 = cpu_knn_rate(tree, qry, 1, 3.0, 20_000)
    out = {"value": rate, "unit": "queries/s", "cores": cores, "kind": kind,
           "sample": "first %d of the 1M queries in %.1fs on %d threads (KD-tree leaf 32, build %.2fs untimed); "
                     "1 thread: %.3g queries/s over %d queries" % (done, dt, cores, build_s, rate1, done1)}
    # the other operations of the path, same box, same clouds (bounded samples)
    try:
        t = time.perf_counter()
        oc = orc.build(po.OCTREE, pts, 32, 1e-4)
        oc_build = time.perf_counter() - t
        nqr = 20_000
        t = time.perf_counter()
        oc.radius_checksum(qry[:nqr], 0.5, nthreads=cores)
        dt_r = time.perf_counter() - t
        out["radius_r0.5"] = {"queries_per_s": nqr / dt_r, "cores": cores, "kind": kind,
                              "sample": "octree (leaf 32, min_extent 1e-4), first %d queries, build %.2fs untimed" % (nqr, oc_build)}
        src, tar, tnrm, _ = pc_synth().icp_pair(100_000, seed=pc_synth().BASE_SEED + 3)
        for mode, name in ((0, "icp_p2p_100k"), (1, "icp_p2plane_100k")):
            t = time.perf_counter()
            r = orc.icp(src, tar, tnrm, max_iter=10, mode=mode, accum=1, nthreads=cores)
            dt_i = time.perf_counter() - t
            out[name] = {"iters_per_s": r["iters"] / dt_i, "cores": cores, "kind": "port",
                         "sample": "restated loop (nanoflann 1-NN over %d threads), %d iterations incl. one kd-tree build" % (cores, r["iters"])}
        # ISS keypoints: the restated detector on a 100k-point slice of the cloud
        port = po.Oracle("port")
        sub = np.ascontiguousarray(pts[:100_000, :3])
        t = time.perf_counter()
        kp, _ = port.iss(sub, ISS_LOCAL_R, ISS_NMS_R, 0.9, 0.9, 5, True, nthreads=cores)
        dt_s = time.perf_counter() - t
        out["iss_keypoints"] = {"points_per_s": len(sub) / dt_s, "cores": cores, "kind": "port",
                                "sample": "oracle/port_iss.cpp (grid neighbourhoods) on the first 100k points, "
                                          "%d keypoints" % len(kp)}
        rng = np.random.default_rng(5)
        ks = rng.uniform(-10, 10, (3000, 3)).astype(np.float32)
        kt = (ks @ pc_synth().rot_zx(25.0, 7.0).astype(np.float32).T + np.float32(1.0)).astype(np.float32)
        kp = np.stack([np.arange(3000), np.arange(3000)], 1).astype(np.int32)
        bad = rng.random(3000) < 0.7
        kp[bad, 1] = rng.integers(0, 3000, bad.sum())
        t = time.perf_counter()
        port.ransac(ks, kt, kp, 8000, 0.1, 0.15, 1, nthreads=cores)
        dt_k = time.perf_counter() - t
        out["ransac_hyp_3k_pairs"] = {"hypotheses_per_s": 8000 / dt_k, "cores": cores, "kind": "port",
                                      "sample": "oracle/port_ransac.cpp, 8,000 of the 80,000 hypotheses over 3,000 correspondences"}
    except Exception as exc:  # a baseline leg must never sink the bench line
        out["extra_error"] = repr(exc)
    return out


def extras(ix, pc, torch, dev, pts, nrm, qry, d_q):
    """Other configs of BASELINE.json on the same clouds, bounded; reported beside the headline."""
    out = {}
    peak, _ = peaks()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # C2: radius r = 0.5, both passes timed, all 1M queries against the 1M cloud
    nq = N_QUERIES
    w = ix.radius(d_q[:nq], 0.5, dist="f32")  # same sizes as the timed call: no allocator growth inside it
    del w
    torch.cuda.synchronize()
    ev0.record()
    off, idx, dist = ix.radius(d_q[:nq], 0.5, dist="f32")
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    hits = int(idx.shape[0])
    alg = nq * 24 + 8 * hits + 16 * N_POINTS
    out["radius_r0.5"] = {"queries_per_s": nq / (ms * 1e-3), "n_queries": nq, "hits_per_query": hits / nq,
                          "ms": ms, "algorithmic_GBps": alg / (ms * 1e-3) / 1e9,
                          "hbm_frac": alg / (ms * 1e-3) / 1e9 / peak, "passes": "count + scan + fill, all timed"}
    del off, idx, dist
    # ISS keypoints of the whole 1M cloud: three fused passes + compaction, one host sync
    ix.iss_keypoints(ISS_LOCAL_R, ISS_NMS_R, device_out=True)
    torch.cuda.synchronize()
    ev0.record()
    kp, lam = ix.iss_keypoints(ISS_LOCAL_R, ISS_NMS_R, device_out=True)
    ev1.record()
    torch.cuda.synchronize()
    out["iss_keypoints_1M"] = {"points_per_s": N_POINTS / (ev0.elapsed_time(ev1) * 1e-3), "ms": ev0.elapsed_time(ev1),
                               "keypoints": int(kp.shape[0]), "candidates": int((lam >= 0).sum().item()),
                               "params": {"local_radius": ISS_LOCAL_R, "non_max_radius": ISS_NMS_R, "gamma21": 0.9,
                                          "gamma32": 0.9, "min_neighbors": 5, "weighted": True}}
    del kp, lam
    # global initialisation: 33-D mutual matching of 5k x 5k descriptors, then the reference's 80,000
    # RANSAC hypotheses (main.cpp:86) over 3,000 correspondences with 70 % outliers
    rng = np.random.default_rng(5)
    sd = torch.from_numpy((rng.random((5000, 33)) * 100).astype(np.float32)).to(dev)
    td = sd[torch.randperm(5000, device=dev)] + 0.5 * torch.randn((5000, 33), device=dev)
    pc.match_descriptors(sd, td, "inter", 0.2)
    torch.cuda.synchronize()
    ev0.record()
    mm = pc.match_descriptors(sd, td, "inter", 0.2)
    ev1.record()
    torch.cuda.synchronize()
    out["match_5k_x_5k_33d"] = {"ms": ev0.elapsed_time(ev1), "pairs": int(mm.shape[0])}
    ks = rng.uniform(-10, 10, (3000, 3)).astype(np.float32)
    Rk = pc.synth.rot_zx(25.0, 7.0).astype(np.float32)
    kt = (ks @ Rk.T + np.array([1.5, -2.0, 0.7], np.float32) + rng.normal(0, 0.01, (3000, 3))).astype(np.float32)
    kp = np.stack([np.arange(3000), np.arange(3000)], 1).astype(np.int32)
    bad = rng.random(3000) < 0.7
    kp[bad, 1] = rng.integers(0, 3000, bad.sum())
    d_ks, d_kt, d_kp = torch.from_numpy(ks).to(dev), torch.from_numpy(kt).to(dev), torch.from_numpy(kp).to(dev)
    pc.ransac(d_ks, d_kt, d_kp, 80000, 0.1, seed=1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rr = pc.ransac(d_ks, d_kt, d_kp, 80000, 0.1, seed=1)
    dt = time.perf_counter() - t0
    out["ransac_80k_hyp_3k_pairs"] = {"ms": 1e3 * dt, "hypotheses_per_s": 80000 / dt, "consensus": rr["consensus"],
                                      "rot_err_vs_truth": float(np.abs(rr["R"] - Rk).max())}
    # normals by PCA over the 10 nearest neighbours within 1 m, all 1M points (pca_normal.py:93's loop)
    ix.estimate_normals(10, 1.0, device_out=True)
    torch.cuda.synchronize()
    ev0.record()
    nn_ = ix.estimate_normals(10, 1.0, device_out=True)
    ev1.record()
    torch.cuda.synchronize()
    out["pca_normals_1M"] = {"points_per_s": N_POINTS / (ev0.elapsed_time(ev1) * 1e-3), "ms": ev0.elapsed_time(ev1),
                             "max_nn": 10, "radius": 1.0,
                             "median_abs_dot_with_generator_normals": float(
                                 (nn_ * torch.from_numpy(nrm).to(dev)).sum(1).abs().median().item())}
    del nn_
    # voxel-grid downsample of the 1M cloud with normals, leaf 0.5 m (HBM-bound byte work)
    d_p, d_n = torch.from_numpy(pts).to(dev), torch.from_numpy(nrm).to(dev)
    pc.voxel_downsample(d_p, d_n, 0.5)
    torch.cuda.synchronize()
    ev0.record()
    vx, vn = pc.voxel_downsample(d_p, d_n, 0.5)
    ev1.record()
    torch.cuda.synchronize()
    vms = ev0.elapsed_time(ev1)
    # algorithmic bytes: 24 B read per point for the means + 12 B for the keys, 24 B written per voxel
    valg = 36 * N_POINTS + 24 * int(vx.shape[0])
    out["voxel_grid_1M"] = {"points_per_s": N_POINTS / (vms * 1e-3), "ms": vms, "voxels": int(vx.shape[0]), "leaf": 0.5,
                            "algorithmic_GBps": valg / (vms * 1e-3) / 1e9, "hbm_frac": valg / (vms * 1e-3) / 1e9 / peak,
                            "timed": "whole call: bounding box + keys + radix sort + scan + means, two host syncs"}
    del d_p, d_n, vx, vn
    # k = 32 (lock-step heap kernel), all 1M queries
    ix.knn(d_q[:1000], 32, dist="f32")
    torch.cuda.synchronize()
    ev0.record()
    ix.knn(d_q[:nq], 32, dist="f32")
    ev1.record()
    torch.cuda.synchronize()
    out["knn_k32"] = {"queries_per_s": nq / (ev0.elapsed_time(ev1) * 1e-3), "n_queries": nq}
    # C5's per-GPU share: k = 32 against the 10M-point map (84 frames, 2 m apart), 8 frames of 120k queries
    t0 = time.perf_counter()
    world_ = pc.synth.World(pc.synth.BASE_SEED + 5)
    big, _ = pc.synth.aggregate(10_000_000, 84, seed=pc.synth.BASE_SEED + 5, world=world_)
    frames = [pc.synth.lidar_frame(120_000, seed=pc.synth.BASE_SEED + 5 + 1000 + f, origin=(20.0 * f + 7.0, 0.5, 0.0),
                                   world=world_)[0] for f in range(8)]
    gen_s = time.perf_counter() - t0
    bix = pc.Index(torch.from_numpy(big).to(dev))
    d_f = torch.from_numpy(np.concatenate(frames)).to(dev)
    o_i = torch.empty((d_f.shape[0], 32), dtype=torch.int32, device=dev)
    o_d = torch.empty((d_f.shape[0], 32), dtype=torch.float32, device=dev)
    bix.knn(d_f, 32, out=(o_i, o_d))
    torch.cuda.synchronize()
    ev0.record()
    bix.knn(d_f, 32, out=(o_i, o_d))
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    nqf = int(d_f.shape[0])
    alg5 = nqf * (16 + 8 * 32) + 16 * 10_000_000
    out["c5_knn_k32_10M_map"] = {"queries_per_s": nqf / (ms * 1e-3), "ms": ms, "n_queries": nqf, "n_points": 10_000_000,
                                 "index_build_ms": bix.info()["build_ms"], "index_bytes": bix.info()["device_bytes"],
                                 "algorithmic_GBps": alg5 / (ms * 1e-3) / 1e9, "hbm_frac": alg5 / (ms * 1e-3) / 1e9 / peak,
                                 "host_generation_s": gen_s,
                                 "note": "one GPU's share of config C5 (8 of the 64 frames); the map is replicated per GPU"}
    bix.close()
    del big, d_f, o_i, o_d
    # C3: point-to-point ICP, 30 iterations, 100k-point pair
    src, tar, tnrm, T = pc.synth.icp_pair(100_000, seed=pc.synth.BASE_SEED + 3)
    tix = pc.Index(tar)
    d_src, d_nrm = torch.from_numpy(src).to(dev), torch.from_numpy(tnrm).to(dev)
    for mode, name in ((0, "icp_p2p_100k"), (1, "icp_p2plane_100k")):
        tix.icp(d_src, tar_normals=d_nrm, max_iter=3, mode=mode)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = tix.icp(d_src, tar_normals=d_nrm, max_iter=30, mode=mode)
        dt = time.perf_counter() - t0
        out[name] = {"iters_per_s": r["iters"] / dt, "iters": r["iters"], "ms_total": 1e3 * dt,
                     "loop_iters_per_s": r["iters"] / (r["loop_ms"] * 1e-3) if r["loop_ms"] > 0 else None,
                     "loop_us_per_iter": 1e3 * r["loop_ms"] / max(r["iters"], 1),
                     "rot_err_vs_truth": float(np.abs(r["T"][:3, :3] - T[:3, :3]).max()),
                     "trans_err_vs_truth": float(np.abs(r["T"][:3, 3] - T[:3, 3]).max()),
                     "timed": "whole call: source ordering + 30 fused iterations + one result read-back"}
    tix.close()
    # C4 shape on one GPU: point-to-plane, 1M source x 1M target
    src1m = ((qry.astype(np.float64) - pc.synth.KNOWN_T) @ pc.synth.KNOWN_R).astype(np.float32)
    d_src, d_nrm = torch.from_numpy(src1m).to(dev), torch.from_numpy(nrm).to(dev)
    ix.icp(d_src, tar_normals=d_nrm, max_iter=2, mode=1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    r = ix.icp(d_src, tar_normals=d_nrm, max_iter=30, mode=1)
    dt = time.perf_counter() - t0
    out["icp_p2plane_1M"] = {"iters_per_s": r["iters"] / dt, "iters": r["iters"], "ms_total": 1e3 * dt,
                             "loop_us_per_iter": 1e3 * r["loop_ms"] / max(r["iters"], 1),
                             "pairs": r["pairs"], "status": r["status"]}
    return out


def icp_sharded(ix, pc, torch, dist, dev, nrm, qry, world, rank):
    """C4: point-to-plane ICP, 1M x 1M, source sharded over the ranks.  Two forms, both without
    host synchronisation inside the loop: the fused one (pcnn_icp_step_peer: one kernel per
    iteration does accumulate + all-reduce through peer memory + replicated solve) and the
    split-phase one with NCCL (accumulate kernel, all-reduce of the 31 sums + a 1-element MAX,
    update kernel)."""
    n = N_QUERIES
    d_src = torch.empty((n, 3), dtype=torch.float32, device=dev)
    d_nrm = torch.empty((N_POINTS, 3), dtype=torch.float32, device=dev)
    if rank == 0:
        src = ((qry.astype(np.float64) - pc.synth.KNOWN_T) @ pc.synth.KNOWN_R).astype(np.float32)
        d_src.copy_(torch.from_numpy(src))
        d_nrm.copy_(torch.from_numpy(nrm))
    dist.broadcast(d_src, 0)
    dist.broadcast(d_nrm, 0)
    lo, hi = pc.shard_range(n, world, rank)

    def run(iters, fused):
        st = pc.IcpState(ix, d_src[lo:hi].contiguous(), tar_normals=d_nrm, max_iter=iters, mode=1,
                         src_index_base=lo)
        if fused:
            st.attach_peers(dist)
        dist.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(iters):
            if fused:
                st.step_peer()           # ONE kernel: accumulate + peer-mailbox all-reduce + solve
            else:
                st.accumulate()
                pc.allreduce_icp_sums(st.sums, dist)
                st.update()
        ev1.record()
        torch.cuda.synchronize()
        t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res = st.finish()
        st.close()
        return res, float(t.item())

    out = {"n_gpus": world}
    for fused, name in ((True, "fused_peer"), (False, "nccl_split_phase")):
        try:
            run(2, fused)                # warm-up
            res, ms = run(30, fused)
        except pc.PcnnError as exc:      # e.g. no CUDA IPC between the ranks: keep the NCCL form
            out[name] = {"error": str(exc)}
            continue
        out[name] = {"iters_per_s": res["iters"] / (ms * 1e-3), "iters": res["iters"], "ms_total": ms,
                     "us_per_iter": 1e3 * ms / max(res["iters"], 1), "pairs": res["pairs"], "status": res["status"],
                     "T": [float(x) for x in res["T"].reshape(-1)]}
    out["iters_per_s"] = out["fused_peer"].get("iters_per_s", out["nccl_split_phase"].get("iters_per_s"))
    out["collective"] = ("fused_peer: the last block of each rank stores its 32 sums into every peer's mailbox over "
                         "NVLink (P2P stores + system-scope flag) and adds the world's in rank order, inside the "
                         "iteration kernel; nccl_split_phase: accumulate kernel, NCCL all-reduce of 31 doubles (SUM) "
                         "+ 1 int64 (MAX), update kernel")
    return out


# ---------------------------------------------------------------------------- GPU side
def run_gpu(args):
    import torch
    import torch.distributed as dist
    import pcnn_b200 as pc

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or pc.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- inputs: reference cloud from rank 0 (broadcast over NCCL), own query batch per rank
    import pcnn_b200 as pc  # noqa: F811
    if rank == 0:
        pts, nrm, qry = make_clouds(0)
        d_pts = torch.from_numpy(pts).to(dev)
    else:
        qry, _ = pc.synth.aggregate(N_QUERIES, FRAMES, seed=pc.synth.BASE_SEED + 2, lateral=1.0,
                                    noise_seed=1000 + 37 * rank)
        pts = nrm = None
        d_pts = torch.empty((N_POINTS, 3), dtype=torch.float32, device=dev)
    if world > 1:
        dist.broadcast(d_pts, 0)
    pc.Index(d_pts, kind=pc.KDTREE, leaf_size=32).close()   # first build pays module load + pool growth
    ix = pc.Index(d_pts, kind=pc.KDTREE, leaf_size=32)
    info = ix.info()
    ix.set_option("profile", 1)
    d_q = torch.from_numpy(qry).to(dev)
    o_idx = torch.empty((N_QUERIES, K), dtype=torch.int32, device=dev)
    o_dist = torch.empty((N_QUERIES, K), dtype=torch.float32, device=dev)
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)  # > 126 MB L2

    def step():
        ix.knn(d_q, K, out=(o_idx, o_dist))

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kern_ms = []
    for i in range(args.steps):
        flush.zero_()                       # L2 flush between timed iterations (not timed)
        ev[i][0].record()
        step()
        ev[i][1].record()
        kern_ms.append(ix.last_kernel_ms())  # waits for this step's kernel
    barrier()
    clocks = sampler.stop()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = max_over_ranks(sum(step_ms))
    value = world * N_QUERIES * args.steps / (total_ms * 1e-3)

    # ---- end to end through the C-ABI host entry: pinned host buffers, H2D + D2H timed
    h_q = torch.from_numpy(qry).pin_memory().numpy()
    h_idx = torch.empty((N_QUERIES, K), dtype=torch.int32).pin_memory().numpy()
    h_dist = torch.empty((N_QUERIES, K), dtype=torch.float32).pin_memory().numpy()
    ix.knn(h_q, K, out=(h_idx, h_dist))
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ix.knn(h_q, K, out=(h_idx, h_dist))
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e = {"value": world * N_QUERIES * args.steps / e2e_s, "unit": "queries/s",
           "h2d_bytes_per_step": int(h_q.nbytes), "d2h_bytes_per_step": int(h_idx.nbytes + h_dist.nbytes),
           "ms_per_step": 1e3 * e2e_s / args.steps,
           "api": "pcnn_knn_host (C ABI), pinned host buffers, 8 stages on 8 streams"}
    assert np.array_equal(h_idx, o_idx.cpu().numpy()), "host and device entry points disagree"

    # ---- roofline of the dominant kernel (knn_dfs_kernel<8>)
    peak, peak_src = peaks()
    alg_bytes = N_QUERIES * (16 + 8 * K) + 16 * N_POINTS   # Q*(16+8k) + 16*N
    k_ms = float(np.mean(kern_ms))
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    traffic = diag = None
    tpath = os.path.join(ROOT, "profiles", "knn_kernel_traffic.json")
    if os.path.exists(tpath):
        try:
            prof = json.load(open(tpath))
            traffic = prof.get("dram_bytes_per_launch")
            diag = prof.get("candidate_traffic")   # from the committed ncu capture, not measured in this run
        except Exception:
            traffic = diag = None
    roofline = {"bound": "hbm", "kernel": "knn_dfs_kernel<8>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": k_ms,
                "kernel_share_of_step": k_ms / (sum(step_ms) / args.steps),
                "note": "kNN is not HBM-bound: candidates are served from L2/L1 (SURVEY 8d, DESIGN.md)",
                "candidate_traffic_from_ncu_capture": diag}

    line = {
        "metric": "kNN queries/s (k=8)", "value": value, "unit": "queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_points": N_POINTS, "n_queries_per_gpu": N_QUERIES, "k": K,
                   "l2": "flushed between timed steps (256 MiB write)", "parallelism": "queries sharded x%d, index replicated" % world,
                   "index_build_ms": info["build_ms"], "index_bytes": info["device_bytes"]},
        "clocks": clocks, "e2e": e2e, "roofline": roofline,
        "gpu_launches": args.steps * 14,  # per step: 1 query-code + 4x3 radix-sort + 1 search kernel
    }
    sharded = icp_sharded(ix, pc, torch, dist, dev, nrm, qry, world, rank) if world > 1 else None
    if rank == 0:
        line["extra"] = {} if args.no_extras else extras(ix, pc, torch, dev, pts, nrm, qry, d_q)
        if sharded:
            line["extra"]["icp_p2plane_1M_sharded"] = sharded
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(pts, qry)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="headline only (short runs under ncu)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: relaunch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
