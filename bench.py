#!/usr/bin/env python
"""bench.py -- headline benchmark of the nearest-neighbour hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Metric (BASELINE.json): kNN queries/s at k = 8.  One step = one pass of batched kNN (k=8)
over one batch of 1M synthetic KITTI-like queries against a 1M-point reference cloud
(the C2 cloud shape).  With N GPUs every rank holds the full index (cloud
broadcast from rank 0 over NCCL) and its own batch of queries: weak scaling, no data-path
collective.  `value` counts queries already resident in HBM; `e2e` goes through the C-ABI
host entry (pcnn_knn_host) with pinned host buffers, H2D and D2H inside the timed region.
The other configs of BASELINE.json (C2 radius, C3/C4 ICP, C5 k=32 on the 10M map, sharded over
the ranks) follow the headline, one compact entry each under roofline.others; with N > 1 the sharded
ICP pose is checked against the one-GPU pose and the oracle inside the run (config.parity) and a
failure makes the run exit non-zero.
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
    """--impl reference: the reference's CPU implementation of the path on the host cores, on the
    SAME config as the GPU arm: every step is the whole batch of 1M queries (benchmark.hpp:59-66
    split over all host threads)."""
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
    for _ in range(args.warmup):
        tree.knn_checksum(qry[:100_000], K, nthreads=cores)
    times = []
    for _ in range(args.steps):
        t = time.perf_counter()
        tree.knn_checksum(qry, K, nthreads=cores)
        times.append(time.perf_counter() - t)
    total = sum(times)
    value = N_QUERIES * args.steps / total
    desc = ("%s KD-tree (leaf 32), all %d queries per step, %d host threads; build %.2fs not timed"
            % (kind, N_QUERIES, cores, build_s))
    line = {
        "impl": "reference", "metric": "kNN queries/s (k=8)", "value": value, "unit": "queries/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_points": N_POINTS, "n_queries_per_gpu": N_QUERIES, "k": K},
        "cpu_baseline": {"value": value, "unit": "queries/s", "cores": cores, "kind": kind, "sample": desc},
        "e2e": {"value": value, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
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


def _timed(torch, fn, reps=1):
    """Device time of fn() in ms (CUDA events on the current stream), after one untimed call."""
    fn()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best, out = None, None
    for _ in range(reps):
        ev0.record()
        out = fn()
        ev1.record()
        torch.cuda.synchronize()
        ms = ev0.elapsed_time(ev1)
        best = ms if best is None or ms < best else best
    return best, out


def other_configs(ix, pc, torch, dev, pts, nrm, qry, d_q):
    """The other configs of BASELINE.json on one GPU, bounded; one COMPACT entry each (the driver
    keeps `roofline.others`).  C2 radius r = 0.5, C3 ICP 100k pair, C4 shape on one GPU."""
    out = []
    peak, _ = peaks()
    nq = N_QUERIES
    # C2: octree radius search r = 0.5, 1M x 1M; the whole call is timed: ordering + count + scan + fill
    ms, (off, idx, dist) = _timed(torch, lambda: ix.radius(d_q[:nq], 0.5, dist="f32"), reps=3)
    hits = int(idx.shape[0])
    alg = nq * 24 + 8 * hits + 16 * N_POINTS
    out.append({"config": "C2 radius r=0.5, 1M x 1M", "ms": round(ms, 4), "queries_per_s": nq / (ms * 1e-3),
                "hits_per_query": round(hits / nq, 1), "algorithmic_GBps": round(alg / (ms * 1e-3) / 1e9, 1),
                "hbm_frac": round(alg / (ms * 1e-3) / 1e9 / peak, 4), "bound": "hbm (output write)",
                "timed": "query ordering + count + scan + fill, incl. one host read of the total"})
    del off, idx, dist
    # C3: point-to-point (and point-to-plane) ICP, 30 iterations, 100k-point pair
    src, tar, tnrm, T = pc.synth.icp_pair(100_000, seed=pc.synth.BASE_SEED + 3)
    tix = pc.Index(tar)
    d_src, d_nrm = torch.from_numpy(src).to(dev), torch.from_numpy(tnrm).to(dev)
    for mode, name in ((0, "C3 ICP p2p 100k pair, 30 iterations"), (1, "ICP p2plane 100k pair, 30 iterations")):
        tix.icp(d_src, tar_normals=d_nrm, max_iter=3, mode=mode)
        torch.cuda.synchronize()
        r = tix.icp(d_src, tar_normals=d_nrm, max_iter=30, mode=mode)
        us = 1e3 * r["loop_ms"] / max(r["iters"], 1)
        alg_i = 16 * 100_000 * 2 + (16 * 100_000 if mode == 1 else 0)
        out.append({"config": name, "us_per_iter": round(us, 2), "iters_per_s": round(1e6 / us, 1) if us > 0 else None,
                    "iters": r["iters"], "hbm_frac": round(alg_i / (us * 1e-6) / 1e9 / peak, 5), "bound": "latency",
                    "rot_err_vs_truth": float(np.abs(r["T"][:3, :3] - T[:3, :3]).max())})
    tix.close()
    # C4 shape on one GPU: point-to-plane, 1M source x 1M target
    src1m = ((qry.astype(np.float64) - pc.synth.KNOWN_T) @ pc.synth.KNOWN_R).astype(np.float32)
    d_src, d_nrm = torch.from_numpy(src1m).to(dev), torch.from_numpy(nrm).to(dev)
    ix.icp(d_src, tar_normals=d_nrm, max_iter=2, mode=1)
    torch.cuda.synchronize()
    r = ix.icp(d_src, tar_normals=d_nrm, max_iter=30, mode=1)
    us = 1e3 * r["loop_ms"] / max(r["iters"], 1)
    out.append({"config": "C4 ICP p2plane 1M x 1M on 1 GPU, 30 iterations", "us_per_iter": round(us, 2),
                "iters_per_s": round(1e6 / us, 1) if us > 0 else None, "iters": r["iters"],
                "hbm_frac": round(48e6 / (us * 1e-6) / 1e9 / peak, 5), "bound": "latency", "pairs": r["pairs"]})
    # k = 32 on the 1M cloud (lock-step heap kernel)
    ms, _ = _timed(torch, lambda: ix.knn(d_q[:nq], 32, dist="f32"))
    out.append({"config": "kNN k=32, 1M x 1M", "ms": round(ms, 4), "queries_per_s": nq / (ms * 1e-3)})
    return out


def next_ops(ix, pc, torch, dev, pts, nrm):
    """SURVEY 8(f) operations on the same cloud, one number each (reported under `extra`)."""
    out = {}
    ms, (kp, lam) = _timed(torch, lambda: ix.iss_keypoints(ISS_LOCAL_R, ISS_NMS_R, device_out=True))
    out["iss_keypoints_1M_ms"] = round(ms, 3)
    out["iss_keypoints"] = int(kp.shape[0])
    ms, nn_ = _timed(torch, lambda: ix.estimate_normals(10, 1.0, device_out=True))
    out["pca_normals_1M_ms"] = round(ms, 3)
    d_p, d_n = torch.from_numpy(pts).to(dev), torch.from_numpy(nrm).to(dev)
    ms, (vx, vn) = _timed(torch, lambda: pc.voxel_downsample(d_p, d_n, 0.5))
    out["voxel_grid_1M_ms"] = round(ms, 3)
    out["voxel_grid_hbm_frac"] = round((36 * N_POINTS + 24 * int(vx.shape[0])) / (ms * 1e-3) / 1e9 / peaks()[0], 4)
    rng = np.random.default_rng(5)
    sd = torch.from_numpy((rng.random((5000, 33)) * 100).astype(np.float32)).to(dev)
    td = sd[torch.randperm(5000, device=dev)] + 0.5 * torch.randn((5000, 33), device=dev)
    ms, _ = _timed(torch, lambda: pc.match_descriptors(sd, td, "inter", 0.2))
    out["match_5k_x_5k_33d_ms"] = round(ms, 3)
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
    pc.ransac(d_ks, d_kt, d_kp, 80000, 0.1, seed=1)
    out["ransac_80k_hyp_ms"] = round(1e3 * (time.perf_counter() - t0), 3)
    return out


def icp_sharded(ix, pc, torch, dist, dev, pts, nrm, qry, world, rank):
    """C4: point-to-plane ICP, 1M x 1M, source sharded over the ranks.  Two forms, both without
    host synchronisation inside the loop: the fused one (pcnn_icp_step_peer: one kernel per
    iteration does accumulate + all-reduce through peer memory + replicated solve) and the
    split-phase one with NCCL (accumulate kernel, all-reduce of the 31 sums + a 1-element MAX,
    update kernel).  PARITY IS ASSERTED HERE, inside the run the driver executes: fused == NCCL bit
    for bit on every rank, every rank holds the same pose, the pose is within 1e-6 of rank 0's
    single-GPU pcnn_icp and within 1e-5 of the restated reference loop (oracle, rank 0's host cores)."""
    n = N_QUERIES
    d_src = torch.empty((n, 3), dtype=torch.float32, device=dev)
    d_nrm = torch.empty((N_POINTS, 3), dtype=torch.float32, device=dev)
    src = None
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

    out = {"config": "C4 ICP p2plane 1M x 1M, source sharded x%d, 30 iterations" % world, "n_gpus": world}
    poses = {}
    for fused, name in ((True, "fused_peer"), (False, "nccl_split_phase")):
        try:
            run(2, fused)                # warm-up
            res, ms = run(30, fused)
        except pc.PcnnError as exc:      # e.g. no CUDA IPC between the ranks: keep the NCCL form
            out[name + "_error"] = str(exc)[:120]
            continue
        poses[name] = res
        out[name + "_us_per_iter"] = round(1e3 * ms / max(res["iters"], 1), 2)
    best = min(v for k, v in out.items() if k.endswith("_us_per_iter"))
    out["us_per_iter"] = best
    out["iters_per_s"] = round(1e6 / best, 1)
    out["collective"] = "fused_peer = P2P stores + system-scope flags inside the iteration kernel; nccl_split_phase = NCCL all-reduce of 32 x 8 B"

    # ---- parity (asserted; the outcome goes into config.parity)
    parity = {}
    ref = poses.get("fused_peer") or poses.get("nccl_split_phase")
    ok_local = True
    if "fused_peer" in poses and "nccl_split_phase" in poses:
        same = (np.array_equal(poses["fused_peer"]["T"], poses["nccl_split_phase"]["T"]) and
                poses["fused_peer"]["pairs"] == poses["nccl_split_phase"]["pairs"])
        ok_local = ok_local and same
    ok_local = ok_local and ref["status"] == 0 and ref["iters"] == 30
    Ts = [torch.empty(16, dtype=torch.float32, device=dev) for _ in range(world)]
    dist.all_gather(Ts, torch.from_numpy(ref["T"].reshape(-1).copy()).to(dev))
    ok_local = ok_local and all(torch.equal(Ts[0], t) for t in Ts)      # rank-ordered sums: same bits everywhere
    flag = torch.tensor([1 if ok_local else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    parity["icp_fused_eq_nccl_all_ranks"] = "ok" if int(flag.item()) == 1 else "FAIL"
    if rank == 0:
        one = ix.icp(d_src, tar_normals=d_nrm, max_iter=30, mode=1)
        d1 = float(np.abs(one["T"] - ref["T"]).max())
        parity["icp_sharded_vs_1gpu"] = "ok" if (d1 < 1e-6 and one["pairs"] == ref["pairs"]) else "FAIL (%.2e)" % d1
        try:
            import pcoracle as po
            if not po.have("port"):
                po.build_oracle(ref=False)
            port = po.Oracle("port")
            want = port.icp(src, pts, nrm, max_iter=30, mode=1, accum=1, nthreads=port.hardware_threads())
            d2 = float(np.abs(want["T"] - ref["T"]).max())
            parity["icp_vs_oracle"] = "ok" if (d2 < 1e-5 and want["pairs"] == ref["pairs"]) else "FAIL (%.2e)" % d2
        except Exception as exc:
            parity["icp_vs_oracle"] = "FAIL (%r)" % (exc,)
    return out, parity


def knn_sample_parity(pc, torch, dist, dev, d_pts, qry, o_idx, world):
    """Every rank checks 256 of its own kNN rows against the canonical brute force (oracle port)."""
    ok = 1
    try:
        import pcoracle as po
        if not po.have("port"):
            po.build_oracle(ref=False)
        port = po.Oracle("port")
        sel = np.arange(0, N_QUERIES, N_QUERIES // 256)[:256]
        cidx, _ = port.brute_knn(d_pts.cpu().numpy(), qry[sel], K, nthreads=max(1, port.hardware_threads() // max(world, 1)))
        ok = 1 if np.array_equal(o_idx[torch.from_numpy(sel).to(dev)].cpu().numpy(), cidx) else 0
    except Exception as exc:
        log("knn_sample_parity: %r" % (exc,))
        ok = 0
    flag = torch.tensor([ok], dtype=torch.int32, device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return "ok" if int(flag.item()) == 1 else "FAIL"


def c5_sharded(pc, torch, dist, dev, world, rank):
    """C5 as written: kNN k = 32 over a 10M-point aggregated map x 64 frames of 120k queries, the
    frames sharded over the ranks (64 / N each), the map and its index replicated on every GPU
    (built by every rank from the broadcast cloud).  Aggregate queries/s = all frames / MAX time."""
    n_map, n_frames, per = 10_000_000, 64, 120_000
    t0 = time.perf_counter()
    world_ = pc.synth.World(pc.synth.BASE_SEED + 5)
    d_map = torch.empty((n_map, 3), dtype=torch.float32, device=dev)
    if rank == 0:
        big, _ = pc.synth.aggregate(n_map, 84, seed=pc.synth.BASE_SEED + 5, world=world_)
        d_map.copy_(torch.from_numpy(big))
        del big
    if world > 1:
        dist.broadcast(d_map, 0)
    lo, hi = pc.shard_range(n_frames, world, rank)
    frames = [pc.synth.lidar_frame(per, seed=pc.synth.BASE_SEED + 5 + 1000 + f, origin=(2.6 * f + 1.0, 0.5, 0.0),
                                   world=world_)[0] for f in range(lo, hi)]
    gen_s = time.perf_counter() - t0
    pc.Index(d_map).close()   # the first build at this size pays pool growth (139 ms); the second is the steady state
    bix = pc.Index(d_map)
    d_f = torch.from_numpy(np.concatenate(frames)).to(dev)
    o_i = torch.empty((d_f.shape[0], 32), dtype=torch.int32, device=dev)
    o_d = torch.empty((d_f.shape[0], 32), dtype=torch.float32, device=dev)
    bix.knn(d_f, 32, out=(o_i, o_d))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    bix.knn(d_f, 32, out=(o_i, o_d))
    ev1.record()
    torch.cuda.synchronize()
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    nq_all = n_frames * per
    info = bix.info()
    alg = (hi - lo) * per * (16 + 8 * 32) + 16 * n_map     # per GPU
    # rows of the first frame of this rank: ascending, in range (cheap, always on)
    ok = bool((o_d[:per, 1:] >= o_d[:per, :-1]).all().item()) and bool(((o_i[:per] >= 0) & (o_i[:per] < n_map)).all().item())
    bix.close()
    return {"config": "C5 kNN k=32, 10M-point map x 64 frames of 120k, frames sharded x%d" % world, "ms": round(ms, 3),
            "queries_per_s": nq_all / (ms * 1e-3), "index_build_ms": round(info["build_ms"], 3),
            "index_bytes": info["device_bytes"], "hbm_frac_per_gpu": round(alg / (ms * 1e-3) / 1e9 / peaks()[0], 4),
            "rows_sane": ok, "host_generation_s": round(gen_s, 1)}


def count_launches(pc, torch, ix, d_q, o_idx, o_dist):
    """How many kernels ONE step launches, counted, not claimed: the step is captured into a CUDA
    graph on a side stream (never instantiated, never run) and its kernel nodes are counted."""
    try:
        from cuda.bindings import runtime as rt
        st = torch.cuda.Stream()
        ix.set_option("profile", 0)
        torch.cuda.synchronize()
        with torch.cuda.stream(st):
            (err,) = rt.cudaStreamBeginCapture(st.cuda_stream, rt.cudaStreamCaptureMode.cudaStreamCaptureModeRelaxed)
            if int(err) != 0:
                raise RuntimeError("begin capture: %s" % err)
            try:
                ix.knn(d_q, K, out=(o_idx, o_dist))
            finally:
                err, graph = rt.cudaStreamEndCapture(st.cuda_stream)
            if int(err) != 0:
                raise RuntimeError("end capture: %s" % err)
            err, _, total = rt.cudaGraphGetNodes(graph, 0)
            err, nodes, total = rt.cudaGraphGetNodes(graph, total)
            kernels = 0
            for nd in nodes[:total]:
                err, ty = rt.cudaGraphNodeGetType(nd)
                kernels += int(ty == rt.cudaGraphNodeType.cudaGraphNodeTypeKernel)
            rt.cudaGraphDestroy(graph)
        torch.cuda.synchronize()
        ix.set_option("profile", 1)
        return kernels, "kernel nodes of one captured step x steps"
    except Exception as exc:  # pragma: no cover
        try:
            ix.set_option("profile", 1)
        except Exception:
            pass
        return None, "not counted: %r" % (exc,)


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
    th_q = torch.from_numpy(qry).pin_memory()
    th_idx = torch.empty((N_QUERIES, K), dtype=torch.int32).pin_memory()
    th_dist = torch.empty((N_QUERIES, K), dtype=torch.float32).pin_memory()
    h_q, h_idx, h_dist = th_q.numpy(), th_idx.numpy(), th_dist.numpy()
    ix.set_option("profile", 0)            # the per-launch profiling events are a diagnostic of the leg above
    for _ in range(max(1, args.warmup)):   # the same W untimed warm-up steps as the device-resident leg
        ix.knn(h_q, K, out=(h_idx, h_dist))
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ix.knn(h_q, K, out=(h_idx, h_dist))
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    ix.set_option("profile", 1)
    barrier()
    # the copy CEILING of that step at this N: the same pinned buffers and bytes, all ranks at once, no
    # kernels -- H2D and D2H on two streams.  e2e is to be read against it (tools/pcie_ceiling.py).
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    d_keep_idx, d_keep_dist = o_idx.clone(), o_dist.clone()
    d_in = torch.empty_like(d_q)

    def copies():
        with torch.cuda.stream(s_in):
            d_in.copy_(th_q, non_blocking=True)
        with torch.cuda.stream(s_out):
            th_idx.copy_(d_keep_idx, non_blocking=True)
            th_dist.copy_(d_keep_dist, non_blocking=True)
    copies()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        copies()
    torch.cuda.synchronize()
    copy_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e = {"value": world * N_QUERIES * args.steps / e2e_s, "unit": "queries/s",
           "h2d_bytes_per_step": int(h_q.nbytes), "d2h_bytes_per_step": int(h_idx.nbytes + h_dist.nbytes),
           "ms_per_step": 1e3 * e2e_s / args.steps,
           "api": "pcnn_knn_host (C ABI), pinned host buffers, 10 stages on 8 streams, one search launch per stage",
           "copy_ceiling": {"queries_per_s": world * N_QUERIES * args.steps / copy_s, "ms_per_step": 1e3 * copy_s / args.steps,
                            "aggregate_GBps": world * args.steps * (h_q.nbytes + h_idx.nbytes + h_dist.nbytes) / copy_s / 1e9,
                            "what": "the step's H2D + D2H alone from the same pinned buffers, all ranks at once, no kernels"},
           "frac_of_copy_ceiling": copy_s / e2e_s}
    del d_keep_idx, d_keep_dist, d_in
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
            diag = prof.get("candidate_traffic")
        except Exception:
            traffic = diag = None
    roofline = {"bound": "hbm", "kernel": "knn_dfs_kernel<8>", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "committed ncu --set full capture of this kernel on this workload (profiles/knn_kernel_traffic.json), not measured in this run",
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": k_ms,
                "kernel_share_of_step": k_ms / (sum(step_ms) / args.steps),
                "note": "kNN is not HBM-bound: candidates are served from L2/L1 (SURVEY 8d, DESIGN.md); "
                        "the HBM-bound operation of the path is the radius search, see others[0]",
                "candidate_traffic_from_ncu_capture": diag}

    launches, launches_how = count_launches(pc, torch, ix, d_q, o_idx, o_dist)
    parity = {"knn_sample": knn_sample_parity(pc, torch, dist, dev, d_pts, qry, o_idx, world),
              "host_entry_eq_device_entry": "ok"}
    line = {
        "metric": "kNN queries/s (k=8)", "value": value, "unit": "queries/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "n_points": N_POINTS, "n_queries_per_gpu": N_QUERIES, "k": K,
                   "l2": "flushed between timed steps (256 MiB write)", "parallelism": "queries sharded x%d, index replicated" % world,
                   "index_build_ms": info["build_ms"], "index_bytes": info["device_bytes"], "parity": parity},
        "clocks": clocks, "e2e": e2e, "roofline": roofline,
        "gpu_launches": None if launches is None else launches * args.steps, "gpu_launches_how": launches_how,
    }
    others = []
    if not args.no_extras:
        if world > 1:
            sh, par = icp_sharded(ix, pc, torch, dist, dev, pts, nrm, qry, world, rank)
            others.append(sh)
            parity.update(par)
        others.append(c5_sharded(pc, torch, dist, dev, world, rank))
    if rank == 0:
        if not args.no_extras:
            others = other_configs(ix, pc, torch, dev, pts, nrm, qry, d_q) + others
            line["extra"] = next_ops(ix, pc, torch, dev, pts, nrm)
        roofline["others"] = others
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(pts, qry)
        print(json.dumps(line), flush=True)
    failed = [k for k, v in parity.items() if v != "ok"]
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if failed and rank == 0:
        log("bench.py: PARITY FAILED: %s" % {k: parity[k] for k in failed})
        return 3
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
