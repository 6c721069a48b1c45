"""Developer tool (GPU box, under torch.distributed.run): the host <-> device copy CEILING of the
end-to-end kNN step at N ranks.  Every rank does nothing but the step's copies from / to pinned host
memory -- 12 MB H2D (1M queries x 12 B) and 64 MB D2H (1M x 8 x (int32 + float32)) -- all ranks at
once, and rank 0 prints the aggregate GB/s and the queries/s that bandwidth would allow.  Three
patterns: both directions at once on two streams (what a pipelined step can approach), D2H only,
H2D only.  bench.py's e2e at the same N is to be read against the first line.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/pcie_ceiling.py
"""
import json
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
NQ, K = 1_000_000, 8
h_in = torch.empty(NQ * 3, dtype=torch.float32).pin_memory()
h_out = torch.empty(NQ * K * 2, dtype=torch.int32).pin_memory()
h_in.fill_(1.0); h_out.fill_(0)          # first touch
d_in = torch.empty_like(h_in, device=dev)
d_out = torch.ones(NQ * K * 2, dtype=torch.int32, device=dev)
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
STEPS = int(os.environ.get("STEPS", "20"))


def run(h2d, d2h):
    for _ in range(2):
        if h2d: d_in.copy_(h_in, non_blocking=True)
        if d2h: h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(STEPS):
        if h2d:
            with torch.cuda.stream(s_in):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s_out):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item())


out = {"n_gpus": world, "steps": STEPS, "h2d_bytes_per_step": h_in.numel() * 4, "d2h_bytes_per_step": h_out.numel() * 4}
for name, a, b in (("both", True, True), ("d2h_only", False, True), ("h2d_only", True, False)):
    dt = run(a, b)
    nbytes = world * STEPS * ((h_in.numel() * 4 if a else 0) + (h_out.numel() * 4 if b else 0))
    out[name] = {"aggregate_GBps": round(nbytes / dt / 1e9, 1), "ms_per_step": round(1e3 * dt / STEPS, 3)}
    if name == "both":
        out[name]["ceiling_queries_per_s"] = round(world * NQ * STEPS / dt)
if rank == 0:
    print("PCIE_CEILING " + json.dumps(out), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
