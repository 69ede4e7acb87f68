#!/usr/bin/env python
"""Host<->device copy rates on this box: H2D alone, D2H alone, both directions at once (two streams).
Explains the e2e leg of bench.py, which moves 609 MB each way per step (fp32).

Under torchrun (one rank per GPU) every rank copies at the same time and rank 0 prints the box's aggregate: the host-side
ceiling that bounds the e2e leg at N > 1 (8 GPUs x 1.22 GB per step topped out near 131 GB/s).
    python tools/pcie_probe.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_probe.py"""
import os
import time

import torch

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

n = 320 * 1024 * 1024 // 4
h_in = torch.empty(n, dtype=torch.float32).pin_memory()
h_out = torch.empty(n, dtype=torch.float32).pin_memory()
d_a = torch.empty(n, dtype=torch.float32, device="cuda")
d_b = torch.randn(n, dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, iters=5):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(iters):
        if h2d:
            with torch.cuda.stream(s1):
                d_a.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_b, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / iters
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    gb = n * 4 / 1e9
    return dt * 1e3, gb / dt


if rank == 0:
    try:
        print("host: %d logical CPUs; affinity of this rank: %d CPUs" % (os.cpu_count(), len(os.sched_getaffinity(0))))
    except Exception:
        pass
for name, a, b in (("H2D alone", 1, 0), ("D2H alone", 0, 1), ("H2D + D2H concurrently", 1, 1)):
    run(a, b, 2)
    ms, gbs = run(a, b)
    if rank == 0:
        print("%-26s %7.2f ms per 335 MB%s  -> %.1f GB/s per direction per GPU, %.1f GB/s per direction over %d GPU(s)%s"
              % (name, ms, " each way" if a and b else "", gbs, gbs * world, world,
                 " = %.1f GB/s both directions" % (2 * gbs * world) if a and b else ""))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
