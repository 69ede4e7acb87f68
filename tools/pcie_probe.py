#!/usr/bin/env python
"""Host<->device copy rates on this box: H2D alone, D2H alone, both directions at once (two streams).
Explains the e2e leg of bench.py, which moves 609 MB each way per step."""
import time
import torch

n = 320 * 1024 * 1024 // 4
h_in = torch.empty(n, dtype=torch.float32).pin_memory()
h_out = torch.empty(n, dtype=torch.float32).pin_memory()
d_a = torch.empty(n, dtype=torch.float32, device="cuda")
d_b = torch.randn(n, dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(h2d, d2h, iters=5):
    torch.cuda.synchronize()
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
    gb = n * 4 / 1e9
    return dt * 1e3, gb / dt


for name, a, b in (("H2D alone", 1, 0), ("D2H alone", 0, 1), ("H2D + D2H concurrently", 1, 1)):
    run(a, b, 2)
    ms, gbs = run(a, b)
    print("%-26s %7.2f ms per 335 MB%s  -> %.1f GB/s per direction" % (name, ms, " each way" if a and b else "", gbs))
