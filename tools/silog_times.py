#!/usr/bin/env python
"""silog_loss (pytorch/bts.py:41-48) as one streaming reduction vs the reference's operator sequence in eager
PyTorch on the same GPU: device time of forward + backward (CUDA events; ours replayed from a CUDA graph, the eager sequence has a dynamic-shape gather and cannot be captured) and the
HBM fraction of the fused kernels (algorithmic bytes: 9 B/px forward, 13 B/px backward)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import bts_b200  # noqa: E402


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def graph_timed(fn, iters=20):
    """Device time of fn replayed from a CUDA graph (the fused loss has no dynamic shape and no host sync)."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    peak, _ = bench.hbm_peak()
    crit = bts_b200.silog_loss(0.85)
    for shape in ((8, 1, 480, 640), (16, 1, 352, 1216), (64, 1, 480, 640)):
        est = (torch.rand(shape, device="cuda") * 9.5 + 0.5).requires_grad_(True)
        gt = torch.rand(shape, device="cuda") * 9.5 + 0.5
        mask = torch.rand(shape, device="cuda") < 0.4
        n = est.numel()

        def ours():
            est.grad = None
            crit(est, gt, mask).backward()

        def eager():
            est.grad = None
            d = torch.log(est[mask]) - torch.log(gt[mask])
            (torch.sqrt((d ** 2).mean() - 0.85 * (d.mean() ** 2)) * 10.0).backward()

        to, te = graph_timed(ours), timed(eager)
        print("%-20s fused fwd+bwd %7.1f us (%4.1f%% of HBM peak on 22 B/px)   eager PyTorch ops %7.1f us   x%.1f" % (
            "x".join(map(str, shape)), to, 100 * 22.0 * n / to / 1e3 / peak, te, te / to))


if __name__ == "__main__":
    main()
