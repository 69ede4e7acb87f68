#!/usr/bin/env python
"""Post-LPG glue (pytorch/bts.py:255-256, 269-270, 283) folded into the fused stage vs left to PyTorch:
device time of forward + backward of each LPG stage through the module API, replayed from a CUDA graph.
Usage: glue_times.py [workload] [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
import bench  # noqa: E402
import bts_b200  # noqa: E402


def graph_time(fn, iters):
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
    wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "nyu_b8"]
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    dev = torch.device("cuda", 0)
    md = wl["max_depth"]
    for st in bench.build_stages(wl, dev, 0):
        if st["final"]:
            continue
        r, red, lpg, x = st["r"], st["red"], st["lpg"], st["x"]
        B, _, h, w = x.shape
        ds = r >= 4
        g_full = torch.randn(B, 1, h * r, w * r, device=dev)
        g_ds = torch.randn(B, 1, 2 * h, 2 * w, device=dev) if ds else None

        def clear():
            x.grad = None
            for p in red.parameters():
                p.grad = None

        def fused():
            clear()
            out = bts_b200.fused_stage(red, lpg, x, scaled=True, downsample=ds)
            if ds:
                torch.autograd.backward([out[0], out[1]], [g_full, g_ds])
            else:
                out.backward(g_full)

        def eager_glue():
            clear()
            depth = bts_b200.fused_stage(red, lpg, x)
            scaled = depth.unsqueeze(1) / md
            if ds:
                small = F.interpolate(scaled, scale_factor=2.0 / r, mode="nearest")
                torch.autograd.backward([scaled, small], [g_full, g_ds])
            else:
                scaled.backward(g_full)

        tf, te = graph_time(fused, iters), graph_time(eager_glue, iters)
        print("%-18s fwd+bwd: glue fused %7.1f us   glue in PyTorch %7.1f us   (saves %5.1f us, %4.1f%%)" % (
            st["name"], tf, te, te - tf, 100 * (te - tf) / te))


if __name__ == "__main__":
    main()
