#!/usr/bin/env python
"""Per-stage DEVICE times of the forward and backward C-ABI calls at a bench workload's shapes: N calls
back to back between two CUDA events (host launch cost is pipelined away), so the figure is what the
kernels of one call take. The quick loop used while tuning kernels.
Usage: stage_times.py [workload] [iters] [stage-name substring]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from bts_b200 import _cabi  # noqa: E402
from bts_b200 import functional as Fn  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "nyu_b8"]
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    dev = torch.device("cuda", 0)
    stages = bench.build_stages(wl, dev, 0)
    px = wl["B"] * wl["H"] * wl["W"]
    peak, _ = bench.hbm_peak()
    lib = _cabi.lib()
    stream = torch.cuda.current_stream().cuda_stream
    tot = 0.0
    only = sys.argv[3] if len(sys.argv) > 3 else ""
    for st in stages:
        if only not in st["name"]:
            continue
        x, g, r, final = st["x"].detach(), st["g"], st["r"], st["final"]
        ws = [w.detach() for w in st["red"].chain_weights()]
        dims = [x.shape[1]] + [w.shape[0] for w in ws]
        B, _, h, w = x.shape
        W = _cabi.ptr_array([t.data_ptr() for t in ws])
        D = _cabi.int_array(dims)
        nws = lib.bts_workspace_bytes(D, len(ws), B, h, w, 0 if final else r)
        wsb = torch.empty((nws + 3) // 4, dtype=torch.float32, device=dev)
        gx = torch.empty_like(x)
        gws = [torch.zeros_like(t) for t in ws]
        GW = _cabi.ptr_array([t.data_ptr() for t in gws])
        md = float(wl["max_depth"])
        if final:
            out = torch.empty(B, 1, h, w, device=dev)
            f = lambda: _cabi.check(lib.bts_reduc_fwd(x.data_ptr(), W, D, len(ws), 1, md, B, h, w, out.data_ptr(),
                                                      wsb.data_ptr(), nws, stream), "fwd")
            b = lambda: _cabi.check(lib.bts_reduc_bwd(x.data_ptr(), W, D, len(ws), 1, md, B, h, w, g.data_ptr(),
                                                      gx.data_ptr(), GW, wsb.data_ptr(), nws, stream), "bwd")
        else:
            depth = torch.empty(B, h * r, w * r, device=dev)
            amin = torch.full((), float("inf"), device=dev)
            f = lambda: _cabi.check(lib.bts_reduc_lpg_fwd(x.data_ptr(), W, D, len(ws), md, B, h, w, r, Fn.EPS, None,
                                                          depth.data_ptr(), amin.data_ptr(), wsb.data_ptr(), nws, stream), "fwd")
            b = lambda: _cabi.check(lib.bts_reduc_lpg_bwd(g.data_ptr(), x.data_ptr(), W, D, len(ws), md, B, h, w, r, Fn.EPS,
                                                          gx.data_ptr(), GW, wsb.data_ptr(), nws, stream), "bwd")
        tf, tb = timed(f, iters), timed(b, iters)
        fb, bb = bench.ALGO_BYTES[st["name"]]
        tot += tf + tb
        print("%-18s fwd %7.1f us (%5.1f%% of HBM peak)   bwd %7.1f us (%5.1f%%)" % (
            st["name"], tf * 1e3, 100 * fb * px / tf / 1e6 / peak, tb * 1e3, 100 * bb * px / tb / 1e6 / peak))
    allb = sum(sum(v) for v in bench.ALGO_BYTES.values()) * px
    print("sum %.1f us -> %.1f Mpix/s, %.1f%% of HBM peak" % (tot * 1e3, px / tot / 1e3, 100 * allb / tot / 1e6 / peak))


if __name__ == "__main__":
    main()
