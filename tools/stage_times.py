#!/usr/bin/env python
"""Per-stage DEVICE times of the forward and backward C-ABI calls at a bench workload's shapes: N calls
back to back between two CUDA events (host launch cost is pipelined away), so the figure is what the
kernels of one call take. The quick loop used while tuning kernels.
Usage: stage_times.py [workload] [iters] [stage-name substring] [f32|bf16]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402


def main():
    wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "nyu_b8"]
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    dev = torch.device("cuda", 0)
    dtype = sys.argv[4] if len(sys.argv) > 4 else "f32"
    sc = 0.5 if dtype == "bf16" else 1.0
    stages = bench.build_stages(wl, dev, 0, dtype)
    px = wl["B"] * wl["H"] * wl["W"]
    peak, _ = bench.hbm_peak()
    tot = 0.0
    only = sys.argv[3] if len(sys.argv) > 3 else ""
    for st in stages:
        if only not in st["name"]:
            continue
        if dtype == "bf16":
            f, fbk = bench.stage_calls_api(st)
            tf = bench.time_back_to_back(f, iters)
            tb = bench.time_back_to_back(fbk, iters) - tf
        else:
            f, b = bench.stage_calls(st, wl, dev)
            tf, tb = bench.time_back_to_back(f, iters), bench.time_back_to_back(b, iters)
        fb, bb = (sc * v for v in bench.ALGO_BYTES[st["name"]])
        tot += tf + tb
        print("%-18s fwd %7.1f us (%5.1f%% of HBM peak)   bwd %7.1f us (%5.1f%%)" % (
            st["name"], tf * 1e3, 100 * fb * px / tf / 1e6 / peak, tb * 1e3, 100 * bb * px / tb / 1e6 / peak))
    if dtype == "f32" and (not only or "lpg" in only):
        for name, d in bench.time_standalone_lpg(wl, dev, iters).items():
            print("%-18s fwd %7.1f us (%5.1f%% of HBM peak)   bwd %7.1f us (%5.1f%%)" % (
                name, d["fwd_ms"] * 1e3, 100 * d["fwd_gbs"] / peak, d["bwd_ms"] * 1e3, 100 * d["bwd_gbs"] / peak))
    allb = sc * sum(sum(v) for v in bench.ALGO_BYTES.values()) * px
    print("sum %.1f us -> %.1f Mpix/s, %.1f%% of HBM peak" % (tot * 1e3, px / tot / 1e3, 100 * allb / tot / 1e6 / peak))


if __name__ == "__main__":
    main()
