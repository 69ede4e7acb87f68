#!/usr/bin/env python
"""Error of every output of the four stages against the fp64 oracle at a bench workload's FULL shapes
(max|a-b| / max|b|; the bar is 1e-5). The tuning companion of tests/test_gpu_baseline_shapes.py: prints the margins.
Usage: tests/perf_parity_report.py [workload] [stage-substring]   (test infrastructure: uses oracle/; not collected by pytest)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import bench  # noqa: E402
import bts_b200  # noqa: E402
from conftest import max_rel_to_peak  # noqa: E402
from oracle import torch_port as tp  # noqa: E402
import test_gpu_baseline_shapes as T  # noqa: E402


def main():
    wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "nyu_b8"]
    only = sys.argv[2] if len(sys.argv) > 2 else ""
    B, H, W, md = wl["B"], wl["H"], wl["W"], wl["max_depth"]
    for name, cin, nout, r, final in bench.STAGES:
        if only not in name:
            continue
        gen = torch.Generator().manual_seed(1)
        ws = tp.xavier_weights(tp.chain_dims(cin, nout, final), final, gen)
        x = torch.randn(B, cin, H // r, W // r, generator=gen)
        g = torch.randn((B, 1, H, W) if final else (B, H, W), generator=gen)
        ref_out, ref_gx, ref_gw, _, _ = T._oracle_fp64(x, ws, md, r, final, g)
        red = bts_b200.reduction_1x1(cin, nout, md, is_final=final).cuda()
        with torch.no_grad():
            for p, wt in zip(red.chain_weights(), ws):
                p.copy_(wt.cuda())
        xd = x.cuda().requires_grad_(True)
        out = red(xd) if final else bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), xd)
        out.backward(g.cuda())
        torch.cuda.synchronize()
        eo = max_rel_to_peak(out.detach().cpu().numpy(), ref_out)
        ex = max_rel_to_peak(xd.grad.cpu().numpy(), ref_gx)
        ew = [max_rel_to_peak(p.grad.cpu().numpy().reshape(gr.shape), gr) for p, gr in zip(red.chain_weights(), ref_gw)]
        print("%-18s out %.2e  grad_x %.2e  grad_W %s" % (name, eo, ex, " ".join("%.2e" % e for e in ew)), flush=True)


if __name__ == "__main__":
    main()
