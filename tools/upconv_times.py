#!/usr/bin/env python
"""Device times of the five decoder upconvs (pytorch/bts.py:226-285, bts_size 512 on densenet161) at a bench workload's
shapes, forward and forward + backward: the folded tensor-core kernels (bts_upconv_fwd / bts_upconv_bwd) against the
reference's operator sequence on cuDNN, with TF32 off (the reference's arithmetic) and on (PyTorch's default).
Usage: upconv_times.py [workload] [iters]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
import bench  # noqa: E402
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
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    B, H, W = wl["B"], wl["H"], wl["W"]
    tot = [0.0] * 6
    for name, cin, cout, div in (("upconv5", 2208, 512, 32), ("upconv4", 512, 256, 16), ("upconv3", 256, 128, 8),
                                 ("upconv2", 128, 64, 4), ("upconv1", 64, 32, 2)):
        h, w = H // div, W // div
        x = torch.randn(B, cin, h, w, device="cuda", requires_grad=True)
        wt = (torch.randn(cout, cin, 3, 3, device="cuda") * (2.0 / (9 * cin)) ** 0.5).requires_grad_(True)
        gy = torch.randn(B, cout, 2 * h, 2 * w, device="cuda")
        ref = lambda: F.elu(F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), wt, padding=1))  # noqa: E731

        def fb(f):
            x.grad = wt.grad = None
            f().backward(gy)

        def fwd_only(f):
            with torch.no_grad():
                return f()

        ours = lambda: Fn.upconv_forward(x, wt)  # noqa: E731
        t = [timed(lambda: fwd_only(ours), iters), timed(lambda: fb(ours), iters)]
        torch.backends.cudnn.allow_tf32 = False
        t += [timed(lambda: fwd_only(ref), iters), timed(lambda: fb(ref), iters)]
        fb(ref)
        gx_ref, gw_ref, y_ref = x.grad.clone(), wt.grad.clone(), fwd_only(ref)
        torch.backends.cudnn.allow_tf32 = True
        t += [timed(lambda: fwd_only(ref), iters), timed(lambda: fb(ref), iters)]
        fb(ours)
        errs = (float((fwd_only(ours) - y_ref).abs().max() / y_ref.abs().max()), float((x.grad - gx_ref).abs().max() / gx_ref.abs().max()),
                float((wt.grad - gw_ref).abs().max() / gw_ref.abs().max()))
        tot = [a + b for a, b in zip(tot, t)]
        print("%s %4d->%3d [%d,%d,%d]  ours fwd %7.3f fwd+bwd %7.3f ms | cuDNN fp32 %7.3f / %7.3f | cuDNN TF32 %7.3f / %7.3f | "
              "ours vs cuDNN fp32: y %.1e gx %.1e gW %.1e" % ((name, cin, cout, B, h, w) + tuple(t) + errs), flush=True)
    print("sum  ours fwd %.3f fwd+bwd %.3f ms | cuDNN fp32 %.3f / %.3f | cuDNN TF32 %.3f / %.3f" % tuple(tot))


if __name__ == "__main__":
    main()
