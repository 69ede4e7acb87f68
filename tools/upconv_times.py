#!/usr/bin/env python
"""Device times of the five decoder upconvs (pytorch/bts.py:226-285, bts_size 512 on densenet161) at a bench workload's
shapes: the folded tensor-core kernel (bts_upconv_fwd) against the reference's operator sequence on cuDNN, with TF32 off
(the reference's arithmetic) and on (PyTorch's default). Forward only.  Usage: upconv_times.py [workload] [iters]"""
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
    tot = [0.0, 0.0, 0.0]
    with torch.no_grad():
        for name, cin, cout, div in (("upconv5", 2208, 512, 32), ("upconv4", 512, 256, 16), ("upconv3", 256, 128, 8),
                                     ("upconv2", 128, 64, 4), ("upconv1", 64, 32, 2)):
            h, w = H // div, W // div
            x = torch.randn(B, cin, h, w, device="cuda")
            wt = torch.randn(cout, cin, 3, 3, device="cuda") * (2.0 / (9 * cin)) ** 0.5
            ref = lambda: F.elu(F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), wt, padding=1))  # noqa: E731
            t_ours = timed(lambda: Fn.upconv_forward(x, wt), iters)
            torch.backends.cudnn.allow_tf32 = False
            t_fp32 = timed(ref, iters)
            y_ref = ref()
            torch.backends.cudnn.allow_tf32 = True
            t_tf32 = timed(ref, iters)
            err = float((Fn.upconv_forward(x, wt) - y_ref).abs().max() / y_ref.abs().max())
            err_tf32 = float((ref() - y_ref).abs().max() / y_ref.abs().max())
            macs = 16.0 * cin * cout * B * h * w
            tot = [tot[0] + t_ours, tot[1] + t_fp32, tot[2] + t_tf32]
            print("%s %4d->%3d [%d,%d,%d]  ours %8.3f ms (%5.1f T MAC/s folded)   cuDNN fp32 %8.3f ms   cuDNN TF32 %8.3f ms   "
                  "|ours - cuDNN fp32| / peak %.1e (TF32: %.1e)" % (name, cin, cout, B, h, w, t_ours, macs / t_ours / 1e9, t_fp32, t_tf32, err, err_tf32),
                  flush=True)
    print("sum  ours %.3f ms   cuDNN fp32 %.3f ms   cuDNN TF32 %.3f ms" % tuple(tot))


if __name__ == "__main__":
    main()
