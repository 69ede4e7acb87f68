#!/usr/bin/env python
"""BASELINE.json configs[4]: LPG decoder throughput sweep, batch 1..256 at the NYU and KITTI shapes (fp32), one GPU.
Per (shape, batch): device time of forward + backward of the four stages through the C ABI (back-to-back calls
between CUDA events, as bench.py's per-call timing), Mpixels/s and fraction of the HBM roofline on 733.25 B/px.
Usage: sweep.py [max_batch]  ->  markdown table on stdout"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402


def main():
    max_b = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    dev = torch.device("cuda", 0)
    peak, _ = bench.hbm_peak()
    bpp = sum(sum(v) for v in bench.ALGO_BYTES.values())
    shapes = (("NYU 416x544", 416, 544, 10.0), ("NYU 480x640", 480, 640, 10.0), ("KITTI 352x704", 352, 704, 80.0),
              ("KITTI 352x1216", 352, 1216, 80.0))
    print("| shape | batch | us fwd+bwd | Mpixels/s | HBM GB/s | of %.0f GB/s |" % peak)
    print("|---|---|---|---|---|---|")
    for name, H, W, md in shapes:
        b = 1
        while b <= max_b:
            wl = dict(H=H, W=W, B=b, max_depth=md, focal=0.0, desc=name)
            try:
                stages = bench.build_stages(wl, dev, 0)
                us = 0.0
                for st in stages:
                    f, bw = bench.stage_calls(st, wl, dev)
                    iters = 20 if b <= 32 else 5
                    us += (bench.time_back_to_back(f, iters) + bench.time_back_to_back(bw, iters)) * 1e3
                    del f, bw
                px = b * H * W
                gbs = bpp * px / us / 1e3
                print("| %s | %d | %.0f | %.0f | %.0f | %.1f %% |" % (name, b, us, px / us, gbs, 100 * gbs / peak), flush=True)
            except torch.OutOfMemoryError:
                print("| %s | %d | out of memory | | | |" % (name, b), flush=True)
                break
            finally:
                stages = None
                torch.cuda.empty_cache()
            b *= 2


if __name__ == "__main__":
    main()
