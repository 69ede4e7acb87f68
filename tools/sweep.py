#!/usr/bin/env python
"""BASELINE.json configs[4]: LPG decoder throughput sweep, batch 1..256 at the NYU and KITTI shapes, fp32 and bf16, at
1 / 2 / 4 / 8 B200 vs the HBM roofline.

Per (shape, per-GPU batch, dtype): device time of forward + backward of the four stages (fp32: back-to-back C-ABI calls
between CUDA events, as bench.py's per-call timing; bf16: the same through the module API, because its 128-channel
backward runs the fp32 kernels on upcast tensors), Mpixels/s and the fraction of the measured HBM peak on 733.25 B/px
(fp32) / 366.6 B/px (bf16).

Multi-GPU (batch sharding, no data-path collective): launch under torchrun -- every rank times its own batch on its own
GPU, the slowest rank's time counts (max over ranks), the throughput is the whole job's:

    python tools/sweep.py [--max-batch 256] [--dtype f32,bf16] [--shapes nyu480,kitti1216]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/sweep.py ...

-> markdown table on stdout (rank 0)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402

SHAPES = {"nyu416": ("NYU 416x544", 416, 544, 10.0), "nyu480": ("NYU 480x640", 480, 640, 10.0),
          "kitti704": ("KITTI 352x704", 352, 704, 80.0), "kitti1216": ("KITTI 352x1216", 352, 1216, 80.0)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-batch", type=int, default=256)
    ap.add_argument("--dtype", default="f32,bf16")
    ap.add_argument("--shapes", default=",".join(SHAPES))
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=dev)
        bench._DIST["on"] = True
    peak, _ = bench.hbm_peak()
    bpp32 = sum(sum(v) for v in bench.ALGO_BYTES.values())
    if rank == 0:
        print("| shape | dtype | GPUs | batch / GPU | us fwd+bwd (max over ranks) | Mpixels/s (whole job) | HBM GB/s per GPU | of %.0f GB/s |" % peak)
        print("|---|---|---|---|---|---|---|---|")
    for key in args.shapes.split(","):
        name, H, W, md = SHAPES[key]
        for dtype in args.dtype.split(","):
            bpp = bpp32 * (0.5 if dtype == "bf16" else 1.0)
            b = 1
            while b <= args.max_batch:
                wl = dict(H=H, W=W, B=b, max_depth=md, focal=0.0, desc=name)
                oom = 0.0
                us = 0.0
                stages = None
                try:
                    stages = bench.build_stages(wl, dev, rank, dtype)
                    iters = 20 if b <= 32 else 5
                    for st in stages:
                        if dtype == "bf16":
                            f, fb = bench.stage_calls_api(st)
                            us += bench.time_back_to_back(fb, iters) * 1e3
                        else:
                            f, bw = bench.stage_calls(st, wl, dev)
                            us += (bench.time_back_to_back(f, iters) + bench.time_back_to_back(bw, iters)) * 1e3
                except torch.OutOfMemoryError:
                    oom = 1.0
                finally:
                    stages = None
                    f = fb = bw = None
                    torch.cuda.empty_cache()
                oom = bench.max_over_ranks(oom, dev)              # every rank leaves the loop together
                if oom:
                    if rank == 0:
                        print("| %s | %s | %d | %d | out of memory | | | |" % (name, dtype, world, b), flush=True)
                    break
                us = bench.max_over_ranks(us, dev)
                px = b * H * W
                gbs = bpp * px / us / 1e3
                if rank == 0:
                    print("| %s | %s | %d | %d | %.0f | %.0f | %.0f | %.1f %% |" % (name, dtype, world, b, us, world * px / us, gbs,
                                                                                    100 * gbs / peak), flush=True)
                b *= 2
    if world > 1:
        import torch.distributed as dist
        bench.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
