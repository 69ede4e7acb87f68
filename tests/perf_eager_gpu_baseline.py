#!/usr/bin/env python
"""Second reported baseline of SURVEY.md 8(d): the reference's operator sequence (oracle/torch_port.py, the
restatement of pytorch/bts.py:97-173 + glue) run by eager PyTorch ON THE B200, TF32 disabled, same four stages
and shapes as bench.py. Test infrastructure (lives under tests/, not collected by pytest): prints one JSON line.
    python tests/perf_eager_gpu_baseline.py [workload] [iters]"""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from oracle import torch_port as tp  # noqa: E402


def main():
    wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "nyu_b8"]
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.set_float32_matmul_precision("highest")
    dev = torch.device("cuda", 0)
    gen = torch.Generator().manual_seed(0)
    items = []
    for name, cin, nout, r, final in bench.STAGES:
        dims = tp.chain_dims(cin, nout, final)
        ws = [w.to(dev).requires_grad_(True) for w in tp.xavier_weights(dims, final, gen)]
        x = torch.randn(wl["B"], cin, wl["H"] // r, wl["W"] // r, device=dev, requires_grad=True)
        g = torch.randn((wl["B"], 1, wl["H"], wl["W"]) if final else (wl["B"], wl["H"], wl["W"]), device=dev)
        items.append((name, r, final, ws, x, g))

    def step():
        for name, r, final, ws, x, g in items:
            for t in ws + [x]:
                t.grad = None
            out = tp.chain_forward(x, ws, wl["max_depth"], True) if final else tp.stage_forward(x, ws, wl["max_depth"], r)[0]
            out.backward(g)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = statistics.median(ts)
    px = wl["B"] * wl["H"] * wl["W"]
    print(json.dumps({"baseline": "reference operator sequence, eager PyTorch %s on %s, TF32 off" % (
        torch.__version__, torch.cuda.get_device_name(0)), "workload": wl["desc"], "ms_per_step": ms,
        "value": px / ms / 1e3, "unit": "Mpixels/s", "iters": iters}))


if __name__ == "__main__":
    main()
