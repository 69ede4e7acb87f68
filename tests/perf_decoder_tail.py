#!/usr/bin/env python
"""the hot path in context. The decoder from daspp_feat on (pytorch/bts.py:249-293: reduc8x8 -> lpg8x8 ->
upconv3 -> conv3 -> reduc4x4 -> ... -> reduc1x1 -> conv1 -> get_depth) at a bench workload's shape and the real widths
(bts_size 512, densenet161 skips), forward + backward, timed on the B200 before and after bts_b200.graft(). The stand-in
decoder is the one tests/test_gpu_dropin.py grafts (classes NAMED like the reference's, arithmetic = oracle/torch_port.py,
because the reference tree does not travel to the GPU box). Test infrastructure (lives under tests/, not collected by
pytest): prints one JSON line.      python tests/perf_decoder_tail.py [workload] [iters]"""
import copy
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.nn as nn  # noqa: E402
import bench  # noqa: E402
import bts_b200  # noqa: E402
from test_gpu_dropin import DecoderTail  # noqa: E402


class FusedDecoderTail(DecoderTail):
    """The same decoder with INTEGRATION.md section 3's edit: each `reduc -> normalize/cat -> lpg -> / max_depth
    [-> interpolate]` group is one bts_b200.fused_stage call."""

    def forward(self, daspp_feat, skip1, skip0, focal):
        F = torch.nn.functional
        d8, d8_ds = bts_b200.fused_stage(self.reduc8x8, self.lpg8x8, daspp_feat, scaled=True, downsample=True)
        iconv3 = self.conv3(torch.cat([self.bn3(self.upconv3(daspp_feat)), skip1, d8_ds], 1))
        d4, d4_ds = bts_b200.fused_stage(self.reduc4x4, self.lpg4x4, iconv3, scaled=True, downsample=True)
        iconv2 = self.conv2(torch.cat([self.bn2(self.upconv2(iconv3)), skip0, d4_ds], 1))
        d2 = bts_b200.fused_stage(self.reduc2x2, self.lpg2x2, iconv2, scaled=True)
        upconv1 = self.upconv1(iconv2)
        reduc1x1 = self.reduc1x1(upconv1)
        iconv1 = self.conv1(torch.cat([upconv1, reduc1x1, d2, d4, d8], 1))
        return d8, d4, d2, reduc1x1, self.max_depth * self.get_depth(iconv1)


def timed(model, inputs, gouts, iters):
    def step():
        model.zero_grad(set_to_none=True)
        ins = [t.detach().requires_grad_(True) for t in inputs[:3]]
        outs = model(*ins, inputs[3])
        torch.autograd.backward(outs, gouts)
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
    return statistics.median(ts)


def main():
    wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "nyu_b8"]
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    B, H, W, md = wl["B"], wl["H"], wl["W"], wl["max_depth"]
    torch.manual_seed(0)
    model = DecoderTail(md, nf=512, skips=(96, 96)).cuda()
    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.xavier_uniform_(m.weight)
    inputs = [torch.randn(B, 128, H // 8, W // 8, device="cuda"), torch.randn(B, 96, H // 4, W // 4, device="cuda"),
              torch.randn(B, 96, H // 2, W // 2, device="cuda"), torch.full((B,), wl["focal"], dtype=torch.float64, device="cuda")]
    gouts = [torch.randn(B, 1, H, W, device="cuda") for _ in range(5)]
    res = {}
    for tf32 in (False, True):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        tag = "tf32_on" if tf32 else "tf32_off"
        res["stock_" + tag] = timed(copy.deepcopy(model), inputs, gouts, iters)
        res["graft_lpg_path_" + tag] = timed(bts_b200.graft(copy.deepcopy(model)), inputs, gouts, iters)
        res["graft_lpg_path_and_upconvs_" + tag] = timed(bts_b200.graft(copy.deepcopy(model), upconvs=True), inputs, gouts, iters)
        fused = bts_b200.graft(copy.deepcopy(model))
        fused.__class__ = FusedDecoderTail
        res["graft_fused_stages_" + tag] = timed(fused, inputs, gouts, iters)
    px = B * H * W
    print(json.dumps({"what": "decoder from daspp_feat on (bts.py:249-293), forward + backward, ms per step (median of %d)" % iters,
                      "workload": wl["desc"], "device": torch.cuda.get_device_name(0), "ms": res,
                      "mpixels_per_s": {k: px / v / 1e3 for k, v in res.items()},
                      "note": "tf32_off = the reference's fp32 arithmetic for the stock cuDNN layers; tf32_on = PyTorch's default. "
                              "graft_lpg_path: reduction_1x1 + local_planar_guidance swapped (strict drop-in, bts.forward untouched); "
                              "graft_fused_stages: plus the three fused_stage edits of INTEGRATION.md section 3."}))


if __name__ == "__main__":
    main()
