"""GPU (-m gpu): the drop-in where it is meant to live.

* graft(): a stand-in decoder with the exact call sequence of the reference's bts.forward (pytorch/bts.py:249-293:
  reduc -> Python normalise/cat -> lpg -> /max_depth -> nearest down-sampling -> stock convolutions -> next stage),
  built from classes NAMED reduction_1x1 / local_planar_guidance whose arithmetic is oracle/torch_port.py (the
  reference tree does not travel to the GPU box), is run before and after bts_b200.graft(); the five depth outputs
  and every gradient must agree.
* re-entrancy (include/bts_b200.h: no global mutable state; nn.DataParallel = one Python thread per replica,
  pytorch/bts_test.py:91): two threads, two streams, concurrent fused_stage forward + backward == serial results.
* environment routes of INTEGRATION.md (BTS_DETERMINISTIC=1, BTS_HEAD_NWG=3, BTS_PDL=0, BTS_HEAD_MMA=0/1/3): read once per process, so each
  runs in a subprocess against the C oracle.
"""
import copy
import os
import subprocess
import sys
import threading

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from conftest import max_rel_to_peak, l2_rel, ROOT
from oracle import torch_port as tp

pytestmark = pytest.mark.gpu


# ---- stand-ins for the reference classes (same names, constructor signatures, child-module names, state_dict keys)
class reduction_1x1(nn.Sequential):          # noqa: N801 -- the reference's class name is the contract (pytorch/bts.py:97)
    def __init__(self, num_in_filters, num_out_filters, max_depth, is_final=False):
        super().__init__()
        self.max_depth, self.is_final = max_depth, is_final
        self.sigmoid = nn.Sigmoid()
        self.reduc = nn.Sequential()
        while num_out_filters >= 4:          # the loop of bts.py:105-122
            if num_out_filters < 8:
                if is_final:
                    self.reduc.add_module("final", nn.Sequential(nn.Conv2d(num_in_filters, 1, 1, bias=False), nn.Sigmoid()))
                else:
                    self.reduc.add_module("plane_params", nn.Conv2d(num_in_filters, 3, 1, bias=False))
                break
            self.reduc.add_module("inter_{}_{}".format(num_in_filters, num_out_filters),
                                  nn.Sequential(nn.Conv2d(num_in_filters, num_out_filters, 1, bias=False), nn.ELU()))
            num_in_filters, num_out_filters = num_out_filters, num_out_filters // 2

    def forward(self, net):
        ws = [m.weight for m in self.reduc.modules() if isinstance(m, nn.Conv2d)]
        return tp.chain_forward(net, ws, self.max_depth, self.is_final)


class local_planar_guidance(nn.Module):      # noqa: N801 (pytorch/bts.py:138)
    def __init__(self, upratio):
        super().__init__()
        self.upratio = float(upratio)
        self.abs_min = None

    def forward(self, plane_eq, focal):
        depth, self.abs_min = tp.lpg_forward(plane_eq, int(self.upratio))
        return depth


class upconv(nn.Module):                     # noqa: N801 (pytorch/bts.py:83-94)
    def __init__(self, cin, cout):
        super().__init__()
        self.conv, self.elu, self.ratio = nn.Conv2d(cin, cout, 3, 1, 1, bias=False), nn.ELU(), 2

    def forward(self, x):
        return self.elu(self.conv(F.interpolate(x, scale_factor=self.ratio, mode="nearest")))


class DecoderTail(nn.Module):
    """bts.forward from daspp_feat on (pytorch/bts.py:249-293), num_features = 512, with stand-in skip widths."""

    def __init__(self, max_depth, nf=512, skips=(8, 8)):
        super().__init__()
        self.max_depth = max_depth
        self.reduc8x8, self.lpg8x8 = reduction_1x1(nf // 4, nf // 4, max_depth), local_planar_guidance(8)
        self.upconv3, self.bn3 = upconv(nf // 4, nf // 4), nn.BatchNorm2d(nf // 4, momentum=0.01, eps=1.1e-5)
        self.conv3 = nn.Sequential(nn.Conv2d(nf // 4 + skips[1] + 1, nf // 4, 3, 1, 1, bias=False), nn.ELU())
        self.reduc4x4, self.lpg4x4 = reduction_1x1(nf // 4, nf // 8, max_depth), local_planar_guidance(4)
        self.upconv2, self.bn2 = upconv(nf // 4, nf // 8), nn.BatchNorm2d(nf // 8, momentum=0.01, eps=1.1e-5)
        self.conv2 = nn.Sequential(nn.Conv2d(nf // 8 + skips[0] + 1, nf // 8, 3, 1, 1, bias=False), nn.ELU())
        self.reduc2x2, self.lpg2x2 = reduction_1x1(nf // 8, nf // 16, max_depth), local_planar_guidance(2)
        self.upconv1 = upconv(nf // 8, nf // 16)
        self.reduc1x1 = reduction_1x1(nf // 16, nf // 32, max_depth, is_final=True)
        self.conv1 = nn.Sequential(nn.Conv2d(nf // 16 + 4, nf // 16, 3, 1, 1, bias=False), nn.ELU())
        self.get_depth = nn.Sequential(nn.Conv2d(nf // 16, 1, 3, 1, 1, bias=False), nn.Sigmoid())

    def _stage(self, reduc, lpg, feat, focal):
        r = reduc(feat)
        n = F.normalize(r[:, :3, :, :], 2, 1)
        pe = torch.cat([n, r[:, 3, :, :].unsqueeze(1)], 1)
        return lpg(pe, focal).unsqueeze(1) / self.max_depth

    def forward(self, daspp_feat, skip1, skip0, focal):
        d8 = self._stage(self.reduc8x8, self.lpg8x8, daspp_feat, focal)
        d8_ds = F.interpolate(d8, scale_factor=0.25, mode="nearest")
        iconv3 = self.conv3(torch.cat([self.bn3(self.upconv3(daspp_feat)), skip1, d8_ds], 1))
        d4 = self._stage(self.reduc4x4, self.lpg4x4, iconv3, focal)
        d4_ds = F.interpolate(d4, scale_factor=0.5, mode="nearest")
        iconv2 = self.conv2(torch.cat([self.bn2(self.upconv2(iconv3)), skip0, d4_ds], 1))
        d2 = self._stage(self.reduc2x2, self.lpg2x2, iconv2, focal)
        upconv1 = self.upconv1(iconv2)
        reduc1x1 = self.reduc1x1(upconv1)
        iconv1 = self.conv1(torch.cat([upconv1, reduc1x1, d2, d4, d8], 1))
        return d8, d4, d2, reduc1x1, self.max_depth * self.get_depth(iconv1)


def _run(model, inputs, gouts, dtype):
    model = model.to(dtype)
    ins = [t.to(dtype).detach().clone().requires_grad_(True) for t in inputs[:3]]
    model.zero_grad()
    outs = model(*ins, inputs[3])
    loss = sum((o * g.to(dtype)).sum() for o, g in zip(outs, gouts))
    loss.backward()
    grads = [t.grad for t in ins] + [p.grad for p in model.parameters()]
    return [o.detach().double().cpu().numpy() for o in outs], [g.double().cpu().numpy() for g in grads]


@pytest.mark.parametrize("upconvs", [False, True], ids=["lpg-path", "lpg-path+upconvs"])
def test_graft_inside_the_decoder_call_sequence(upconvs):
    import bts_b200
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    B, H, W, md = 2, 96, 128, 10.0
    model = DecoderTail(md).cuda()
    for m in model.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.xavier_uniform_(m.weight)                      # weights_init_xavier, bts.py:34-38
    inputs = [torch.randn(B, 128, H // 8, W // 8, device="cuda"), torch.randn(B, 8, H // 4, W // 4, device="cuda"),
              torch.randn(B, 8, H // 2, W // 2, device="cuda"), torch.full((B,), 518.8579, dtype=torch.float64, device="cuda")]
    shapes = [(B, 1, H, W)] * 5
    gouts = [torch.randn(s, device="cuda") for s in shapes]
    keys_before = list(model.state_dict().keys())

    ref64 = _run(copy.deepcopy(model), inputs, gouts, torch.float64)
    ref32 = _run(copy.deepcopy(model), inputs, gouts, torch.float32)
    grafted = bts_b200.graft(copy.deepcopy(model), upconvs=upconvs)
    assert list(grafted.state_dict().keys()) == keys_before
    for name in ("upconv3", "upconv2", "upconv1"):                 # (opt-in: SURVEY.md 8(f)-3)
        assert isinstance(getattr(grafted, name), bts_b200.upconv) == upconvs
    for name in ("reduc8x8", "reduc4x4", "reduc2x2", "reduc1x1"):
        assert isinstance(getattr(grafted, name), bts_b200.reduction_1x1)
        for a, b in zip(getattr(grafted, name).parameters(), getattr(model, name).parameters()):
            assert torch.equal(a, b)
    for name in ("lpg8x8", "lpg4x4", "lpg2x2"):
        assert isinstance(getattr(grafted, name), bts_b200.local_planar_guidance)
    ours = _run(grafted, inputs, gouts, torch.float32)
    assert float(grafted.lpg8x8.abs_min) > 0 and grafted.lpg2x2.abs_min.dim() == 0   # the monitor of bts_main.py:484-486

    # both fp32 models share the stock cuDNN layers; the bar is the fp32 stand-in's own distance from fp64
    names = ["depth_8x8_scaled", "depth_4x4_scaled", "depth_2x2_scaled", "reduc1x1", "final_depth"]
    for kind, got, r32, r64 in (("out", ours[0], ref32[0], ref64[0]), ("grad", ours[1], ref32[1], ref64[1])):
        for i, (a, b32, b64) in enumerate(zip(got, r32, r64)):
            own = max_rel_to_peak(b32, b64)
            err = max_rel_to_peak(a, b64)
            assert err <= max(2e-5, 3 * own), (kind, names[i] if kind == "out" else i, err, own)
            assert l2_rel(a, b64) <= max(2e-5, 3 * l2_rel(b32, b64)), (kind, i)


def test_weights_at_unaligned_offsets_take_the_generic_route():
    """Weights that are views at odd offsets of a flat buffer (flattened parameters) are not 16-byte aligned: the
    tensor-core kernels pack them with 128-bit loads, so the library must route such a call to the generic kernels."""
    import bts_b200
    torch.manual_seed(2)
    red = bts_b200.reduction_1x1(64, 32, 10.0).cuda()
    for p in red.parameters():
        nn.init.xavier_uniform_(p)
    x = torch.randn(2, 64, 12, 16, device="cuda")
    g = torch.randn(2, 24, 32, device="cuda")
    lpg = bts_b200.local_planar_guidance(2)
    xa = x.clone().requires_grad_(True)
    ref = bts_b200.fused_stage(red, lpg, xa)
    ref.backward(g)
    ref_gw = [p.grad.clone() for p in red.parameters()]
    shifted = copy.deepcopy(red)
    flat = torch.empty(sum(p.numel() + 8 for p in red.parameters()) + 8, device="cuda")
    off = 1                                                       # every view starts 4 bytes past a 16-byte boundary + k
    for p, q in zip(red.parameters(), shifted.parameters()):
        view = flat[off:off + p.numel()].view_as(p)
        view.copy_(p.detach())
        q.data = view
        off += p.numel() + (5 - p.numel() % 4)
    assert any(q.data_ptr() % 16 for q in shifted.parameters())
    xb = x.clone().requires_grad_(True)
    out = bts_b200.fused_stage(shifted, lpg, xb)
    out.backward(g)
    assert torch.allclose(out, ref, rtol=2e-5, atol=1e-5)
    assert max_rel_to_peak(xb.grad.cpu().numpy(), xa.grad.cpu().numpy()) <= 2e-5
    for q, r in zip(shifted.parameters(), ref_gw):
        assert max_rel_to_peak(q.grad.cpu().numpy(), r.cpu().numpy()) <= 2e-5


def test_downsample_is_rejected_for_upratio_2():
    import bts_b200
    red = bts_b200.reduction_1x1(64, 32, 10.0).cuda()
    x = torch.randn(1, 64, 8, 12, device="cuda")
    with pytest.raises(RuntimeError):
        bts_b200.fused_stage(red, bts_b200.local_planar_guidance(2), x, scaled=True, downsample=True)


@pytest.mark.parametrize("cfg", [(128, 64, 4, 2, 60, 80), (64, 32, 2, 2, 120, 160), (32, 16, 0, 2, 240, 320)], ids=str)
def test_two_threads_two_streams_match_serial(cfg):
    """DataParallel contract: replicas of the modules driven from two Python threads on their own streams."""
    import bts_b200
    cin, nout, r, B, h, w = cfg
    final = r == 0
    torch.manual_seed(17)
    base = bts_b200.reduction_1x1(cin, nout, 10.0, is_final=final).cuda()
    for p in base.parameters():
        nn.init.xavier_uniform_(p)
    data = [(torch.randn(B, cin, h, w, device="cuda"),
             torch.randn((B, 1, h, w) if final else (B, h * r, w * r), device="cuda")) for _ in range(2)]

    def work(i, stream, out, iters):
        red = copy.deepcopy(base)
        with torch.cuda.stream(stream):
            for _ in range(iters):
                x = data[i][0].clone().requires_grad_(True)
                red.zero_grad()
                o = red(x) if final else bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), x)
                o.backward(data[i][1])
            out[i] = (o.detach(), x.grad, [p.grad.clone() for p in red.parameters()])
        stream.synchronize()

    torch.cuda.synchronize()
    serial, conc = [None, None], [None, None]
    for i in range(2):
        work(i, torch.cuda.current_stream(), serial, 1)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for s in streams:
        s.wait_stream(torch.cuda.current_stream())
    threads = [threading.Thread(target=work, args=(i, streams[i], conc, 6)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    for i in range(2):
        assert conc[i] is not None, "worker thread %d raised" % i
        assert torch.equal(conc[i][0], serial[i][0])               # outputs and grad_x: fixed summation order
        assert torch.equal(conc[i][1], serial[i][1])
        for a, b in zip(conc[i][2], serial[i][2]):                 # grad_W: reproducible to fp32 rounding
            assert max_rel_to_peak(a.cpu().numpy(), b.cpu().numpy()) <= 2e-6


_ENV_CHECK = r"""
import sys, numpy as np, torch
sys.path.insert(0, %(root)r)
import bts_b200
from oracle import c_oracle as co, torch_port as tp
gen = torch.Generator().manual_seed(99)
worst = 0.0
for cin, nout, r, final, B, h, w in ((32, 16, 1, True, 2, 72, 100), (64, 32, 2, False, 2, 40, 52), (128, 64, 4, False, 2, 24, 36)):
    dims = tp.chain_dims(cin, nout, final)
    ws = [t[:, :, 0, 0].numpy() for t in tp.xavier_weights(dims, final, gen)]
    x = torch.randn(B, cin, h, w, generator=gen).numpy()
    g = torch.randn((B, 1, h, w) if final else (B, h * r, w * r), generator=gen).numpy()
    red = bts_b200.reduction_1x1(cin, nout, 10.0, is_final=final).cuda()
    with torch.no_grad():
        for p, wt in zip(red.chain_weights(), ws):
            p.copy_(torch.from_numpy(wt).cuda().view_as(p))
    runs = []
    for rep in range(2):
        xd = torch.from_numpy(x).cuda().requires_grad_(True)
        red.zero_grad()
        out = red(xd) if final else bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), xd)
        out.backward(torch.from_numpy(g).cuda())
        runs.append([xd.grad.clone()] + [p.grad.clone() for p in red.parameters()])
    gx, gw = (co.reduc_bwd(x, ws, 10.0, True, g, np.float64) if final else co.fused_bwd(x, ws, 10.0, r, g, np.float64))
    for a, b in zip(runs[0], [gx] + list(gw)):
        a = a.cpu().numpy().reshape(b.shape).astype(np.float64)
        worst = max(worst, float(np.abs(a - b).max() / np.abs(b).max()))
    same = all(torch.equal(a, b) for a, b in zip(runs[0][:1], runs[1][:1]))
    assert same, "grad_x differs between two identical calls"
print("WORST %%.3e" %% worst)
assert worst <= 1e-5, worst
"""


@pytest.mark.parametrize("env", [{"BTS_DETERMINISTIC": "1"}, {"BTS_HEAD_NWG": "3"}, {"BTS_PDL": "0"}, {"BTS_HEAD_MMA": "0"},
                                 {"BTS_HEAD_MMA": "1"}, {"BTS_HEAD_MMA": "3"}, {"BTS_HEAD_EXACT": "1"}], ids=lambda e: "-".join("%s=%s" % kv for kv in e.items()))
def test_environment_routes_vs_c_oracle(env):
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, "-c", _ENV_CHECK % {"root": ROOT}], env=e, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "WORST" in r.stdout
