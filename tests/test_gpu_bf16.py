"""GPU (-m gpu): the bf16 twin of the path (BASELINE.json north_star: "1e-2 for an optional bf16 path"; SURVEY.md 8(c):
"same metrics at 1e-2 against the fp32 reference fed the bf16-rounded inputs").

Oracle = oracle/torch_port.py (the reference's operator sequence, pinned to pytorch/bts.py by the golden vectors) in fp64
on the host, fed the SAME bf16-rounded features / plane coefficients / upstream gradients the GPU gets, fp32 weights.
Covered: native bf16 kernels (forward of every tensor-core chain; backward of the head and of the 2x2 stage; standalone
LPG at upratio 1, 2, 4, 8), and the shapes / chains the library does not serve natively, where the wrappers run the fp32
CUDA path on upcast tensors (128-channel chains' backward, h*w % 8 != 0).
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import max_rel_to_peak, l2_rel
from oracle import torch_port as tp

pytestmark = pytest.mark.gpu

TOL = 1e-2


def _bf(t):
    return t.to(torch.bfloat16)


def _close(a, b, what, tol=TOL):
    a = a.detach().float().cpu().numpy().astype(np.float64)
    e1, e2 = max_rel_to_peak(a, b), l2_rel(a, b)
    assert e1 <= tol and e2 <= tol, (what, "max-rel-to-peak", e1, "l2-rel", e2)


def _close_depth(a, b, md, what):
    """element-wise rtol 1e-2 (+1e-3 max_depth); steep planes (|den| << 1, upratio 8) amplify the bf16 rounding of the plane
    itself, so at most one pixel in 1e4 may miss it"""
    a = a.detach().float().cpu().numpy().astype(np.float64)
    bad = np.abs(a - b) > (1e-3 * md + TOL * np.abs(b))
    assert bad.mean() <= 1e-4, (what, int(bad.sum()), float(np.abs(a - b).max()))


def _oracle(xb, ws, md, r, final, gb):
    wd = [w.double().requires_grad_(True) for w in ws]
    x = xb.double().requires_grad_(True)
    out = tp.chain_forward(x, wd, md, True) if final else tp.stage_forward(x, wd, md, r)[0]
    out.backward(gb.double())
    return out.detach().numpy(), x.grad.numpy(), [w.grad[:, :, 0, 0].numpy() for w in wd]


# (C_in, num_out, upratio, is_final, B, h, w): native one-pass backward | native stash backward (128-channel chains, also at
# BASELINE shapes and with a padded pixel axis) | fp32 route (h*w % 8 != 0)
CASES = [(32, 16, 1, True, 2, 96, 128), (64, 32, 2, False, 2, 48, 64), (32, 16, 1, True, 8, 480, 640),
         (64, 32, 2, False, 8, 240, 320), (128, 64, 4, False, 2, 24, 32), (128, 128, 8, False, 2, 12, 16),
         (128, 64, 4, False, 8, 120, 160), (128, 128, 8, False, 8, 60, 80), (128, 128, 8, False, 3, 19, 20),
         (64, 32, 2, False, 1, 9, 7), (32, 16, 1, True, 1, 15, 17),
         # the head's bf16 twin on the warp-level MMA kernels: 32-pixel groups that straddle images, a ragged last group
         (32, 16, 1, True, 3, 4, 10), (32, 16, 1, True, 5, 6, 12)]


@pytest.mark.parametrize("cfg", CASES, ids=lambda c: "x".join(map(str, c)))
def test_bf16_stage_vs_fp32_reference_on_rounded_inputs(cfg):
    import bts_b200
    from bts_b200 import functional as Fn
    cin, nout, r, final, B, h, w = cfg
    md = 10.0
    gen = torch.Generator().manual_seed(4000 + cin + r + B)
    dims = tp.chain_dims(cin, nout, final)
    ws = tp.xavier_weights(dims, final, gen)
    xb = _bf(torch.randn(B, cin, h, w, generator=gen))
    gb = _bf(torch.randn((B, 1, h, w) if final else (B, h * r, w * r), generator=gen))
    ref_out, ref_gx, ref_gw = _oracle(xb.float(), ws, md, r, final, gb.float())

    red = bts_b200.reduction_1x1(cin, nout, md, is_final=final).cuda()
    with torch.no_grad():
        for p, wt in zip(red.chain_weights(), ws):
            p.copy_(wt.cuda())
    # inference: the forward is native for every tensor-core chain whatever the shape
    with torch.no_grad():
        xd = xb.cuda()
        out = red(xd) if final else bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), xd)
    assert out.dtype == torch.bfloat16
    if final:
        _close(out, ref_out, "out (no grad)")
    else:
        _close_depth(out, ref_out, md, "depth (no grad)")
    # training
    xd = xb.cuda().requires_grad_(True)
    lpg = None if final else bts_b200.local_planar_guidance(r)
    out = red(xd) if final else bts_b200.fused_stage(red, lpg, xd)
    assert out.dtype == torch.bfloat16
    if final:
        _close(out, ref_out, "out")
    else:
        _close_depth(out, ref_out, md, "depth")
        assert lpg.abs_min.dtype == torch.float32 and float(lpg.abs_min) > 0
    out.backward(gb.cuda())
    torch.cuda.synchronize()
    assert xd.grad.dtype == torch.bfloat16
    _close(xd.grad, ref_gx, "grad_x")
    for i, (p, gr) in enumerate(zip(red.chain_weights(), ref_gw)):
        assert p.grad.dtype == torch.float32
        _close(p.grad.view(p.shape[0], p.shape[1]), gr, "grad_w%d" % i)
    # one-pass chains want h*w % 8 == 0, the stash route of the 128-channel chains h*w % 4 == 0
    assert Fn.bf16_native(dims, final, True, h, w, None, r) == ((h * w) % (8 if cin <= 64 else 4) == 0)


@pytest.mark.parametrize("r,B,h,w", [(8, 2, 12, 20), (4, 2, 24, 28), (2, 3, 30, 44), (1, 1, 16, 24), (8, 8, 60, 80)])
def test_bf16_standalone_lpg(r, B, h, w):
    """strict drop-in route: local_planar_guidance on a bf16 plane_eq (as reduction_1x1 + the Python glue produce it)"""
    import bts_b200
    gen = torch.Generator().manual_seed(77 + r)
    th = torch.rand(B, h, w, generator=gen) * (np.pi / 3) * 0.9
    ph = torch.rand(B, h, w, generator=gen) * 2 * np.pi
    pe = _bf(torch.stack([th.sin() * ph.cos(), th.sin() * ph.sin(), th.cos(), torch.rand(B, h, w, generator=gen) * 9 + 1], 1))
    gb = _bf(torch.randn(B, h * r, w * r, generator=gen))
    pr = pe.double().requires_grad_(True)
    ref, ref_amin = tp.lpg_forward(pr, r)
    ref.backward(gb.double())
    m = bts_b200.local_planar_guidance(r)
    pd = pe.cuda().requires_grad_(True)
    depth = m(pd, None)
    assert depth.dtype == torch.bfloat16 and depth.shape == (B, h * r, w * r)
    _close_depth(depth, ref.detach().numpy(), 10.0, "depth")
    assert abs(float(m.abs_min) - float(ref_amin)) <= 1e-3
    depth.backward(gb.cuda())
    _close(pd.grad, pr.grad.numpy(), "grad_plane_eq")


def test_bf16_strict_dropin_route():
    """reduction_1x1 -> the reference's own glue lines (bts.py:250-253) in bf16 torch ops -> local_planar_guidance"""
    import bts_b200
    gen = torch.Generator().manual_seed(5)
    cin, nout, r, B, h, w, md = 64, 32, 2, 2, 40, 48, 10.0
    ws = tp.xavier_weights(tp.chain_dims(cin, nout, False), False, gen)
    xb = _bf(torch.randn(B, cin, h, w, generator=gen))
    ref_out, _, _ = _oracle(xb.float(), ws, md, r, False, torch.zeros(B, h * r, w * r))
    red = bts_b200.reduction_1x1(cin, nout, md).cuda()
    with torch.no_grad():
        for p, wt in zip(red.chain_weights(), ws):
            p.copy_(wt.cuda())
        pe4 = red(xb.cuda())
        n = F.normalize(pe4[:, :3, :, :].float(), 2, 1).to(torch.bfloat16)
        depth = bts_b200.local_planar_guidance(r)(torch.cat([n, pe4[:, 3, :, :].unsqueeze(1)], 1), None)
    assert pe4.dtype == torch.bfloat16 and depth.dtype == torch.bfloat16
    # two bf16 roundings of the plane (after the chain, after the glue) sit between the features and the depth
    a = depth.float().cpu().numpy().astype(np.float64)
    assert np.mean(np.abs(a - ref_out) > 2e-2 * np.abs(ref_out) + 1e-2) <= 1e-4


@pytest.mark.parametrize("cfg", [(64, 32, 2, 2, 40, 48), (128, 64, 4, 2, 24, 32), (128, 128, 8, 2, 12, 20)], ids=str)
def test_bf16_scaled_depth(cfg):
    """fused_stage(..., scaled=True): depth.unsqueeze(1) / max_depth (bts.py:255) folded in, on the one-pass and the stash
    backward of the bf16 twin"""
    import bts_b200
    cin, nout, r, B, h, w = cfg
    md = 80.0
    gen = torch.Generator().manual_seed(900 + cin + r)
    ws = tp.xavier_weights(tp.chain_dims(cin, nout, False), False, gen)
    xb = _bf(torch.randn(B, cin, h, w, generator=gen))
    gb = _bf(torch.randn(B, 1, h * r, w * r, generator=gen))
    wd = [t.double().requires_grad_(True) for t in ws]
    xr = xb.double().requires_grad_(True)
    ref = tp.stage_forward(xr, wd, md, r)[0].unsqueeze(1) / md
    ref.backward(gb.double())
    red = bts_b200.reduction_1x1(cin, nout, md).cuda()
    with torch.no_grad():
        for p, wt in zip(red.chain_weights(), ws):
            p.copy_(wt.cuda())
    xd = xb.cuda().requires_grad_(True)
    out = bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), xd, scaled=True)
    assert out.dtype == torch.bfloat16 and out.shape == (B, 1, h * r, w * r)
    _close_depth(out, ref.detach().numpy(), 1.0, "scaled depth")
    out.backward(gb.cuda())
    _close(xd.grad, xr.grad.numpy(), "grad_x")
    for i, (p, g) in enumerate(zip(red.chain_weights(), wd)):
        _close(p.grad.view(p.shape[0], p.shape[1]), g.grad[:, :, 0, 0].numpy(), "grad_w%d" % i)
