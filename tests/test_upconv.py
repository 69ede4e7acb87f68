"""upconv (pytorch/bts.py:83-94), forward: golden vectors of the imported reference class
(tests/golden/upconv_*.npz by make_golden.py). CPU: the oracle's operator sequence and the folded restatement the kernel
implements both reproduce the reference; GPU: the kernel through the C ABI / module mirror."""
import numpy as np
import pytest
import torch

from conftest import golden_cases, max_rel_to_peak
from oracle import torch_port as tp

CASES = golden_cases("upconv")


@pytest.mark.parametrize("g", CASES, ids=lambda g: g.name)
def test_upconv_oracle_vs_golden(g):
    x, w = torch.from_numpy(g["x"]), torch.from_numpy(g["w"])
    assert np.array_equal(tp.upconv_forward(x, w).numpy(), g["f32_y"])                     # same ATen operators: bit-exact
    y64 = tp.upconv_forward_folded(x.double(), w.double()).numpy()
    assert max_rel_to_peak(y64, g["f64_y"]) <= 1e-12                                       # the fold is an identity
    assert max_rel_to_peak(tp.upconv_forward_folded(x, w).numpy(), g["f64_y"]) <= 2e-6


def test_upconv_fold_and_unfold_are_adjoint():
    """<fold(W), G> == <W, unfold(G)> and autograd through the folded restatement gives unfold(grad of the folded weights):
    pins the un-fold step of the weight-gradient kernel to the oracle."""
    gen = torch.Generator().manual_seed(5)
    w = torch.randn(6, 5, 3, 3, generator=gen, dtype=torch.float64)
    g = torch.randn(2, 2, 2, 2, 6, 5, generator=gen, dtype=torch.float64)
    assert abs(float((tp.upconv_fold_weights(w) * g).sum() - (w * tp.upconv_unfold_grads(g)).sum())) < 1e-10
    wl = w.clone().requires_grad_(True)
    wf = tp.upconv_fold_weights(wl)
    (wf * g).sum().backward()
    assert torch.allclose(wl.grad, tp.upconv_unfold_grads(g), atol=1e-12)


def _oracle_grads(x, w, gy, dtype):
    xd = x.detach().clone().to(dtype).requires_grad_(True)
    wd = w.detach().clone().to(dtype).requires_grad_(True)
    y = tp.upconv_forward(xd, wd)
    y.backward(gy.to(dtype))
    return y.detach(), xd.grad, wd.grad


@pytest.mark.parametrize("g", CASES, ids=lambda g: g.name)
def test_upconv_oracle_backward_vs_golden(g):
    y, gx, gw = _oracle_grads(torch.from_numpy(g["x"]), torch.from_numpy(g["w"]), torch.from_numpy(g["grad_y"]), torch.float64)
    assert max_rel_to_peak(gx.numpy(), g["f64_grad_x"]) <= 1e-12 and max_rel_to_peak(gw.numpy(), g["f64_grad_w"]) <= 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("g", CASES, ids=lambda g: g.name)
def test_upconv_kernel_vs_golden(g):
    import bts_b200
    m = g.meta
    mod = bts_b200.upconv(m["cin"], m["cout"]).cuda()
    with torch.no_grad():
        mod.conv.weight.copy_(torch.from_numpy(g["w"]).cuda())
        x = torch.from_numpy(g["x"]).cuda()
        assert mod.native(x)
        y = mod(x)
    torch.cuda.synchronize()
    assert tuple(y.shape) == g["f64_y"].shape
    assert max_rel_to_peak(y.cpu().numpy(), g["f64_y"]) <= 1e-5
    assert np.allclose(y.cpu().numpy(), g["f64_y"], rtol=1e-4, atol=1e-5 * np.abs(g["f64_y"]).max())
    # backward (served for cin % 32 == 0; otherwise the module runs the reference's operator sequence, also checked)
    xg = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
    assert mod.native(xg) == (m["cin"] % 32 == 0)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False              # (the stock sequence is 3e-4 off with PyTorch's TF32 default)
    try:
        yg = mod(xg)
        yg.backward(torch.from_numpy(g["grad_y"]).cuda())
        torch.cuda.synchronize()
    finally:
        torch.backends.cudnn.allow_tf32 = tf32
    assert max_rel_to_peak(yg.detach().cpu().numpy(), g["f64_y"]) <= 1e-5
    assert max_rel_to_peak(xg.grad.cpu().numpy(), g["f64_grad_x"]) <= 1e-5
    assert max_rel_to_peak(mod.conv.weight.grad.cpu().numpy(), g["f64_grad_w"]) <= 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("cin,cout,B,h,w", [(64, 32, 2, 60, 80), (128, 64, 1, 30, 40), (512, 256, 1, 15, 20), (2208, 512, 1, 7, 10)], ids=str)
def test_upconv_kernel_vs_oracle_decoder_shapes(cin, cout, B, h, w):
    """The decoder's own channel counts (bts.py:226-285, bts_size 512 on densenet161), forward and backward, against the
    fp64 oracle (autograd of the reference's operator sequence) on the host."""
    from bts_b200 import functional as Fn
    gen = torch.Generator().manual_seed(7 + cin)
    x = torch.randn(B, cin, h, w, generator=gen)
    wt = torch.randn(cout, cin, 3, 3, generator=gen) * (2.0 / (9 * cin)) ** 0.5
    gy = torch.randn(B, cout, 2 * h, 2 * w, generator=gen)
    ref_y, ref_gx, ref_gw = _oracle_grads(x, wt, gy, torch.float64)
    xd, wd = x.cuda().requires_grad_(True), wt.cuda().requires_grad_(True)
    y = Fn.upconv_forward(xd, wd)
    y.backward(gy.cuda())
    torch.cuda.synchronize()
    assert max_rel_to_peak(y.detach().cpu().numpy(), ref_y.numpy()) <= 1e-5
    assert max_rel_to_peak(xd.grad.cpu().numpy(), ref_gx.numpy()) <= 1e-5
    assert max_rel_to_peak(wd.grad.cpu().numpy(), ref_gw.numpy()) <= 1e-5
    with torch.no_grad():
        assert torch.equal(Fn.upconv_forward(x.cuda(), wt.cuda()), y.detach())


@pytest.mark.gpu
def test_upconv_full_size_configs1():
    """upconv2 of the decoder at BASELINE.json configs[1] (NYU 480x640, batch 8): [8,128,120,160] -> [8,64,240,320], forward
    and backward against the fp64 oracle on the host (5 G multiply-adds per pass)."""
    from bts_b200 import functional as Fn
    gen = torch.Generator().manual_seed(11)
    x = torch.randn(8, 128, 120, 160, generator=gen)
    wt = torch.randn(64, 128, 3, 3, generator=gen) * (2.0 / (9 * 128)) ** 0.5
    gy = torch.randn(8, 64, 240, 320, generator=gen)
    ref_y, ref_gx, ref_gw = _oracle_grads(x, wt, gy, torch.float64)
    xd, wd = x.cuda().requires_grad_(True), wt.cuda().requires_grad_(True)
    y = Fn.upconv_forward(xd, wd)
    y.backward(gy.cuda())
    torch.cuda.synchronize()
    assert max_rel_to_peak(y.detach().cpu().numpy(), ref_y.numpy()) <= 1e-5
    assert max_rel_to_peak(xd.grad.cpu().numpy(), ref_gx.numpy()) <= 1e-5
    assert max_rel_to_peak(wd.grad.cpu().numpy(), ref_gw.numpy()) <= 1e-5


@pytest.mark.gpu
def test_upconv_module_contract():
    import bts_b200
    from bts_b200 import functional as Fn
    mod = bts_b200.upconv(64, 32)
    assert list(mod.state_dict().keys()) == ["conv.weight"] and mod.ratio == 2 and isinstance(mod.elu, torch.nn.ELU)
    mod = mod.cuda()
    x = torch.randn(1, 64, 6, 10, device="cuda", requires_grad=True)
    assert mod.native(x)
    mod(x).sum().backward()
    assert x.grad is not None and mod.conv.weight.grad is not None
    only_x = torch.randn(1, 64, 6, 10, device="cuda", requires_grad=True)       # frozen weights: grad_x only
    w_frozen = mod.conv.weight.detach()
    Fn.upconv_forward(only_x, w_frozen).sum().backward()
    ref_in = only_x.detach().double().cpu().requires_grad_(True)
    ref_gx = torch.autograd.grad(tp.upconv_forward(ref_in, w_frozen.double().cpu()).sum(), ref_in)[0]
    assert max_rel_to_peak(only_x.grad.cpu().numpy(), ref_gx.numpy()) <= 1e-5
    odd = bts_b200.upconv(24, 32).cuda()                                         # not served: the reference's sequence
    xo = torch.randn(1, 24, 4, 4, device="cuda")
    assert not odd.native(xo) and odd(xo).shape == (1, 32, 8, 8)
    with torch.no_grad(), pytest.raises(RuntimeError):
        Fn.upconv_forward(xo, odd.conv.weight)
