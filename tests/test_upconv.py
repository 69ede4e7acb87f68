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


@pytest.mark.gpu
@pytest.mark.parametrize("g", CASES, ids=lambda g: g.name)
def test_upconv_kernel_vs_golden(g):
    import bts_b200
    m = g.meta
    mod = bts_b200.upconv(m["cin"], m["cout"]).cuda()
    with torch.no_grad():
        mod.conv.weight.copy_(torch.from_numpy(g["w"]).cuda())
        x = torch.from_numpy(g["x"]).cuda()
        if m["cin"] % 16 == 0 and m["cout"] % 32 == 0:
            assert mod.native(x)
        y = mod(x)
    torch.cuda.synchronize()
    assert tuple(y.shape) == g["f64_y"].shape
    assert max_rel_to_peak(y.cpu().numpy(), g["f64_y"]) <= 1e-5
    assert np.allclose(y.cpu().numpy(), g["f64_y"], rtol=1e-4, atol=1e-5 * np.abs(g["f64_y"]).max())


@pytest.mark.gpu
@pytest.mark.parametrize("cin,cout,B,h,w", [(64, 32, 2, 60, 80), (128, 64, 1, 30, 40), (512, 256, 1, 15, 20), (2208, 512, 1, 7, 10)], ids=str)
def test_upconv_kernel_vs_oracle_decoder_shapes(cin, cout, B, h, w):
    """The decoder's own channel counts (bts.py:226-285, bts_size 512 on densenet161) against the fp64 oracle on the host."""
    from bts_b200 import functional as Fn
    gen = torch.Generator().manual_seed(7 + cin)
    x = torch.randn(B, cin, h, w, generator=gen)
    wt = torch.randn(cout, cin, 3, 3, generator=gen) * (2.0 / (9 * cin)) ** 0.5
    ref = tp.upconv_forward(x.double(), wt.double()).numpy()
    with torch.no_grad():
        y = Fn.upconv_forward(x.cuda(), wt.cuda())
    assert max_rel_to_peak(y.cpu().numpy(), ref) <= 1e-5


@pytest.mark.gpu
def test_upconv_module_contract():
    import bts_b200
    mod = bts_b200.upconv(64, 32)
    assert list(mod.state_dict().keys()) == ["conv.weight"] and mod.ratio == 2 and isinstance(mod.elu, torch.nn.ELU)
    mod = mod.cuda()
    x = torch.randn(1, 64, 6, 10, device="cuda", requires_grad=True)
    y = mod(x)                                     # a gradient is required: the reference's operator sequence (backward not built)
    y.sum().backward()
    assert x.grad is not None and not mod.native(x)
    from bts_b200 import functional as Fn
    with pytest.raises(RuntimeError):
        Fn.upconv_forward(x, mod.conv.weight)
    with torch.no_grad(), pytest.raises(RuntimeError):
        Fn.upconv_forward(torch.randn(1, 24, 4, 4, device="cuda"), torch.randn(32, 24, 3, 3, device="cuda"))
