"""CPU: pin both oracle restatements to the golden vectors produced by the unmodified reference
(tests/golden/make_golden.py ran /root/reference/pytorch/bts.py)."""
import numpy as np
import pytest
import torch

from conftest import golden_cases, max_rel_to_peak
from oracle import c_oracle as co
from oracle import torch_port as tp

CHAIN = golden_cases("chain")
LPG = golden_cases("lpg")


def test_golden_inventory():
    assert len(CHAIN) >= 10 and len(LPG) >= 7
    names = {g.name for g in CHAIN}
    # the four decoder instances of bts_size=512 (pytorch/bts.py:198-217) must be covered
    assert {"chain_8x8_128_128", "chain_4x4_128_64", "chain_2x2_64_32", "chain_head_32_16"} <= names


@pytest.mark.parametrize("g", LPG, ids=lambda g: g.name)
@pytest.mark.parametrize("tag,dt", [("f32", np.float32), ("f64", np.float64)])
def test_c_oracle_lpg(g, tag, dt):
    r = g.meta["upratio"]
    depth, amin = co.lpg_fwd(g["plane_eq"], r, dt)
    # same operation order as the reference => bit-identical, including inf at den == 0
    assert np.array_equal(depth, g[tag + "_depth"], equal_nan=True)
    assert np.array_equal(np.asarray(amin), g[tag + "_abs_min"], equal_nan=True)
    grad = co.lpg_bwd(g["grad_out"], g["plane_eq"], r, dt)
    ref = g[tag + "_grad_plane_eq"]
    assert np.array_equal(np.isnan(grad), np.isnan(ref))
    assert np.array_equal(np.isinf(grad), np.isinf(ref))
    fin = np.isfinite(ref)
    assert max_rel_to_peak(grad[fin], ref[fin]) < (2e-6 if dt == np.float32 else 1e-13)


@pytest.mark.parametrize("g", CHAIN, ids=lambda g: g.name)
@pytest.mark.parametrize("tag,dt", [("f32", np.float32), ("f64", np.float64)])
def test_c_oracle_chain(g, tag, dt):
    m = g.meta
    tol = 5e-6 if dt == np.float32 else 1e-13
    if m["is_final"]:
        out = co.reduc_fwd(g["x"], g.weights, m["max_depth"], True, dt)
        gx, gw = co.reduc_bwd(g["x"], g.weights, m["max_depth"], True, g["grad_out"], dt)
    else:
        res = co.fused_fwd(g["x"], g.weights, m["max_depth"], m["upratio"], dt)
        out = res["out"]
        ref_depth = g[tag + "_depth"]
        assert np.abs(res["depth"] - ref_depth).max() <= tol * np.abs(ref_depth).max()
        assert np.allclose(res["depth"], ref_depth, rtol=4 * tol, atol=0)
        assert max_rel_to_peak(res["plane_eq"], g[tag + "_plane_eq"]) < tol
        assert abs(float(res["abs_min"]) - float(g[tag + "_abs_min"])) < tol
        gx, gw = co.fused_bwd(g["x"], g.weights, m["max_depth"], m["upratio"], g["grad_out"], dt)
    assert max_rel_to_peak(out, g[tag + "_out"]) < tol
    assert max_rel_to_peak(gx, g[tag + "_grad_x"]) < tol
    for i, a in enumerate(gw):
        assert max_rel_to_peak(a, g[tag + "_grad_w%d" % i]) < tol


@pytest.mark.parametrize("g", CHAIN, ids=lambda g: g.name)
def test_torch_port_chain_bit_exact(g):
    """The torch port issues the reference's ATen ops in the reference's order: bit-equal on CPU."""
    torch.set_num_threads(1)
    m = g.meta
    assert tp.chain_dims(m["num_in"], m["num_out"], m["is_final"]) == m["dims"]
    ws = [torch.from_numpy(w)[:, :, None, None].clone().requires_grad_(True) for w in g.weights]
    x = torch.from_numpy(g["x"]).clone().requires_grad_(True)
    out = tp.chain_forward(x, ws, m["max_depth"], m["is_final"])
    assert np.array_equal(out.detach().numpy(), g["f32_out"])
    if m["is_final"]:
        final = out
    else:
        final, amin = tp.lpg_forward(tp.glue(out), m["upratio"])
        assert np.array_equal(final.detach().numpy(), g["f32_depth"])
        assert float(amin) == float(g["f32_abs_min"])
    final.backward(torch.from_numpy(g["grad_out"]))
    assert max_rel_to_peak(x.grad.numpy(), g["f32_grad_x"]) < 1e-6
    for i, w in enumerate(ws):
        assert max_rel_to_peak(w.grad.numpy()[:, :, 0, 0], g["f32_grad_w%d" % i]) < 1e-6


@pytest.mark.parametrize("g", LPG, ids=lambda g: g.name)
def test_torch_port_lpg_bit_exact(g):
    depth, amin = tp.lpg_forward(torch.from_numpy(g["plane_eq"]), g.meta["upratio"])
    assert np.array_equal(depth.numpy(), g["f32_depth"], equal_nan=True)
    assert np.array_equal(np.asarray(amin.numpy()), g["f32_abs_min"], equal_nan=True)


def test_known_answers_lpg():
    """Hand-computable cases: clamp edges, zero denominator, r = 1."""
    pe = np.zeros((1, 4, 1, 3), dtype=np.float64)
    pe[0, 2, 0] = [5e-4, -5e-4, 0.0]
    pe[0, 3, 0] = [1.0, 1.0, 1.0]
    depth, amin = co.lpg_fwd(pe, 2, np.float64)
    assert np.allclose(depth[0, :, 0:2], 1000.0) and np.allclose(depth[0, :, 2:4], -1000.0)
    assert np.all(np.isinf(depth[0, :, 4:6])) and amin == 0.0
    # clamped pixels pass no gradient to n1..n3 and g/(+-eps) to n4
    g = np.ones((1, 2, 6))
    gr = co.lpg_bwd(g, pe, 2, np.float64)
    assert np.all(gr[0, :3, 0, :2] == 0.0)
    assert np.allclose(gr[0, 3, 0, :2], [4 * 1000.0, -4 * 1000.0])
    # r = 1: depth = n4 / n3
    pe1 = np.random.default_rng(0).uniform(0.5, 1.0, (2, 4, 3, 5))
    d1, _ = co.lpg_fwd(pe1, 1, np.float64)
    assert np.allclose(d1, pe1[:, 3] / pe1[:, 2], rtol=1e-15)
    # sub-pixel offsets: r=2 -> +-0.25 ; r=8 -> +-0.0625 ... +-0.4375
    pe2 = np.zeros((1, 4, 1, 1)); pe2[0, 0] = 1.0; pe2[0, 3] = 1.0
    d2, _ = co.lpg_fwd(pe2, 2, np.float64)
    assert np.allclose(1.0 / d2[0, 0], [-0.25, 0.25])
    d8, _ = co.lpg_fwd(pe2, 8, np.float64)
    assert np.allclose(1.0 / d8[0, 0], (np.arange(8) - 3.5) / 8)


SILOG = golden_cases("silog")


@pytest.mark.parametrize("g", SILOG, ids=lambda g: g.name)
@pytest.mark.parametrize("tag,dt", [("f32", torch.float32), ("f64", torch.float64)])
def test_torch_port_silog(g, tag, dt):
    """oracle silog_forward == the imported reference silog_loss (pytorch/bts.py:41-48) on its golden vectors."""
    assert len(SILOG) == 3
    est = torch.from_numpy(g["depth_est"]).to(dt).requires_grad_(True)
    loss = tp.silog_forward(est, torch.from_numpy(g["depth_gt"]).to(dt), torch.from_numpy(g["mask"]), g.meta["variance_focus"])
    loss.backward()
    tol = 1e-6 if dt == torch.float32 else 1e-13
    assert abs(float(loss) - float(g[tag + "_loss"])) <= tol * abs(float(g[tag + "_loss"]))
    assert max_rel_to_peak(est.grad.numpy(), g[tag + "_grad_est"]) <= tol
