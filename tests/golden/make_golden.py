#!/usr/bin/env python
"""Generate golden input/output vectors for the reduction_1x1 + local_planar_guidance hot path
by RUNNING THE UNMODIFIED REFERENCE (/root/reference/pytorch/bts.py) on CPU.

Run in the build container only (the reference tree does not travel to the GPU box):

    python tests/golden/make_golden.py            # rewrites tests/golden/*.npz

The reference hard-codes `.cuda()` at pytorch/bts.py:157,160; on a CPU-only machine we run it
under a context manager that makes Tensor.cuda the identity. Nothing else
is patched. Each case is run twice: fp32 (the canonical oracle) and fp64 (the arbiter).

Every .npz holds, for one case:
  meta (json string), x, w0..wL-1, grad_out (upstream gradient),
  f32_* / f64_* : out (reduction_1x1 output), plane_eq (after bts.py:250-253 glue), depth,
                  abs_min, grad_x, grad_w0.., and for standalone-LPG cases grad_plane_eq.
"""
import contextlib
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference/pytorch"
HERE = os.path.dirname(os.path.abspath(__file__))


@contextlib.contextmanager
def cuda_is_identity():
    orig = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.Tensor.cuda = orig


def load_ref():
    sys.path.insert(0, REF)
    import bts as ref  # noqa: the reference module itself
    return ref


def run_chain_case(ref, name, num_in, num_out, max_depth, is_final, upratio, B, h, w, seed, dtype,
                   x_np=None, w_np=None, g_np=None):
    torch.manual_seed(seed)
    red = ref.reduction_1x1(num_in, num_out, max_depth, is_final=is_final)
    red.apply(ref.weights_init_xavier)  # as bts_main.py:300
    if w_np is not None:
        with torch.no_grad():
            for p, wn in zip(red.parameters(), w_np):
                p.copy_(torch.from_numpy(wn).reshape(p.shape))
    red = red.to(dtype)
    x = torch.randn(B, num_in, h, w) if x_np is None else torch.from_numpy(x_np)
    x = x.to(dtype).requires_grad_(True)
    res = {}
    with cuda_is_identity():
        out = red(x)
        if is_final:
            g = torch.randn(B, 1, h, w) if g_np is None else torch.from_numpy(g_np)
            final = out
        else:
            lpg = ref.local_planar_guidance(upratio)
            n = F.normalize(out[:, :3, :, :], 2, 1)           # bts.py:250-251
            d = out[:, 3, :, :]                                # bts.py:252
            plane_eq = torch.cat([n, d.unsqueeze(1)], 1)       # bts.py:253
            depth = lpg(plane_eq, torch.full((B,), 518.8579, dtype=torch.float64))  # bts.py:254
            g = torch.randn(B, h * upratio, w * upratio) if g_np is None else torch.from_numpy(g_np)
            final = depth
            res["plane_eq"] = plane_eq.detach().numpy()
            res["depth"] = depth.detach().numpy()
            res["abs_min"] = np.asarray(lpg.abs_min.detach().numpy())
        final.backward(g.to(dtype))
    res["out"] = out.detach().numpy()
    res["grad_x"] = x.grad.numpy()
    ws = [p.detach().numpy().reshape(p.shape[0], p.shape[1]) for p in red.parameters()]
    for i, p in enumerate(red.parameters()):
        res["grad_w%d" % i] = p.grad.numpy().reshape(p.shape[0], p.shape[1])
    return x.detach().numpy(), ws, g.numpy(), res, [k for k, _ in red.state_dict().items()]


def chain_case(ref, name, num_in, num_out, max_depth, is_final, upratio, B, h, w, seed):
    x32, w32, g32, r32, keys = run_chain_case(ref, name, num_in, num_out, max_depth, is_final, upratio,
                                              B, h, w, seed, torch.float32)
    _, _, _, r64, _ = run_chain_case(ref, name, num_in, num_out, max_depth, is_final, upratio,
                                     B, h, w, seed, torch.float64, x_np=x32, w_np=w32, g_np=g32)
    meta = dict(kind="chain", name=name, num_in=num_in, num_out=num_out, max_depth=max_depth,
                is_final=is_final, upratio=upratio, B=B, h=h, w=w, seed=seed,
                dims=[num_in] + [int(a.shape[0]) for a in w32], state_dict_keys=keys,
                reference="pytorch/bts.py:97-173,249-254", torch=torch.__version__)
    out = dict(meta=json.dumps(meta), x=x32, grad_out=g32)
    for i, a in enumerate(w32):
        out["w%d" % i] = a
    for k, v in r32.items():
        out["f32_" + k] = v
    for k, v in r64.items():
        out["f64_" + k] = v
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if k != "meta"})


def lpg_case(ref, name, upratio, plane_eq, g, note):
    out = dict(plane_eq=plane_eq.astype(np.float32), grad_out=g.astype(np.float32))
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        pe = torch.from_numpy(out["plane_eq"]).to(dt).requires_grad_(True)
        lpg = ref.local_planar_guidance(upratio)
        with cuda_is_identity():
            depth = lpg(pe, None)
            # non-finite upstream products are the caller's problem; the reference autograd
            # still defines them, so keep whatever it returns.
            depth.backward(torch.from_numpy(out["grad_out"]).to(dt))
        out[tag + "_depth"] = depth.detach().numpy()
        out[tag + "_abs_min"] = np.asarray(lpg.abs_min.detach().numpy())
        out[tag + "_grad_plane_eq"] = pe.grad.numpy()
    B, _, h, w = plane_eq.shape
    meta = dict(kind="lpg", name=name, upratio=upratio, B=B, h=h, w=w, note=note,
                reference="pytorch/bts.py:138-173", torch=torch.__version__)
    out["meta"] = json.dumps(meta)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if k != "meta"})


def planes_from_angles(theta, phi, dist):
    return np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta), dist], 1)


def main():
    ref = load_ref()
    torch.set_num_threads(1)  # deterministic summation order in oneDNN
    # --- fused chain + glue + LPG cases: the four decoder instances (bts.py:198-217), small spatial size
    chain_case(ref, "chain_8x8_128_128", 128, 128, 10.0, False, 8, 2, 5, 7, 11)
    chain_case(ref, "chain_4x4_128_64", 128, 64, 10.0, False, 4, 2, 6, 10, 12)
    chain_case(ref, "chain_2x2_64_32", 64, 32, 80.0, False, 2, 2, 8, 14, 13)
    chain_case(ref, "chain_head_32_16", 32, 16, 80.0, True, 1, 2, 16, 24, 14)
    # bts_size=256 variants (chain is data) and a short chain
    chain_case(ref, "chain_8x8_64_64", 64, 64, 10.0, False, 8, 1, 4, 6, 15)
    chain_case(ref, "chain_2x2_32_16", 32, 16, 10.0, False, 2, 1, 5, 9, 16)
    chain_case(ref, "chain_head_16_8", 16, 8, 10.0, True, 1, 1, 7, 12, 17)
    chain_case(ref, "chain_head_8_4", 8, 4, 10.0, True, 1, 1, 4, 8, 18)   # 8->1 directly
    # ragged: tile-unfriendly spatial sizes (odd w) exercise the masked / unaligned kernel paths
    chain_case(ref, "chain_4x4_128_64_ragged", 128, 64, 10.0, False, 4, 1, 3, 5, 19)
    chain_case(ref, "chain_2x2_64_32_ragged", 64, 32, 10.0, False, 2, 3, 3, 7, 20)

    # --- standalone LPG cases
    rng = np.random.default_rng(21)
    for r in (8, 4, 2, 1):
        B, h, w = 2, 5, 6
        theta = rng.uniform(0, 1, (B, h, w)) * np.pi / 3
        phi = rng.uniform(0, 1, (B, h, w)) * 2 * np.pi
        dist = rng.uniform(0.1, 10, (B, h, w))
        pe = planes_from_angles(theta, phi, dist)
        g = rng.standard_normal((B, h * r, w * r))
        lpg_case(ref, "lpg_r%d_random" % r, r, pe, g, "random in-range planes")
    # adversarial: theta ~ U(0.95,1)*pi/3 at r=8 makes the denominator cross zero
    B, h, w = 2, 12, 16
    theta = rng.uniform(0.95, 1.0, (B, h, w)) * np.pi / 3
    phi = rng.uniform(0, 2 * np.pi, (B, h, w))
    dist = rng.uniform(0.1, 80, (B, h, w))
    pe = planes_from_angles(theta, phi, dist)
    g = rng.standard_normal((B, h * 8, w * 8))
    lpg_case(ref, "lpg_r8_adversarial", 8, pe, g, "denominator crosses zero; clamp fires")
    # planted exact clamp / zero / sign cases: n1=n2=0 so den == n3 exactly for every sub-pixel
    vals = np.array([5e-4, -5e-4, 1e-3, -1e-3, 0.0, 9.999e-4, -9.999e-4, 2e-3, -2e-3, 1.0, 1e-30, -1e-30],
                    dtype=np.float64)
    pe = np.zeros((1, 4, 3, 4))
    pe[0, 2] = vals.reshape(3, 4)
    pe[0, 3] = np.arange(1, 13).reshape(3, 4)
    for r in (8, 2):
        g = np.ones((1, 3 * r, 4 * r))
        lpg_case(ref, "lpg_r%d_planted" % r, r, pe, g, "den=n3 exactly: clamp edges, zero -> inf")

    # --- silog_loss: the imported reference class, pytorch/bts.py:41-48, three mask densities
    for name, shape, frac, vf, seed in (("silog_dense", (2, 1, 24, 32), 0.9, 0.85, 31), ("silog_half", (3, 1, 17, 23), 0.5, 0.85, 32),
                                        ("silog_sparse", (2, 1, 40, 56), 0.05, 0.5, 33)):
        silog_case(ref, name, shape, frac, vf, seed)

    # --- upconv: the imported reference class, pytorch/bts.py:83-94, forward; ragged tiles, odd
    # widths, one tile smaller than the kernel's 8 x 16
    for name, cin, cout, B, h, w, seed in (("upconv_64_32", 64, 32, 2, 11, 19, 41), ("upconv_128_64", 128, 64, 1, 8, 16, 42),
                                           ("upconv_32_32_tiny", 32, 32, 2, 3, 5, 43), ("upconv_48_96_odd", 48, 96, 1, 9, 17, 44)):
        upconv_case(ref, name, cin, cout, B, h, w, seed)


def upconv_case(ref, name, cin, cout, B, h, w, seed):
    torch.manual_seed(seed)
    mod = ref.upconv(cin, cout)                                  # the reference module itself (default Conv2d init)
    x = torch.randn(B, cin, h, w)
    gy = torch.randn(B, cout, 2 * h, 2 * w)
    out = dict(x=x.numpy(), w=mod.conv.weight.detach().numpy(), grad_y=gy.numpy())
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        m = mod.to(dt)
        m.zero_grad()
        xd = x.detach().clone().to(dt).requires_grad_(True)
        y = m(xd)
        y.backward(gy.to(dt))
        out[tag + "_y"] = y.detach().numpy()
        if tag == "f64":                                         # the arbiter's gradients (the fp32 ones are not used by any test)
            out[tag + "_grad_x"] = xd.grad.numpy()
            out[tag + "_grad_w"] = m.conv.weight.grad.numpy().copy()
    meta = dict(kind="upconv", name=name, cin=cin, cout=cout, B=B, h=h, w=w, seed=seed, reference="pytorch/bts.py:83-94",
                torch=torch.__version__)
    out["meta"] = json.dumps(meta)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if k != "meta"})


def silog_case(ref, name, shape, frac, vf, seed):
    gen = torch.Generator().manual_seed(seed)
    est = (torch.rand(shape, generator=gen) * 9.5 + 0.5)
    gt = (torch.rand(shape, generator=gen) * 9.5 + 0.5)
    mask = torch.rand(shape, generator=gen) < frac
    out = dict(depth_est=est.numpy(), depth_gt=gt.numpy(), mask=mask.numpy())
    for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
        e = est.detach().clone().to(dt).requires_grad_(True)
        loss = ref.silog_loss(variance_focus=vf)(e, gt.to(dt), mask)        # the reference module itself
        loss.backward()
        out[tag + "_loss"] = np.asarray(loss.detach().numpy())
        out[tag + "_grad_est"] = e.grad.numpy()
    meta = dict(kind="silog", name=name, shape=list(shape), mask_fraction=frac, variance_focus=vf, seed=seed,
                reference="pytorch/bts.py:41-48", torch=torch.__version__)
    out["meta"] = json.dumps(meta)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if k != "meta"})


if __name__ == "__main__":
    main()
