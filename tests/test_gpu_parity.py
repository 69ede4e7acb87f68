"""GPU (-m gpu): the CUDA path, called through the C ABI, against the golden vectors of the
unmodified reference, against the oracle on fresh seeded inputs, and against size-independent
properties at BASELINE.json's full sizes.

Tolerances (north_star: 1e-5 relative in fp32):
  depth ............ element-wise rtol 1e-5 (+ atol 1e-6*max_depth)
  plane_eq, grads .. max|a-b| <= 1e-5 * max|b|  and  L2-relative <= 1e-5   (they cross zero)
  standalone LPG forward: BIT-IDENTICAL (same IEEE operation order as the reference).
"""
import numpy as np
import pytest
import torch

from conftest import golden_cases, max_rel_to_peak, l2_rel
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu

RTOL = 1e-5
CHAIN = golden_cases("chain")
LPG = golden_cases("lpg")


def dev(a, requires_grad=False):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t.requires_grad_(True) if requires_grad else t


def close_peak(a, b, tol=RTOL):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a
    assert max_rel_to_peak(a, b) <= tol, ("max-rel-to-peak", max_rel_to_peak(a, b))
    assert l2_rel(a, b) <= tol, ("l2-rel", l2_rel(a, b))


def close_depth(a, b, max_depth):
    a = a.detach().cpu().numpy()
    assert np.allclose(a, b, rtol=RTOL, atol=1e-6 * max_depth), float(np.abs(a - b).max())


@pytest.mark.parametrize("g", LPG, ids=lambda g: g.name)
def test_lpg_module_vs_golden(g):
    import bts_b200
    r = g.meta["upratio"]
    pe = dev(g["plane_eq"], True)
    m = bts_b200.local_planar_guidance(r)
    depth = m(pe, torch.full((pe.shape[0],), 518.8579, dtype=torch.float64).cuda())
    assert np.array_equal(depth.detach().cpu().numpy(), g["f32_depth"], equal_nan=True)   # bit-exact
    assert np.array_equal(m.abs_min.cpu().numpy(), g["f32_abs_min"], equal_nan=True)
    depth.backward(dev(g["grad_out"]))
    got, ref = pe.grad.cpu().numpy(), g["f64_grad_plane_eq"].astype(np.float64)
    assert np.array_equal(np.isnan(got), np.isnan(ref)) and np.array_equal(np.isinf(got), np.isinf(ref))
    fin = np.isfinite(ref)
    # vs the reference's own fp32 result: 1e-5. vs fp64: 1e-5, or the reference-fp32's own distance
    # from fp64 when the case is ill-conditioned (adversarial planes put |den| near the 1e-3 clamp,
    # where the reference itself is 2e-5 away from its fp64 run).
    ref32 = g["f32_grad_plane_eq"].astype(np.float64)
    close_peak(got[fin], ref32[fin])
    own = max_rel_to_peak(ref32[fin], ref[fin])
    close_peak(got[fin], ref[fin], tol=max(RTOL, 2 * own))


@pytest.mark.parametrize("g", CHAIN, ids=lambda g: g.name)
def test_strict_dropin_vs_golden(g):
    """reduction_1x1 -> (reference's own Python glue, bts.py:250-253) -> local_planar_guidance."""
    import bts_b200
    import torch.nn.functional as F
    m = g.meta
    red = bts_b200.reduction_1x1(m["num_in"], m["num_out"], m["max_depth"], m["is_final"]).cuda()
    with torch.no_grad():
        for p, w in zip(red.chain_weights(), g.weights):
            p.copy_(dev(w).view_as(p))
    x = dev(g["x"], True)
    out = red(x)
    close_peak(out, g["f64_out"])
    if m["is_final"]:
        final = out
    else:
        lpg = bts_b200.local_planar_guidance(m["upratio"])
        n = F.normalize(out[:, :3, :, :], 2, 1)
        pe = torch.cat([n, out[:, 3, :, :].unsqueeze(1)], 1)
        final = lpg(pe, None)
        close_depth(final, g["f32_depth"], m["max_depth"])
        assert abs(float(lpg.abs_min) - float(g["f32_abs_min"])) <= 1e-5
    final.backward(dev(g["grad_out"]))
    close_peak(x.grad, g["f64_grad_x"])
    for i, p in enumerate(red.chain_weights()):
        close_peak(p.grad.view(p.shape[0], p.shape[1]), g["f64_grad_w%d" % i])


@pytest.mark.parametrize("g", [c for c in CHAIN if not c.meta["is_final"]], ids=lambda g: g.name)
def test_fused_stage_vs_golden(g):
    import bts_b200
    m = g.meta
    red = bts_b200.reduction_1x1(m["num_in"], m["num_out"], m["max_depth"]).cuda()
    with torch.no_grad():
        for p, w in zip(red.chain_weights(), g.weights):
            p.copy_(dev(w).view_as(p))
    lpg = bts_b200.local_planar_guidance(m["upratio"])
    x = dev(g["x"], True)
    depth, pe = bts_b200.fused_stage(red, lpg, x, return_plane_eq=True)
    close_depth(depth, g["f32_depth"], m["max_depth"])
    close_peak(pe, g["f64_plane_eq"])
    assert abs(float(lpg.abs_min) - float(g["f32_abs_min"])) <= 1e-5
    depth.backward(dev(g["grad_out"]))
    close_peak(x.grad, g["f64_grad_x"])
    for i, p in enumerate(red.chain_weights()):
        close_peak(p.grad.view(p.shape[0], p.shape[1]), g["f64_grad_w%d" % i])


SEEDED = [  # (num_in, num_out, is_final, r, B, h, w, max_depth): sizes the C oracle finishes in seconds
    (128, 128, False, 8, 2, 15, 20, 10.0),
    (128, 64, False, 4, 2, 30, 40, 10.0),
    (64, 32, False, 2, 2, 44, 76, 80.0),
    (32, 16, True, 1, 2, 88, 152, 80.0),
    (64, 64, False, 8, 1, 11, 13, 10.0),
    (64, 64, False, 8, 1, 17, 9, 10.0),      # hw % 4 != 0: recompute backward + non-TMA weight-gradient kernel
    (48, 24, False, 4, 1, 9, 7, 10.0),       # widths no fast path covers: generic kernel
    (16, 16, True, 1, 3, 33, 31, 10.0),
    # the head's register-operand backward (chain_bwd_mma.cu): 32-pixel groups that straddle images, a ragged last
    # group, images smaller than a group (several wraps inside one group), a tensor smaller than a group
    (32, 16, True, 1, 3, 6, 10, 10.0),
    (32, 16, True, 1, 5, 7, 12, 80.0),
    (32, 16, True, 1, 7, 3, 4, 10.0),
    (32, 16, True, 1, 1, 2, 2, 10.0),
    (32, 16, True, 1, 2, 9, 7, 10.0),        # hw % 4 != 0: no 16-byte copies, the tcgen05 / generic head kernels
]


@pytest.mark.parametrize("cfg", SEEDED, ids=lambda c: "x".join(map(str, c[:7])))
def test_fused_vs_c_oracle_seeded(cfg):
    import bts_b200
    from oracle import torch_port as tp
    num_in, num_out, is_final, r, B, h, w, md = cfg
    gen = torch.Generator().manual_seed(1234 + num_in + r)
    dims = tp.chain_dims(num_in, num_out, is_final)
    ws = [t[:, :, 0, 0].numpy() for t in tp.xavier_weights(dims, is_final, gen)]
    x = torch.randn(B, num_in, h, w, generator=gen).numpy()
    red = bts_b200.reduction_1x1(num_in, num_out, md, is_final).cuda()
    with torch.no_grad():
        for p, wt in zip(red.chain_weights(), ws):
            p.copy_(dev(wt).view_as(p))
    xd = dev(x, True)
    if is_final:
        g = torch.randn(B, 1, h, w, generator=gen).numpy()
        out = red(xd)
        close_peak(out, co.reduc_fwd(x, ws, md, True, np.float64))
        out.backward(dev(g))
        gx, gw = co.reduc_bwd(x, ws, md, True, g, np.float64)
    else:
        g = torch.randn(B, h * r, w * r, generator=gen).numpy()
        lpg = bts_b200.local_planar_guidance(r)
        depth = bts_b200.fused_stage(red, lpg, xd)
        ref = co.fused_fwd(x, ws, md, r, np.float64)
        close_depth(depth, ref["depth"], md)
        assert abs(float(lpg.abs_min) - float(ref["abs_min"])) <= 1e-5
        depth.backward(dev(g))
        gx, gw = co.fused_bwd(x, ws, md, r, g, np.float64)
    close_peak(xd.grad, gx)
    for p, gr in zip(red.chain_weights(), gw):
        close_peak(p.grad.view(p.shape[0], p.shape[1]), gr)


@pytest.mark.parametrize("r,B,h,w", [(8, 8, 60, 80), (4, 8, 120, 160), (2, 8, 240, 320), (8, 16, 44, 152),
                                     (3, 1, 5, 7), (8, 1, 3, 5), (2, 2, 7, 9), (1, 2, 6, 10), (1, 1, 5, 4)])
def test_lpg_full_size_vs_c_oracle(r, B, h, w):
    """BASELINE.json config shapes (NYU 480x640 b8, KITTI 352x1216 b16) and ragged / odd-ratio shapes
    that take the scalar kernels. Adversarial planes: the clamp fires."""
    from bts_b200 import functional as Fn
    rng = np.random.default_rng(r * 1000 + w)
    theta = rng.uniform(0.9, 1.0, (B, h, w)) * np.pi / 3
    phi = rng.uniform(0, 2 * np.pi, (B, h, w))
    pe = np.stack([np.sin(theta) * np.cos(phi), np.sin(theta) * np.sin(phi), np.cos(theta),
                   rng.uniform(0.1, 80, (B, h, w))], 1).astype(np.float32)
    g = rng.standard_normal((B, h * r, w * r)).astype(np.float32)
    ped = dev(pe, True)
    depth, amin = Fn.lpg(ped, r)
    ref_d, ref_a = co.lpg_fwd(pe, r, np.float32)
    assert np.array_equal(depth.detach().cpu().numpy(), ref_d)
    assert float(amin) == float(ref_a)
    depth.backward(dev(g))
    # |den| sits near the 1e-3 clamp for these planes, so fp32 itself is ill-conditioned: allow the
    # fp32 oracle's own distance from the fp64 oracle (both follow the reference's formulas).
    g64, g32 = co.lpg_bwd(g, pe, r, np.float64), co.lpg_bwd(g, pe, r, np.float32)
    close_peak(ped.grad, g64, tol=max(RTOL, 4 * max_rel_to_peak(g32, g64)))


def test_lpg_properties_full_size():
    """Size-independent properties at the largest config (KITTI 352x1216, batch 16)."""
    from bts_b200 import functional as Fn
    B, h, w, r = 16, 44, 152, 8
    gen = torch.Generator(device="cuda").manual_seed(7)
    theta = torch.rand(B, h, w, generator=gen, device="cuda") * (np.pi / 3) * 0.8
    phi = torch.rand(B, h, w, generator=gen, device="cuda") * 2 * np.pi
    n4 = torch.rand(B, h, w, generator=gen, device="cuda") * 80 + 0.1
    pe = torch.stack([theta.sin() * phi.cos(), theta.sin() * phi.sin(), theta.cos(), n4], 1).contiguous()
    d1, a1 = Fn.lpg(pe, r)
    # (1) linear in n4:  lpg(n, 2*n4) == 2*lpg(n, n4) exactly (power-of-two scaling commutes with rounding)
    pe2 = pe.clone(); pe2[:, 3] *= 2
    d2, a2 = Fn.lpg(pe2, r)
    assert torch.equal(d2, 2 * d1) and torch.equal(a1, a2)
    # (2) a fronto-parallel plane (n1=n2=0) yields constant depth n4/n3 over each r x r block
    pe3 = pe.clone(); pe3[:, 0] = 0; pe3[:, 1] = 0
    d3, a3 = Fn.lpg(pe3, r)
    blk = (pe3[:, 3] / pe3[:, 2]).repeat_interleave(r, 1).repeat_interleave(r, 2)
    assert torch.equal(d3, blk) and float(a3) == float(pe3[:, 2].abs().min())
    # (3) the mean of 1/depth over a block is n3/n4 (the sub-pixel offsets are symmetric)
    inv = (1.0 / d1.double()).view(B, h, r, w, r).mean(dim=(2, 4))
    assert torch.allclose(inv, (pe[:, 2] / pe[:, 3]).double(), rtol=1e-5, atol=1e-7)
    # (4) deterministic
    d1b, _ = Fn.lpg(pe, r)
    assert torch.equal(d1, d1b)


def test_abs_min_nan_and_empty():
    from bts_b200 import functional as Fn
    pe = torch.rand(1, 4, 4, 4, device="cuda") + 0.5
    pe[0, 2, 1, 1] = float("nan")
    _, amin = Fn.lpg(pe, 2)
    assert torch.isnan(amin)                      # torch.min propagates NaN (bts.py:167)
    depth, amin = Fn.lpg(torch.empty(0, 4, 4, 4, device="cuda"), 8)
    assert tuple(depth.shape) == (0, 32, 32) and float(amin) == float("inf")


def test_chain_linearity_of_backward_full_size():
    """NYU config-2 head shape [8,32,480,640]: backward is linear in the upstream gradient, and
    grad_W of a batch equals the sum of per-sample grad_W (batch sharding needs no collective)."""
    import bts_b200
    torch.manual_seed(3)
    red = bts_b200.reduction_1x1(32, 16, 10.0, True).cuda()
    for p in red.parameters():
        torch.nn.init.xavier_uniform_(p)
    x = torch.randn(8, 32, 480, 640, device="cuda")
    g = torch.randn(8, 1, 480, 640, device="cuda")

    def run(xx, gg):
        xx = xx.clone().requires_grad_(True)
        red.zero_grad()
        out = red(xx)
        out.backward(gg)
        return out.detach(), xx.grad, [p.grad.clone() for p in red.parameters()]

    out, gx, gw = run(x, g)
    assert out.min() > 0 and out.max() < 1
    _, gx2, gw2 = run(x, 2 * g)
    assert torch.equal(gx2, 2 * gx)
    for a, b in zip(gw2, gw):
        assert torch.allclose(a, 2 * b, rtol=1e-4, atol=1e-4 * float(b.abs().max()))   # fp32 atomics order
    acc = [torch.zeros_like(p) for p in gw]
    for s in range(0, 8, 4):
        _, gxs, gws = run(x[s:s + 4], g[s:s + 4])
        assert torch.equal(gxs, gx[s:s + 4])
        for a, b in zip(acc, gws):
            a += b
    for a, b in zip(acc, gw):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-4 * float(b.abs().max()))


def test_noncontiguous_and_state_dict_roundtrip():
    import bts_b200
    red = bts_b200.reduction_1x1(64, 32, 10.0).cuda()
    sd = {k: v.clone() for k, v in red.state_dict().items()}
    red2 = bts_b200.reduction_1x1(64, 32, 10.0).cuda()
    red2.load_state_dict(sd)
    x = torch.randn(2, 6, 10, 64, device="cuda").permute(0, 3, 1, 2)      # channels-last strides
    assert not x.is_contiguous()
    assert torch.equal(red(x), red2(x.contiguous()))


def test_ddp_wraps_the_modules():
    """DDP(find_unused_parameters=True) as bts_main.py:314 -- weight grads must reach the reducer hooks."""
    import os
    import torch.distributed as dist
    import bts_b200
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    created = False
    if not dist.is_initialized():
        dist.init_process_group("nccl", rank=0, world_size=1)
        created = True
    try:
        class Stage(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.reduc = bts_b200.reduction_1x1(64, 32, 10.0)
                self.lpg = bts_b200.local_planar_guidance(2)

            def forward(self, x):
                return bts_b200.fused_stage(self.reduc, self.lpg, x)

        m = torch.nn.parallel.DistributedDataParallel(Stage().cuda(), device_ids=[0], find_unused_parameters=True)
        out = m(torch.randn(2, 64, 8, 8, device="cuda"))
        out.mean().backward()
        assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())
    finally:
        if created:
            dist.destroy_process_group()


@pytest.mark.parametrize("cfg", [(128, 128, False, 8, 3, 19, 23), (128, 64, False, 4, 2, 33, 41), (64, 32, False, 2, 2, 50, 70),
                                 (32, 16, True, 1, 2, 100, 130), (64, 64, False, 8, 1, 17, 9), (32, 32, False, 4, 1, 40, 40),
                                 (32, 16, False, 2, 1, 31, 33), (128, 128, False, 1, 1, 9, 20), (128, 128, False, 8, 3, 20, 28),
                                 (128, 64, False, 4, 2, 36, 44), (64, 32, False, 2, 3, 60, 84), (32, 16, True, 1, 2, 96, 160)],
                         ids=lambda c: "x".join(map(str, c)))
def test_tensor_core_path_vs_generic_path(cfg):
    """The tcgen05 kernels and the generic SIMT kernels are independent fp32 implementations (different summation
    orders, ex2.approx vs expf in the ELU): each is held to 1e-5 of the fp64 oracle by the tests above, so they must
    agree with EACH OTHER to 2e-5, on shapes with partial tiles and image wraps. (bts_force_path is process-wide
    since round 2, so the generic run really takes the SIMT backward too.)"""
    import bts_b200
    from bts_b200 import functional as Fn
    num_in, num_out, is_final, r, B, h, w = cfg
    dims = bts_b200.chain_channels(num_in, num_out, is_final)
    assert Fn.chain_path(dims, is_final) == Fn.PATH_TCGEN05
    torch.manual_seed(5)
    red = bts_b200.reduction_1x1(num_in, num_out, 10.0, is_final).cuda()
    for p in red.parameters():
        torch.nn.init.xavier_uniform_(p)
    x = torch.randn(B, num_in, h, w, device="cuda")
    g = torch.randn((B, 1, h, w) if is_final else (B, h * r, w * r), device="cuda")

    def run():
        xx = x.clone().requires_grad_(True)
        red.zero_grad()
        if is_final:
            out, amin = red(xx), None
        else:
            lpg = bts_b200.local_planar_guidance(r)
            out = bts_b200.fused_stage(red, lpg, xx)
            amin = float(lpg.abs_min)
        out.backward(g)
        return out.detach(), amin, xx.grad, [p.grad.clone() for p in red.parameters()]

    try:
        Fn.force_path(Fn.PATH_GENERIC)
        ref = run()
    finally:
        Fn.force_path(Fn.PATH_AUTO)
    got = run()
    if is_final:
        close_peak(got[0], ref[0].cpu().numpy(), tol=2 * RTOL)
    else:
        assert torch.allclose(got[0], ref[0], rtol=2 * RTOL, atol=1e-5)
        assert abs(got[1] - ref[1]) <= 1e-5
    close_peak(got[2], ref[2].cpu().numpy(), tol=2 * RTOL)
    for a, b in zip(got[3], ref[3]):
        close_peak(a, b.cpu().numpy(), tol=2 * RTOL)


@pytest.mark.parametrize("cfg", [(128, 128, False, 8, 3, 20, 28, True), (128, 128, False, 8, 3, 19, 23, False),
                                 (128, 64, False, 4, 2, 36, 44, True), (64, 64, False, 8, 1, 16, 24, True),
                                 (32, 32, False, 4, 1, 40, 40, True), (128, 128, False, 0, 2, 24, 20, True),
                                 (64, 64, True, 0, 2, 12, 36, True), (128, 64, False, 1, 1, 8, 16, True),
                                 (64, 32, False, 2, 2, 32, 32, False)],
                         ids=lambda c: "x".join(map(str, c)))
def test_stash_backward_vs_recompute_backward(cfg):
    """ABI 2: the forward may leave the wide chains' activations in a stash so that backward skips the
    recompute. Both backward routes must agree to rounding; shapes cover padded pixel axes (hw % 32 != 0),
    hw % 4 != 0 (no stash: identical route), the plain reduction (r = 0) and a wide sigmoid head."""
    import bts_b200
    from bts_b200 import _cabi, functional as Fn
    num_in, num_out, is_final, r, B, h, w, stashes = cfg
    dims = bts_b200.chain_channels(num_in, num_out, is_final)
    nbytes = _cabi.lib().bts_stash_bytes(_cabi.int_array(dims), len(dims) - 1, B, h, w)
    assert (nbytes > 0) == stashes
    torch.manual_seed(11)
    red = bts_b200.reduction_1x1(num_in, num_out, 10.0, is_final).cuda()
    for p in red.parameters():
        torch.nn.init.xavier_uniform_(p)
    x = torch.randn(B, num_in, h, w, device="cuda")
    if is_final:
        g = torch.randn(B, 1, h, w, device="cuda")
    elif r == 0:
        g = torch.randn(B, 4, h, w, device="cuda")
    else:
        g = torch.randn(B, h * r, w * r, device="cuda")

    def run():
        xx = x.clone().requires_grad_(True)
        red.zero_grad()
        out = red(xx) if (is_final or r == 0) else bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), xx)
        out.backward(g)
        return out.detach(), xx.grad, [p.grad.clone() for p in red.parameters()]

    prev = Fn.set_stash(False)
    try:
        ref = run()
    finally:
        Fn.set_stash(prev)
    got = run()
    assert torch.equal(got[0], ref[0])                      # the stash does not touch the forward result
    close_peak(got[1], ref[1].cpu().numpy())
    for a, b in zip(got[2], ref[2]):
        close_peak(a, b.cpu().numpy())


@pytest.mark.parametrize("cfg", [(128, 128, 8, 2, 28, 36), (128, 64, 4, 2, 36, 44), (64, 32, 2, 2, 40, 56)],
                         ids=lambda c: "x".join(map(str, c)))
def test_cuda_graph_replay_matches_eager_launches(cfg):
    """The C ABI only enqueues on the caller's stream, so a whole forward + backward can be captured into a CUDA
    graph (bench.py does). The wide-chain backward chains its four kernels by programmatic dependent launch
    (common.cuh launch_dependent): stream capture must record those edges and replays must return what eager
    launches return -- bit for bit where the summation order is fixed (depth, grad_x, the wide chains' grad_W),
    to rounding for the narrow chain's atomically accumulated grad_W."""
    import bts_b200
    num_in, num_out, r, B, h, w = cfg
    torch.manual_seed(5)
    red = bts_b200.reduction_1x1(num_in, num_out, 10.0).cuda()
    for p in red.parameters():
        torch.nn.init.xavier_uniform_(p)
    lpg = bts_b200.local_planar_guidance(r)
    x = torch.randn(B, num_in, h, w, device="cuda", requires_grad=True)
    g = torch.randn(B, h * r, w * r, device="cuda")
    params = list(red.parameters())

    def step():
        return torch.autograd.grad(bts_b200.fused_stage(red, lpg, x), [x] + params, g)

    eager = [t.clone() for t in step()]
    depth_eager = bts_b200.fused_stage(red, lpg, x).detach().clone()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):
            step()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        depth_static = bts_b200.fused_stage(red, lpg, x)
        static = torch.autograd.grad(depth_static, [x] + params, g)
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(depth_static.detach(), depth_eager)
    assert torch.equal(static[0], eager[0])
    wide = num_in >= 128
    for a, b in zip(static[1:], eager[1:]):
        if wide:
            assert torch.equal(a, b)
        else:
            close_peak(a, b.cpu().numpy())


def _glue_reference(x, ws, md, r, g_full, g_ds):
    """The reference's own operator sequence (oracle/torch_port.py restates pytorch/bts.py:124-173, 250-254) followed
    by the post-LPG glue lines bts.py:255-256 / 269-270, in fp64 on the CPU, with autograd for the gradients."""
    from oracle import torch_port as tp
    xt = torch.from_numpy(x).double().requires_grad_(True)
    wt = [torch.from_numpy(w).double().reshape(w.shape[0], w.shape[1], 1, 1).requires_grad_(True) for w in ws]
    depth, _ = tp.stage_forward(xt, wt, md, r)
    scaled = depth.unsqueeze(1) / md
    loss = (scaled * torch.from_numpy(g_full).double()).sum()
    ds = None
    if g_ds is not None:
        ds = torch.nn.functional.interpolate(scaled, scale_factor=2.0 / r, mode="nearest")
        loss = loss + (ds * torch.from_numpy(g_ds).double()).sum()
    loss.backward()
    return (scaled.detach().numpy(), None if ds is None else ds.detach().numpy(), xt.grad.numpy(),
            [w.grad[:, :, 0, 0].numpy() for w in wt])


@pytest.mark.parametrize("cfg", [(128, 128, 8, 2, 20, 28, True, None), (128, 128, 8, 2, 19, 23, True, None),
                                 (128, 64, 4, 2, 36, 44, True, None), (64, 32, 2, 2, 40, 56, False, None),
                                 (32, 16, 4, 1, 24, 40, True, None), (128, 128, 8, 1, 16, 24, True, "generic"),
                                 (64, 32, 4, 2, 17, 21, True, "generic"), (128, 64, 4, 2, 36, 44, True, "nostash")],
                         ids=lambda c: "x".join(map(str, c)))
def test_fused_stage_with_post_lpg_glue(cfg):
    """`depth.unsqueeze(1) / max_depth` and the nearest down-sampling folded into the
    fused stage, forward and backward, on every route (stash, recompute, one-pass, generic)."""
    import bts_b200
    from bts_b200 import functional as Fn
    from oracle import torch_port as tp
    num_in, num_out, r, B, h, w, with_ds, route = cfg
    md = 10.0
    gen = torch.Generator().manual_seed(77 + num_in + r)
    dims = tp.chain_dims(num_in, num_out, False)
    ws = [t[:, :, 0, 0].numpy() for t in tp.xavier_weights(dims, False, gen)]
    x = torch.randn(B, num_in, h, w, generator=gen).numpy()
    g_full = torch.randn(B, 1, h * r, w * r, generator=gen).numpy()
    g_ds = torch.randn(B, 1, 2 * h, 2 * w, generator=gen).numpy() if with_ds else None
    ref_scaled, ref_ds, ref_gx, ref_gw = _glue_reference(x, ws, md, r, g_full, g_ds)

    red = bts_b200.reduction_1x1(num_in, num_out, md, False).cuda()
    with torch.no_grad():
        for p, wt in zip(red.chain_weights(), ws):
            p.copy_(dev(wt).view_as(p))
    xd = dev(x, True)
    lpg = bts_b200.local_planar_guidance(r)
    prev = Fn.set_stash(route != "nostash")
    try:
        if route == "generic":
            Fn.force_path(Fn.PATH_GENERIC)
        out = bts_b200.fused_stage(red, lpg, xd, scaled=True, downsample=with_ds)
        scaled, ds = (out if with_ds else (out, None))
        assert scaled.shape == (B, 1, h * r, w * r)
        loss = (scaled * dev(g_full)).sum()
        if with_ds:
            assert ds.shape == (B, 1, 2 * h, 2 * w)
            loss = loss + (ds * dev(g_ds)).sum()
        loss.backward()
    finally:
        Fn.force_path(Fn.PATH_AUTO)
        Fn.set_stash(prev)
    close_depth(scaled, ref_scaled, 1.0)
    if with_ds:
        close_depth(ds, ref_ds, 1.0)
        # the down-sampled map is exactly rows / columns 0 and r/2 of every block of the full map
        assert torch.equal(ds, scaled[:, :, ::r // 2, ::r // 2])
    close_peak(xd.grad, ref_gx)
    for p, gr in zip(red.chain_weights(), ref_gw):
        close_peak(p.grad.view(p.shape[0], p.shape[1]), gr)


def test_glue_writes_into_a_channel_slice():
    """C ABI, batch strides: the scaled depth lands in channel 1 of a [B,3,H,W] buffer (as concat1 of bts.py:287
    would hold it) and the backward reads its gradient from the same slice; neighbours stay untouched."""
    import bts_b200
    from bts_b200 import _cabi, functional as Fn
    torch.manual_seed(9)
    B, h, w, r, md = 2, 24, 32, 4, 10.0
    red = bts_b200.reduction_1x1(128, 64, md, False).cuda()
    for p in red.parameters():
        torch.nn.init.xavier_uniform_(p)
    x = torch.randn(B, 128, h, w, device="cuda")
    H, W = h * r, w * r
    buf = torch.full((B, 3, H, W), 7.0, device="cuda")
    gbuf = torch.randn(B, 3, H, W, device="cuda")
    # dense route through the module API
    xd = x.clone().requires_grad_(True)
    dense = bts_b200.fused_stage(red, bts_b200.local_planar_guidance(r), xd, scaled=True)
    dense.backward(gbuf[:, 1:2].contiguous())
    ref_gw = [p.grad.clone() for p in red.parameters()]
    # strided route through the C ABI
    lib = _cabi.lib()
    wts = [t.detach() for t in red.chain_weights()]
    dims = [128] + [t.shape[0] for t in wts]
    Wp, D = _cabi.ptr_array([t.data_ptr() for t in wts]), _cabi.int_array(dims)
    nws = lib.bts_workspace_bytes(D, len(wts), B, h, w, r)
    wsb = torch.empty((nws + 3) // 4, device="cuda")
    amin = torch.full((), float("inf"), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _cabi.check(lib.bts_reduc_lpg_fwd_glue(x.data_ptr(), Wp, D, len(wts), md, B, h, w, r, Fn.EPS, md, None,
                                           buf[:, 1].data_ptr(), 3 * H * W, None, 0, amin.data_ptr(),
                                           wsb.data_ptr(), nws, None, 0, st), "fwd")
    assert torch.equal(buf[:, 1:2], dense.detach())
    assert bool((buf[:, 0] == 7.0).all()) and bool((buf[:, 2] == 7.0).all())
    gx = torch.empty_like(x)
    gws = [torch.zeros_like(t) for t in wts]
    _cabi.check(lib.bts_reduc_lpg_bwd_glue(gbuf[:, 1].data_ptr(), 3 * H * W, None, 0, md, x.data_ptr(), Wp, D, len(wts),
                                           md, B, h, w, r, Fn.EPS, gx.data_ptr(),
                                           _cabi.ptr_array([t.data_ptr() for t in gws]), wsb.data_ptr(), nws, None, 0, st),
                "bwd")
    close_peak(gx, xd.grad.cpu().numpy())
    for a, b in zip(gws, ref_gw):
        close_peak(a, b.cpu().numpy())


@pytest.mark.parametrize("r,cin,cout", [(8, 128, 128), (4, 128, 64), (2, 64, 32)], ids=str)
def test_fused_stage_writes_into_a_concat_slice(r, cin, cout):
    """SURVEY.md 8(f)-1 through the module API: fused_stage(..., scaled=True, out=concat[:, 1:2]) writes the map in
    place of the torch.cat of bts.py:287 and back-propagates from the slice autograd returns; the torch.cat route
    (whose CatBackward hands a batch-strided gradient slice to the same Function) must give the same numbers."""
    import bts_b200
    torch.manual_seed(21)
    B, h, w, md = 2, 12, 20, 10.0
    H, W = h * r, w * r
    red = bts_b200.reduction_1x1(cin, cout, md, False).cuda()
    for p in red.parameters():
        torch.nn.init.xavier_uniform_(p)
    lpg = bts_b200.local_planar_guidance(r)
    x = torch.randn(B, cin, h, w, device="cuda")
    c0, c2 = torch.randn(B, 1, H, W, device="cuda"), torch.randn(B, 1, H, W, device="cuda")
    gbuf = torch.randn(B, 3, H, W, device="cuda")
    # route 1: torch.cat of a dense map (the reference's own structure)
    x1 = x.clone().requires_grad_(True)
    dense = bts_b200.fused_stage(red, lpg, x1, scaled=True)
    (torch.cat([c0, dense, c2], 1) * gbuf).sum().backward()
    gw1 = [p.grad.clone() for p in red.parameters()]
    for p in red.parameters():
        p.grad = None
    # route 2: caller-owned buffer, the stage writes its channel slice
    x2 = x.clone().requires_grad_(True)
    concat = torch.empty(B, 3, H, W, device="cuda")
    concat[:, 0:1] = c0
    concat[:, 2:3] = c2
    got = bts_b200.fused_stage(red, lpg, x2, scaled=True, out=concat[:, 1:2])
    assert got.data_ptr() == concat[:, 1:2].data_ptr()
    assert torch.equal(concat[:, 1:2].detach(), dense.detach())
    assert torch.equal(concat[:, 0:1].detach(), c0) and torch.equal(concat[:, 2:3].detach(), c2)
    (concat * gbuf).sum().backward()
    close_peak(x2.grad, x1.grad.cpu().numpy())
    for p, ref in zip(red.parameters(), gw1):
        close_peak(p.grad, ref.cpu().numpy())


@pytest.mark.parametrize("shape,frac", [((4, 1, 52, 68), 0.6), ((8, 1, 480, 640), 0.35), ((3, 1, 37, 41), 0.9),
                                        ((2, 1, 16, 16), 1.0)], ids=str)
def test_silog_loss_vs_reference_ops(shape, frac):
    """silog_loss (pytorch/bts.py:41-48) as one masked reduction; the reference's own operator
    sequence (gather by boolean mask, log, mean, sqrt) in fp64 on the CPU is the oracle, autograd its backward."""
    import bts_b200
    gen = torch.Generator().manual_seed(5)
    est = (torch.rand(shape, generator=gen) * 9.5 + 0.5)
    gt = (torch.rand(shape, generator=gen) * 9.5 + 0.5)
    mask = torch.rand(shape, generator=gen) < frac
    e64 = est.double().requires_grad_(True)
    d = torch.log(e64[mask]) - torch.log(gt.double()[mask])
    ref = torch.sqrt((d ** 2).mean() - 0.85 * (d.mean() ** 2)) * 10.0
    ref.backward()
    crit = bts_b200.silog_loss(variance_focus=0.85)
    ed = est.cuda().requires_grad_(True)
    loss = crit(ed, gt.cuda(), mask.cuda())
    assert loss.dim() == 0
    (loss * 1.7).backward()
    assert abs(float(loss) - float(ref)) <= 1e-5 * abs(float(ref))
    close_peak(ed.grad, 1.7 * e64.grad.numpy())
    assert bool((ed.grad[~mask.cuda()] == 0).all())


@pytest.mark.parametrize("g", golden_cases("silog"), ids=lambda g: g.name)
def test_silog_loss_vs_golden(g):
    """bts_b200.silog_loss against the imported reference class (pytorch/bts.py:41-48; fixtures by make_golden.py)."""
    import bts_b200
    est = dev(g["depth_est"], True)
    loss = bts_b200.silog_loss(g.meta["variance_focus"])(est, dev(g["depth_gt"]), torch.from_numpy(g["mask"]).cuda())
    loss.backward()
    ref = float(g["f64_loss"])
    assert abs(float(loss) - ref) <= RTOL * abs(ref), (float(loss), ref)
    close_peak(est.grad, g["f64_grad_est"])
    masked_out = ~g["mask"]
    assert not est.grad.cpu().numpy()[masked_out].any()            # no gradient outside the mask, as the gather gives


def test_silog_loss_edge_cases():
    import bts_b200
    crit = bts_b200.silog_loss(0.85)
    est = torch.rand(2, 1, 9, 7, device="cuda") + 0.5            # odd size: scalar tail, unaligned row ends
    gt = torch.rand(2, 1, 9, 7, device="cuda") + 0.5
    none = torch.zeros(2, 1, 9, 7, dtype=torch.bool, device="cuda")
    assert torch.isnan(crit(est, gt, none))                        # mean of an empty selection, as the reference
    one = none.clone()
    one[1, 0, 3, 4] = True
    # a single pixel: mean(d^2) - vf * mean(d)^2 = (1 - vf) d^2
    d = float(torch.log(est[1, 0, 3, 4]) - torch.log(gt[1, 0, 3, 4]))
    assert abs(float(crit(est, gt, one)) - 10.0 * (0.15 * d * d) ** 0.5) <= 1e-5 * (10.0 * (0.15 * d * d) ** 0.5) + 1e-7
    with pytest.raises(RuntimeError):
        crit(est, gt, one.float())


def _compute_errors_ref(gt, pred):
    """Restatement of compute_errors, pytorch/bts_main.py:87-108 (numpy, evaluated in fp64 here)."""
    gt, pred = gt.astype(np.float64), pred.astype(np.float64)
    thresh = np.maximum(gt / pred, pred / gt)
    d1, d2, d3 = (thresh < 1.25).mean(), (thresh < 1.25 ** 2).mean(), (thresh < 1.25 ** 3).mean()
    rms = np.sqrt(((gt - pred) ** 2).mean())
    log_rms = np.sqrt(((np.log(gt) - np.log(pred)) ** 2).mean())
    abs_rel = np.mean(np.abs(gt - pred) / gt)
    sq_rel = np.mean(((gt - pred) ** 2) / gt)
    err = np.log(pred) - np.log(gt)
    silog = np.sqrt(np.mean(err ** 2) - np.mean(err) ** 2) * 100
    log10 = np.mean(np.abs(np.log10(pred) - np.log10(gt)))
    return [silog, abs_rel, log10, rms, sq_rel, log_rms, d1, d2, d3]


@pytest.mark.parametrize("B,H,W,crop,lo,hi", [(3, 480, 640, (45, 471, 41, 601), 1e-3, 10.0), (2, 352, 1216, None, 1e-3, 80.0),
                                              (1, 37, 53, (5, 30, 7, 50), 0.5, 8.0)], ids=str)
def test_eval_metrics_vs_reference(B, H, W, crop, lo, hi):
    """online_eval's clamp + validity mask + compute_errors (bts_main.py:233-254, 87-108)."""
    from bts_b200 import functional as Fn
    gen = torch.Generator().manual_seed(21)
    gt = torch.rand(B, H, W, generator=gen) * (hi * 1.2)            # some gt beyond max_depth, some ~0
    gt[torch.rand(B, H, W, generator=gen) < 0.3] = 0.0              # sparse ground truth
    pred = gt * (1 + 0.3 * torch.randn(B, H, W, generator=gen)) + 0.05 * torch.rand(B, H, W, generator=gen)
    pred[0, 3, 4], pred[0, 5, 6], pred[0, 7, 8], pred[0, 9, 10] = float("inf"), float("nan"), -1.0, 1e9
    got, cnt = Fn.eval_metrics(pred.cuda(), gt.cuda(), lo, hi, crop)
    for b in range(B):
        p, g = pred[b].numpy().copy(), gt[b].numpy()
        p[p < lo] = lo; p[p > hi] = hi; p[np.isinf(p)] = hi; p[np.isnan(p)] = lo     # bts_main.py:233-236
        valid = np.logical_and(g > lo, g < hi)
        if crop is not None:
            m = np.zeros_like(valid)
            m[crop[0]:crop[1], crop[2]:crop[3]] = 1
            valid = np.logical_and(valid, m)
        ref = _compute_errors_ref(g[valid], p[valid])
        assert int(cnt[b]) == int(valid.sum())
        assert np.allclose(got[b].cpu().numpy(), np.array(ref), rtol=1e-5, atol=1e-7), (got[b].cpu().numpy(), ref)


@pytest.mark.parametrize("B,H,W", [(1, 480, 640), (2, 37, 53), (1, 3, 5)], ids=str)
def test_backproject_vs_viewer_math(B, H, W):
    """posFromDepth / worldCoords / edges of pytorch/bts_live_3d.py:132-136, 414-429 -- the oracle is
    the viewer's own numpy + scipy.ndimage.sobel sequence (restated below line by line)."""
    from scipy import ndimage
    from bts_b200 import functional as Fn
    gen = torch.Generator().manual_seed(4)
    depth = (torch.rand(B, H, W, generator=gen) * 0.2 + torch.linspace(1, 6, W)[None, None, :]).contiguous()
    depth[:, H // 2:, W // 2:] += 2.5                                      # a depth discontinuity -> hidden points
    got = Fn.backproject(depth.cuda()).cpu().numpy()
    for b in range(B):
        d = depth[b].numpy().astype(np.float32).copy()
        cx, cy, fx, fy = W / 2, H / 2, 518.8579, 518.8579                  # bts_live_3d.py:415-417
        xx, yy = np.tile(range(W), H), np.repeat(range(H), W)
        xx, yy = (xx - cx) / fx, (yy - cy) / fy
        e = np.abs(ndimage.sobel(d, 0)) + np.abs(ndimage.sobel(d, 1))      # :133-136
        d[e > 0.3] = 1e6                                                   # :425
        z = d.reshape(H * W)
        ref = np.dstack((xx * z, yy * z, z)).reshape((H * W, 3))
        hidden_ref, hidden_got = ref[:, 2] == 1e6, got[b][:, 2] == 1e6
        # the Sobel sums are evaluated in a different order: only a pixel within rounding of the threshold may flip
        clear = (np.abs(e - 0.3) > 1e-4).reshape(H * W)
        assert (hidden_ref == hidden_got)[clear].all()
        assert np.allclose(got[b][clear], ref[clear], rtol=1e-5, atol=1e-6)
    if H * W > 1000:
        assert (got[..., 2] == 1e6).any() and (got[..., 2] != 1e6).any()
