"""Torch-CPU port of the reference hot path -- TEST INFRASTRUCTURE ONLY (never imported by bts_b200/).

The reference (pytorch/bts.py) is Python on top of ATen, and /root/reference does not travel to
the GPU box, so this functional port is what bench.py times as the CPU baseline
(cpu_baseline.kind = "port", and `bench.py --impl reference`). It issues the same ATen operators
in the same order as the reference modules so that its cost on the host cores is the reference's:

  chain_forward ........ reduction_1x1.forward              bts.py:124-136 (layers built as in 98-122)
  glue ................. normalize + cat in bts.forward     bts.py:250-253
  lpg_forward .......... local_planar_guidance.forward      bts.py:149-173 (the two `.cuda()` calls at
                         :157,:160 become no-ops because the port runs where its input lives)

Parity pin: tests/test_oracle_golden.py compares it with tests/golden/ (outputs of the unmodified
reference) -- forward values bit-exact on CPU with one thread, gradients to 1e-6.
"""
import math

import torch
import torch.nn.functional as F


def xavier_weights(dims, is_final, generator=None, dtype=torch.float32):
    """Weights shaped like reduction_1x1's ([out,in,1,1] each), xavier_uniform as bts.py:34-38."""
    ws = []
    for cin, cout in zip(dims[:-1], dims[1:]):
        bound = math.sqrt(6.0 / (cin + cout))
        ws.append(((torch.rand(cout, cin, 1, 1, generator=generator, dtype=dtype) * 2 - 1) * bound))
    return ws


def chain_dims(num_in, num_out, is_final):
    """Channel list produced by reduction_1x1.__init__'s while loop (bts.py:105-122)."""
    dims = [num_in]
    while num_out >= 4:
        if num_out < 8:
            dims.append(1 if is_final else 3)
            break
        dims.append(num_out)
        num_out //= 2
    return dims


def chain_forward(x, weights, max_depth, is_final):
    net = x
    for i, w in enumerate(weights):
        net = F.conv2d(net, w)
        if i + 1 < len(weights):
            net = F.elu(net)
    if is_final:
        return torch.sigmoid(net)
    theta = torch.sigmoid(net[:, 0, :, :]) * math.pi / 3
    phi = torch.sigmoid(net[:, 1, :, :]) * math.pi * 2
    dist = torch.sigmoid(net[:, 2, :, :]) * max_depth
    n1 = torch.mul(torch.sin(theta), torch.cos(phi)).unsqueeze(1)
    n2 = torch.mul(torch.sin(theta), torch.sin(phi)).unsqueeze(1)
    n3 = torch.cos(theta).unsqueeze(1)
    n4 = dist.unsqueeze(1)
    return torch.cat([n1, n2, n3, n4], dim=1)


def glue(out4):
    normal = F.normalize(out4[:, :3, :, :], 2, 1)
    dist = out4[:, 3, :, :]
    return torch.cat([normal, dist.unsqueeze(1)], 1)


def lpg_forward(plane_eq, upratio):
    """Returns (depth [B,H,W], abs_min 0-dim)."""
    r = int(upratio)
    rf = float(upratio)
    dev = plane_eq.device
    ex = torch.repeat_interleave(plane_eq, r, 2)
    ex = torch.repeat_interleave(ex, r, 3)
    n1, n2, n3, n4 = ex[:, 0, :, :], ex[:, 1, :, :], ex[:, 2, :, :], ex[:, 3, :, :]
    # the reference builds u, v on the host at full size and uploads them every call (bts.py:157,160)
    u = torch.arange(r).reshape([1, 1, r]).float().repeat(plane_eq.size(0), plane_eq.size(2) * r, plane_eq.size(3)).to(dev)
    u = (u - (rf - 1) * 0.5) / rf
    v = torch.arange(r).reshape([1, r, 1]).float().repeat(plane_eq.size(0), plane_eq.size(2), plane_eq.size(3) * r).to(dev)
    v = (v - (rf - 1) * 0.5) / rf
    if plane_eq.dtype != torch.float32:
        u, v = u.to(plane_eq.dtype), v.to(plane_eq.dtype)
    divided = n1 * u + n2 * v + n3
    abs_min = torch.abs(divided).min()
    eps = 1e-3
    dummy_eps = torch.ones_like(divided) * eps
    divided = torch.where((divided > 0) & (divided < eps), dummy_eps, divided)
    divided = torch.where((divided < 0) & (divided > -eps), -dummy_eps, divided)
    return n4 / divided, abs_min


def stage_forward(x, weights, max_depth, upratio):
    """One decoder LPG stage: reducNxN -> glue -> lpgNxN (bts.py:249-254). Returns (depth, abs_min)."""
    return lpg_forward(glue(chain_forward(x, weights, max_depth, False)), upratio)


def silog_forward(depth_est, depth_gt, mask, variance_focus):
    """silog_loss.forward, pytorch/bts.py:46-48 (pinned by tests/golden/silog_*.npz)."""
    d = torch.log(depth_est[mask]) - torch.log(depth_gt[mask])
    return torch.sqrt((d ** 2).mean() - variance_focus * (d.mean() ** 2)) * 10.0


# ---- upconv (pytorch/bts.py:83-94) and the fold the CUDA kernel uses -------------------------------
def upconv_forward(x, weight):
    """The reference's operator sequence: nearest x2 -> conv3x3 (padding 1, no bias) -> ELU   (bts.py:90-94)."""
    import torch.nn.functional as F
    return F.elu(F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), weight, padding=1))


def upconv_fold_weights(weight):
    """[Cout,Cin,3,3] -> [2 a][2 b][2 dy][2 dx][Cout][Cin]: phase (a, b) of the output is a 2x2 convolution of the LOW-RES
    input with sums of the 3x3 taps (rows of (a, dy): a=0: {0} | {1,2}; a=1: {0,1} | {2}; same for columns)."""
    import torch
    rows = {(0, 0): (0,), (0, 1): (1, 2), (1, 0): (0, 1), (1, 1): (2,)}
    wf = torch.zeros((2, 2, 2, 2) + tuple(weight.shape[:2]), dtype=weight.dtype)
    for a in range(2):
        for b in range(2):
            for dy in range(2):
                for dx in range(2):
                    for ky in rows[(a, dy)]:
                        for kx in rows[(b, dx)]:
                            wf[a, b, dy, dx] += weight[:, :, ky, kx]
    return wf


def upconv_forward_folded(x, weight):
    """Restatement of the kernel's arithmetic: y[o][2i+a][2j+b] = sum_c sum_{dy,dx} Wf[a][b][dy][dx][o][c] x[c][i-1+a+dy][j-1+b+dx]."""
    import torch
    import torch.nn.functional as F
    B, cin, h, w = x.shape
    wf = upconv_fold_weights(weight)
    xp = F.pad(x, (1, 1, 1, 1))
    y = torch.zeros((B, weight.shape[0], 2 * h, 2 * w), dtype=x.dtype)
    for a in range(2):
        for b in range(2):
            acc = 0
            for dy in range(2):
                for dx in range(2):
                    win = xp[:, :, a + dy:a + dy + h, b + dx:b + dx + w]           # x[c][i-1+a+dy][j-1+b+dx]
                    acc = acc + torch.einsum("oc,bcij->boij", wf[a, b, dy, dx], win)
            y[:, :, a::2, b::2] = acc
    return F.elu(y)


def upconv_unfold_grads(gwf):
    """Adjoint of upconv_fold_weights: [2][2][2][2][Cout][Cin] -> [Cout,Cin,3,3]; tap (ky, kx) collects the folded entries whose
    tap sets contain it (what upconv_unfold_kernel computes from the folded weight gradient)."""
    import torch
    rows = {(0, 0): (0,), (0, 1): (1, 2), (1, 0): (0, 1), (1, 1): (2,)}
    gw = torch.zeros(tuple(gwf.shape[4:]) + (3, 3), dtype=gwf.dtype)
    for a in range(2):
        for b in range(2):
            for dy in range(2):
                for dx in range(2):
                    for ky in rows[(a, dy)]:
                        for kx in rows[(b, dx)]:
                            gw[:, :, ky, kx] += gwf[a, b, dy, dx]
    return gw
