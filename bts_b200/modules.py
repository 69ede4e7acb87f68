"""Drop-in replacements for the two hot-path modules of minghanz/bts (pytorch/bts.py).

    reduction_1x1(num_in_filters, num_out_filters, max_depth, is_final=False).forward(net)
    local_planar_guidance(upratio).forward(plane_eq, focal)

Constructor and forward signatures, attribute names (`reduc`, `sigmoid`, `max_depth`, `is_final`,
`upratio`, `u`, `v`, `abs_min`) and state_dict keys (`reduc.inter_<in>_<out>.0.weight`,
`reduc.plane_params.weight`, `reduc.final.0.weight`) are the reference's (pytorch/bts.py:97-147,
SURVEY.md 3.5), so `weights_init_xavier` (bts.py:34-38), checkpoints (bts_main.py:368,725-728), the
optimizer grouping (bts_main.py:354-355) and DDP hooks keep working. The arithmetic runs in
libbts_b200.so (hand-written sm_100a CUDA); there is no PyTorch or CPU fallback for these two modules.

`fused_stage(reduc, lpg, net)` evaluates reducNxN -> normalise/cat -> lpgNxN (bts.py:249-254) in one
pass; `graft(decoder)` swaps the modules inside an existing reference `bts` decoder / BtsModel.
"""
import torch
import torch.nn as nn

from . import functional as Fn


def chain_channels(num_in_filters, num_out_filters, is_final):
    """Channel widths [in, ..., 3|1] of the 1x1 chain reduction_1x1 builds (pytorch/bts.py:105-122):
    halve the width down to 8, then project to 3 plane logits (or 1 depth logit when final)."""
    widths = [int(num_in_filters)]
    width = int(num_out_filters)
    while width >= 8:
        widths.append(width)
        width //= 2
    if width >= 4:
        widths.append(1 if is_final else 3)
    return widths


class reduction_1x1(nn.Sequential):
    """pytorch/bts.py:97-136. Holds real nn.Conv2d parameters; forward runs the fused CUDA chain."""

    def __init__(self, num_in_filters, num_out_filters, max_depth, is_final=False):
        super().__init__()
        self.max_depth = max_depth
        self.is_final = is_final
        self.sigmoid = nn.Sigmoid()
        self.reduc = nn.Sequential()
        widths = chain_channels(num_in_filters, num_out_filters, is_final)
        last = len(widths) - 2
        for idx, (cin, cout) in enumerate(zip(widths[:-1], widths[1:])):
            conv = nn.Conv2d(cin, cout, kernel_size=1, stride=1, padding=0, bias=False)
            if idx == last and cout < 4:   # the projection: widths above it never drop below 8
                if is_final:
                    self.reduc.add_module("final", nn.Sequential(conv, nn.Sigmoid()))
                else:
                    self.reduc.add_module("plane_params", conv)
            else:
                self.reduc.add_module("inter_{}_{}".format(cin, cout), nn.Sequential(conv, nn.ELU()))
        self._widths = widths

    def chain_weights(self):
        """Conv weights in chain order ([out,in,1,1] each)."""
        ws = []
        for m in self.reduc.children():
            conv = m if isinstance(m, nn.Conv2d) else m[0]
            ws.append(conv.weight)
        return ws

    def forward(self, net):
        return Fn.reduc(net, self.chain_weights(), self.max_depth, self.is_final)


class local_planar_guidance(nn.Module):
    """pytorch/bts.py:138-173. `focal` is accepted and ignored, exactly like the reference."""

    def __init__(self, upratio):
        super().__init__()
        self.upratio = upratio
        self.u = torch.arange(self.upratio).reshape([1, 1, self.upratio]).float()
        self.v = torch.arange(int(self.upratio)).reshape([1, self.upratio, 1]).float()
        self.upratio = float(upratio)
        self.abs_min = None   # min |den| before the clamp, refreshed by every forward (bts_main.py:484-486)

    def forward(self, plane_eq, focal=None):
        depth, self.abs_min = Fn.lpg(plane_eq, int(self.upratio))
        return depth


class silog_loss(nn.Module):
    """Same constructor / forward as the reference silog_loss (pytorch/bts.py:41-48)."""

    def __init__(self, variance_focus):
        super(silog_loss, self).__init__()
        self.variance_focus = variance_focus

    def forward(self, depth_est, depth_gt, mask):
        return Fn.silog(depth_est, depth_gt, mask, self.variance_focus)


class upconv(nn.Module):
    """Mirror of the reference `upconv` (pytorch/bts.py:83-94): same constructor, attributes (`elu`, `conv`, `ratio`) and
    state_dict key (`conv.weight`). Forward and backward run the folded tensor-core kernels (functional.upconv_forward).
    What the kernels do not serve -- ratio != 2, channel counts that are not multiples of 32 (16 for the input channels
    when no gradient is required), non-CUDA / non-fp32 tensors -- runs the reference's own operator sequence;
    `native(x)` tells which."""

    def __init__(self, in_channels, out_channels, ratio=2):
        super().__init__()
        self.elu = nn.ELU()
        self.conv = nn.Conv2d(in_channels=in_channels, out_channels=out_channels, bias=False, kernel_size=3, stride=1, padding=1)
        self.ratio = ratio

    def native(self, x):
        need_grad = torch.is_grad_enabled() and (x.requires_grad or self.conv.weight.requires_grad)
        return (self.ratio == 2 and x.is_cuda and x.dtype == torch.float32 and x.dim() == 4
                and Fn.upconv_supported(self.conv.in_channels, self.conv.out_channels, need_grad))

    def forward(self, x):
        if self.native(x):
            return Fn.upconv_forward(x, self.conv.weight)
        up_x = torch.nn.functional.interpolate(x, scale_factor=self.ratio, mode='nearest')
        return self.elu(self.conv(up_x))


def fused_stage(reduc, lpg, net, return_plane_eq=False, scaled=False, downsample=False, out=None):
    """reducNxN -> normalise/cat -> lpgNxN (pytorch/bts.py:249-254) as ONE fused call.

    Returns depth [B,H,W] (and plane_eq [B,4,h,w] when asked); sets lpg.abs_min like the module would.
    scaled=True folds `depth.unsqueeze(1) / max_depth` (bts.py:255, 269, 283) in and returns [B,1,H,W];
    downsample=True (upratio 8 / 4) also returns `F.interpolate(depth_scaled, scale_factor=2/upratio,
    mode='nearest')` (bts.py:256, 270), [B,1,2h,2w], right after the depth map.
    out= (with scaled=True): a caller-owned [B,1,H,W] view -- e.g. `concat1[:, 1:2]` of a pre-allocated buffer that
    replaces the `torch.cat` of bts.py:287 -- receives the map in place (no copy forward). Backward: on this out= route
    autograd's CopySlices hands the Function a contiguous clone of the gradient slice (one full-resolution copy made by
    PyTorch, not by us); the zero-copy read of a batch-strided gradient slice applies to the plain `torch.cat` route,
    where CatBackward returns a strided view."""
    if reduc.is_final:
        raise ValueError("fused_stage needs a plane-parameter reduction (is_final=False)")
    res = Fn.reduc_lpg(net, reduc.chain_weights(), reduc.max_depth, int(lpg.upratio), return_plane_eq, scaled, downsample, out)
    lpg.abs_min = res[1]
    out = [res[0]] + list(res[2:])
    return out[0] if len(out) == 1 else tuple(out)


_STAGES = (("reduc8x8", "lpg8x8"), ("reduc4x4", "lpg4x4"), ("reduc2x2", "lpg2x2"))


def graft(model, upconvs=False):
    """Replace every reference reduction_1x1 / local_planar_guidance inside `model` (a reference `bts`
    decoder, BtsModel, or anything containing them) by the B200 modules, keeping the weights.
    Matching is by class NAME so the reference module need not be importable here.
    upconvs=True also swaps the reference's `upconv` blocks (bts.py:83-94) for the folded kernels: fp32-accurate
    (1.5e-6 of the fp64 oracle) and 1.6x the reference's sequence on cuDNN in fp32, but slower than cuDNN with PyTorch's
    TF32 default (which is 3e-4 off) -- hence opt-in."""
    for name, child in list(model.named_children()):
        cls = type(child).__name__
        if upconvs and cls == "upconv" and not isinstance(child, upconv):
            new = upconv(child.conv.in_channels, child.conv.out_channels, child.ratio)
            new.load_state_dict(child.state_dict())
            new.to(child.conv.weight.device)
            setattr(model, name, new)
            continue
        if cls == "reduction_1x1" and not isinstance(child, reduction_1x1):
            convs = [m for m in child.modules() if isinstance(m, nn.Conv2d)]
            new = reduction_1x1(convs[0].in_channels, convs[0].out_channels if len(convs) > 1 else 4,
                                child.max_depth, child.is_final)
            new.load_state_dict(child.state_dict())
            new.to(convs[0].weight.device)
            setattr(model, name, new)
        elif cls == "local_planar_guidance" and not isinstance(child, local_planar_guidance):
            setattr(model, name, local_planar_guidance(int(child.upratio)))
        else:
            graft(child, upconvs)
    return model
