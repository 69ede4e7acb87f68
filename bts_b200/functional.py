"""torch.autograd.Function wrappers over the C ABI (include/bts_b200.h).

PyTorch is plumbing here: it owns device memory, the current stream and the autograd graph. All
arithmetic happens in libbts_b200.so. Tensors must be CUDA fp32; anything else raises (no fallback).
"""
import torch

from . import _cabi

EPS = 1e-3  # pytorch/bts.py:168


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _require_cuda_f32(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise RuntimeError("%s is on %s: bts_b200 runs on CUDA (sm_100a) only and has no CPU fallback" % (name, t.device))
    if t.dtype != torch.float32:
        raise RuntimeError("%s has dtype %s: the fp32 path expects torch.float32" % (name, t.dtype))
    return t.contiguous()


def _batch_strided(t, B, H, W, name):
    """(tensor, pointer, batch stride in elements) of a [B,H,W] / [B,1,H,W] fp32 CUDA map whose rows are dense but
    whose images may be further apart than H*W -- a channel slice of a concat buffer (pytorch/bts.py:260, 274, 287)
    or the slice CatBackward hands back. Anything else is made contiguous (one copy)."""
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda or t.dtype != torch.float32:
        t = _require_cuda_f32(t, name)
    shape_ok = tuple(t.shape) in ((B, H, W), (B, 1, H, W))
    if shape_ok and t.stride(-1) == 1 and t.stride(-2) == W:
        bs = t.stride(0) if B > 1 else H * W
        if bs >= H * W and bs % 4 == 0 and t.data_ptr() % 16 == 0:
            return t, t.data_ptr(), (0 if bs == H * W else bs)
    t = t.contiguous()
    return t, t.data_ptr(), 0


def _weights2d(weights):
    ws = []
    for i, w in enumerate(weights):
        w = _require_cuda_f32(w, "weight[%d]" % i)
        if w.dim() == 4:
            if w.shape[2] != 1 or w.shape[3] != 1:
                raise RuntimeError("weight[%d] must be a 1x1 convolution kernel, got %s" % (i, tuple(w.shape)))
        elif w.dim() != 2:
            raise RuntimeError("weight[%d] must be [out,in,1,1] or [out,in]" % i)
        ws.append(w)
    return ws


def _dims(x, ws):
    dims = [x.shape[1]] + [w.shape[0] for w in ws]
    for l, w in enumerate(ws):
        if w.shape[1] != dims[l]:
            raise RuntimeError("weight[%d] expects %d input channels, chain provides %d" % (l, w.shape[1], dims[l]))
    return dims


def _zero_grads(ws):
    """One zero-filled buffer carved into per-layer views: a single fill launch per backward call."""
    sizes = [t.numel() for t in ws]
    offs = [0]
    for n in sizes:
        offs.append(offs[-1] + (n + 3) // 4 * 4)  # 16-byte aligned views
    flat = torch.zeros(offs[-1], dtype=torch.float32, device=ws[0].device)
    return [flat[o:o + n].view(t.shape) for o, n, t in zip(offs, sizes, ws)]


def _workspace(dims, B, h, w, r, device):
    with torch.cuda.device(device):                                # the library sizes per-CTA partials by THIS device's SM count
        n = _cabi.lib().bts_workspace_bytes(_cabi.int_array(dims), len(dims) - 1, B, h, w, r)
    if n == 0:
        return None, 0, 0
    ws = torch.empty((n + 3) // 4, dtype=torch.float32, device=device)
    return ws, ws.data_ptr(), n


_STASH = {"on": True}


def set_stash(on):
    """Keep (default) or drop the activation stash of the wide chains; without it their backward recomputes
    the chain from x (less memory, one more kernel). Returns the previous setting."""
    prev = _STASH["on"]
    _STASH["on"] = bool(on)
    return prev


def _stash(dims, B, h, w, device, wanted):
    """Activation stash for chains whose backward would otherwise recompute them in a separate kernel
    (include/bts_b200.h, ABI 2). (tensor | None, pointer, bytes); 0 bytes when nothing is stashed."""
    with torch.cuda.device(device):
        n = _cabi.lib().bts_stash_bytes(_cabi.int_array(dims), len(dims) - 1, B, h, w) if (wanted and _STASH["on"]) else 0
    if n == 0:
        return None, None, 0
    st = torch.empty((n + 3) // 4, dtype=torch.float32, device=device)
    return st, st.data_ptr(), n


class _LPG(torch.autograd.Function):
    """local_planar_guidance.forward (pytorch/bts.py:149-173) and its autograd backward."""

    @staticmethod
    def forward(ctx, plane_eq, upratio):
        plane_eq = _require_cuda_f32(plane_eq, "plane_eq")
        if plane_eq.dim() != 4 or plane_eq.shape[1] != 4:
            raise RuntimeError("plane_eq must be [B,4,h,w], got %s" % (tuple(plane_eq.shape),))
        B, _, h, w = plane_eq.shape
        r = int(upratio)
        depth = torch.empty((B, h * r, w * r), dtype=torch.float32, device=plane_eq.device)
        abs_min = torch.full((), float("inf"), dtype=torch.float32, device=plane_eq.device)
        with torch.cuda.device(plane_eq.device):
            _cabi.check(_cabi.lib().bts_lpg_fwd(plane_eq.data_ptr(), depth.data_ptr(), abs_min.data_ptr(),
                                                B, h, w, r, EPS, _stream()), "bts_lpg_fwd")
        ctx.save_for_backward(plane_eq)
        ctx.r = r
        ctx.mark_non_differentiable(abs_min)
        ctx.set_materialize_grads(False)                           # no zero tensor (a fill kernel) for abs_min's gradient
        return depth, abs_min

    @staticmethod
    def backward(ctx, grad_depth, _grad_abs_min):
        (plane_eq,) = ctx.saved_tensors
        grad_depth = _require_cuda_f32(grad_depth, "grad_depth")
        B, _, h, w = plane_eq.shape
        gpe = torch.empty_like(plane_eq)
        with torch.cuda.device(plane_eq.device):
            _cabi.check(_cabi.lib().bts_lpg_bwd(grad_depth.data_ptr(), plane_eq.data_ptr(), gpe.data_ptr(),
                                                B, h, w, ctx.r, EPS, _stream()), "bts_lpg_bwd")
        return gpe, None


class _Reduc(torch.autograd.Function):
    """reduction_1x1.forward (pytorch/bts.py:124-136) and its autograd backward (chain recomputed)."""

    @staticmethod
    def forward(ctx, x, max_depth, is_final, *weights):
        x = _require_cuda_f32(x, "net")
        if x.dim() != 4:
            raise RuntimeError("net must be [B,C,h,w], got %s" % (tuple(x.shape),))
        ws = _weights2d(weights)
        dims = _dims(x, ws)
        B, _, h, w = x.shape
        out = torch.empty((B, 1 if is_final else 4, h, w), dtype=torch.float32, device=x.device)
        wsb, wsp, wsn = _workspace(dims, B, h, w, 0, x.device)
        st, stp, stn = _stash(dims, B, h, w, x.device, any(ctx.needs_input_grad))
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().bts_reduc_fwd_stash(
                x.data_ptr(), _cabi.ptr_array([t.data_ptr() for t in ws]), _cabi.int_array(dims), len(ws),
                int(bool(is_final)), float(max_depth), B, h, w, out.data_ptr(), wsp, wsn, stp, stn, _stream()),
                "bts_reduc_fwd")
        ctx.stash = st
        ctx.save_for_backward(x, *ws)
        ctx.cfg = (float(max_depth), bool(is_final), dims, [tuple(t.shape) for t in weights])
        return out

    @staticmethod
    def backward(ctx, grad_out):
        x, *ws = ctx.saved_tensors
        max_depth, is_final, dims, shapes = ctx.cfg
        grad_out = _require_cuda_f32(grad_out, "grad_out")
        B, _, h, w = x.shape
        need_x = ctx.needs_input_grad[0]
        need_w = any(ctx.needs_input_grad[3:])
        gx = torch.empty_like(x) if need_x else None
        gws = _zero_grads(ws) if need_w else None
        wsb, wsp, wsn = _workspace(dims, B, h, w, 0, x.device)
        st = ctx.stash
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().bts_reduc_bwd_stash(
                x.data_ptr(), _cabi.ptr_array([t.data_ptr() for t in ws]), _cabi.int_array(dims), len(ws),
                int(is_final), max_depth, B, h, w, grad_out.data_ptr(), gx.data_ptr() if need_x else None,
                _cabi.ptr_array([t.data_ptr() for t in gws]) if need_w else None, wsp, wsn,
                st.data_ptr() if st is not None else None, st.numel() * 4 if st is not None else 0, _stream()),
                "bts_reduc_bwd")
        gw_out = [g.view(s) for g, s in zip(gws, shapes)] if need_w else [None] * len(ws)
        return (gx, None, None, *gw_out)


class _ReducLPG(torch.autograd.Function):
    """Fused decoder stage reducNxN -> normalise/cat -> lpgNxN (pytorch/bts.py:249-254), optionally with the
    post-LPG glue `unsqueeze(1) / max_depth` and the nearest down-sampling by upratio/2 (bts.py:255-256, 269-270)."""

    @staticmethod
    def forward(ctx, out, x, max_depth, upratio, want_plane_eq, scaled, downsample, abs_min_buf, *weights):
        # `out` comes FIRST: autograd's CopySlices treats input 0 as the view that is modified in place
        x = _require_cuda_f32(x, "net")
        if x.dim() != 4:
            raise RuntimeError("net must be [B,C,h,w], got %s" % (tuple(x.shape),))
        ws = _weights2d(weights)
        dims = _dims(x, ws)
        B, _, h, w = x.shape
        r = int(upratio)
        if downsample and (not scaled or r not in (4, 8)):
            raise RuntimeError("the down-sampled map is the scaled map's (bts.py:256, 270): scaled=True and upratio 8 or 4")
        depth_bs = 0
        if out is not None:
            # destination owned by the caller: a [B,1,H,W] channel slice of a concat buffer (bts.py:260, 274, 287)
            if not scaled or downsample or want_plane_eq:
                raise RuntimeError("out= receives the scaled map depth / max_depth (bts.py:255) and nothing else: "
                                   "pass scaled=True, downsample=False, return_plane_eq=False")
            if tuple(out.shape) != (B, 1, h * r, w * r) or not out.is_cuda or out.dtype != torch.float32:
                raise RuntimeError("out must be a CUDA fp32 [B,1,H,W] = %s tensor or view, got %s %s on %s"
                                   % ((B, 1, h * r, w * r), tuple(out.shape), out.dtype, out.device))
            depth, _, depth_bs = _batch_strided(out, B, h * r, w * r, "out")
            if depth.data_ptr() != out.data_ptr():
                raise RuntimeError("out must have dense rows, a batch stride that is a multiple of 4 and 16-byte alignment")
            ctx.mark_dirty(out)
            depth = out
        else:
            depth = torch.empty((B, 1, h * r, w * r) if scaled else (B, h * r, w * r), dtype=torch.float32, device=x.device)
        ds = torch.empty((B, 1, 2 * h, 2 * w), dtype=torch.float32, device=x.device) if downsample else None
        # abs_min_buf: pre-set to +inf by the caller (the out= route cannot return a second tensor, see below)
        abs_min = abs_min_buf if abs_min_buf is not None else torch.full((), float("inf"), dtype=torch.float32, device=x.device)
        plane_eq = torch.empty((B, 4, h, w), dtype=torch.float32, device=x.device) if want_plane_eq else None
        wsb, wsp, wsn = _workspace(dims, B, h, w, r, x.device)
        st, stp, stn = _stash(dims, B, h, w, x.device, any(ctx.needs_input_grad))
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().bts_reduc_lpg_fwd_glue(
                x.data_ptr(), _cabi.ptr_array([t.data_ptr() for t in ws]), _cabi.int_array(dims), len(ws),
                float(max_depth), B, h, w, r, EPS, float(max_depth) if scaled else 1.0,
                plane_eq.data_ptr() if want_plane_eq else None, depth.data_ptr(), depth_bs,
                ds.data_ptr() if downsample else None, 0, abs_min.data_ptr(), wsp, wsn, stp, stn, _stream()),
                "bts_reduc_lpg_fwd")
        ctx.stash = st
        ctx.save_for_backward(x, *ws)
        ctx.cfg = (float(max_depth), r, dims, [tuple(t.shape) for t in weights], bool(scaled), bool(downsample))
        ctx.set_materialize_grads(False)                           # no zero tensors (fill kernels) for unused outputs
        if out is not None:
            # autograd allows ONE returned tensor when an input view is modified in place: abs_min went to abs_min_buf
            return out
        ctx.mark_non_differentiable(abs_min)
        outs = [depth, abs_min]
        if downsample:
            outs.append(ds)
        if want_plane_eq:
            ctx.mark_non_differentiable(plane_eq)
            outs.append(plane_eq)
        return tuple(outs)

    @staticmethod
    def backward(ctx, grad_depth, _g_abs_min=None, *rest):
        x, *ws = ctx.saved_tensors
        max_depth, r, dims, shapes, scaled, downsample = ctx.cfg
        B, _, h, w = x.shape
        if grad_depth is None:                                     # only the down-sampled map was used downstream
            grad_depth = torch.zeros((B, h * r, w * r), dtype=torch.float32, device=x.device)
        # the gradient of a map that went into torch.cat (or into a caller-owned slice) comes back as a batch-strided
        # slice: read it where it lies instead of copying it
        grad_depth, gd_ptr, gd_bs = _batch_strided(grad_depth, B, h * r, w * r, "grad_depth")
        grad_ds = rest[0] if downsample and rest and rest[0] is not None else None
        if grad_ds is not None:
            grad_ds = _require_cuda_f32(grad_ds, "grad_depth_ds")
        need_x = ctx.needs_input_grad[1]
        need_w = any(ctx.needs_input_grad[8:])
        gx = torch.empty_like(x) if need_x else None
        gws = _zero_grads(ws) if need_w else None
        wsb, wsp, wsn = _workspace(dims, B, h, w, r, x.device)
        st = ctx.stash
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().bts_reduc_lpg_bwd_glue(
                gd_ptr, gd_bs, grad_ds.data_ptr() if grad_ds is not None else None, 0,
                max_depth if scaled else 1.0, x.data_ptr(), _cabi.ptr_array([t.data_ptr() for t in ws]),
                _cabi.int_array(dims), len(ws), max_depth, B, h, w, r, EPS, gx.data_ptr() if need_x else None,
                _cabi.ptr_array([t.data_ptr() for t in gws]) if need_w else None, wsp, wsn,
                st.data_ptr() if st is not None else None, st.numel() * 4 if st is not None else 0, _stream()),
                "bts_reduc_lpg_bwd")
        gw_out = [g.view(s) for g, s in zip(gws, shapes)] if need_w else [None] * len(ws)
        return (None, gx, None, None, None, None, None, None, *gw_out)


# ------------------------------------------------------------------------------------------------
# bf16 twin (include/bts_b200.h, ABI 4): big tensors in bf16, weights fp32 (bf16 weights are upcast: 41 552 values),
# arithmetic as in the fp32 path. Where the library has no native bf16 kernel (the backward of the two 128-channel
# chains, odd shapes) the wrappers below run the fp32 CUDA path on upcast tensors -- still this library, never PyTorch
# or CPU arithmetic.
# ------------------------------------------------------------------------------------------------
def _require_cuda_bf16(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.bfloat16:
        raise RuntimeError("%s must be a CUDA torch.bfloat16 tensor in the bf16 path" % name)
    return t.contiguous()


def _weights_f32(weights):
    return _weights2d([w if w.dtype == torch.float32 else w.float() for w in weights])


def _workspace_bf16(dims, B, h, w, r, device):
    with torch.cuda.device(device):
        n = max(_cabi.lib().bts_workspace_bytes_bf16(_cabi.int_array(dims), len(dims) - 1, B, h, w, r), 256)
    ws = torch.empty((n + 3) // 4, dtype=torch.float32, device=device)
    return ws, ws.data_ptr(), n


def bf16_native(dims, is_final, backward, h, w, x=None, r=0):
    """Does libbts_b200 serve this chain natively in bf16 (forward, or forward + backward)? Backward: the one-pass kernel
    of the two narrow chains (upratio <= 2, h*w % 8 == 0, aligned x) or the stash route of the wider ones (h*w % 4 == 0)."""
    mode = _cabi.lib().bts_bf16_supported(_cabi.int_array(list(dims)), len(dims) - 1, int(bool(is_final)), int(bool(backward)))
    if not backward:
        return mode == 1
    aligned = x is None or x.data_ptr() % 16 == 0
    if mode == 1:
        return (h * w) % 8 == 0 and aligned and r <= 2
    return mode == 2 and (h * w) % 4 == 0 and aligned


def _stash_bf16(dims, B, h, w, device, wanted):
    n = 0
    if wanted and _STASH["on"]:
        with torch.cuda.device(device):
            n = _cabi.lib().bts_stash_bytes_bf16(_cabi.int_array(dims), len(dims) - 1, B, h, w)
    if n == 0:
        return None, None, 0
    st = torch.empty((n + 3) // 4, dtype=torch.float32, device=device)
    return st, st.data_ptr(), n


class _LPGBF(torch.autograd.Function):
    @staticmethod
    def forward(ctx, plane_eq, upratio):
        plane_eq = _require_cuda_bf16(plane_eq, "plane_eq")
        B, _, h, w = plane_eq.shape
        r = int(upratio)
        depth = torch.empty((B, h * r, w * r), dtype=torch.bfloat16, device=plane_eq.device)
        abs_min = torch.full((), float("inf"), dtype=torch.float32, device=plane_eq.device)
        with torch.cuda.device(plane_eq.device):
            _cabi.check(_cabi.lib().bts_lpg_fwd_bf16(plane_eq.data_ptr(), depth.data_ptr(), abs_min.data_ptr(), B, h, w, r, EPS,
                                                     _stream()), "bts_lpg_fwd_bf16")
        ctx.save_for_backward(plane_eq)
        ctx.r = r
        ctx.mark_non_differentiable(abs_min)
        ctx.set_materialize_grads(False)
        return depth, abs_min

    @staticmethod
    def backward(ctx, grad_depth, _g=None):
        (plane_eq,) = ctx.saved_tensors
        grad_depth = _require_cuda_bf16(grad_depth.to(torch.bfloat16), "grad_depth")
        B, _, h, w = plane_eq.shape
        gpe = torch.empty_like(plane_eq)
        with torch.cuda.device(plane_eq.device):
            _cabi.check(_cabi.lib().bts_lpg_bwd_bf16(grad_depth.data_ptr(), plane_eq.data_ptr(), gpe.data_ptr(), B, h, w, ctx.r,
                                                     EPS, _stream()), "bts_lpg_bwd_bf16")
        return gpe, None


class _ReducBF(torch.autograd.Function):
    """reduction_1x1 (+ optionally the fused LPG stage when upratio > 0) in bf16 storage."""

    @staticmethod
    def forward(ctx, x, max_depth, is_final, upratio, scaled, want_plane_eq, *weights):
        x = _require_cuda_bf16(x, "net")
        ws = _weights_f32(weights)
        dims = _dims(x, ws)
        B, _, h, w = x.shape
        r = int(upratio)
        wsb, wsp, wsn = _workspace_bf16(dims, B, h, w, r, x.device)
        st, stp, stn = _stash_bf16(dims, B, h, w, x.device, any(ctx.needs_input_grad))
        W = _cabi.ptr_array([t.data_ptr() for t in ws])
        outs = []
        with torch.cuda.device(x.device):
            if r == 0:
                out = torch.empty((B, 1 if is_final else 4, h, w), dtype=torch.bfloat16, device=x.device)
                _cabi.check(_cabi.lib().bts_reduc_fwd_bf16(x.data_ptr(), W, _cabi.int_array(dims), len(ws), int(bool(is_final)),
                                                           float(max_depth), B, h, w, out.data_ptr(), wsp, wsn, stp, stn, _stream()),
                            "bts_reduc_fwd_bf16")
                outs = [out]
            else:
                depth = torch.empty((B, 1, h * r, w * r) if scaled else (B, h * r, w * r), dtype=torch.bfloat16, device=x.device)
                abs_min = torch.full((), float("inf"), dtype=torch.float32, device=x.device)
                pe = torch.empty((B, 4, h, w), dtype=torch.bfloat16, device=x.device) if want_plane_eq else None
                _cabi.check(_cabi.lib().bts_reduc_lpg_fwd_bf16(
                    x.data_ptr(), W, _cabi.int_array(dims), len(ws), float(max_depth), B, h, w, r, EPS,
                    float(max_depth) if scaled else 1.0, pe.data_ptr() if want_plane_eq else None, depth.data_ptr(),
                    abs_min.data_ptr(), wsp, wsn, stp, stn, _stream()), "bts_reduc_lpg_fwd_bf16")
                ctx.mark_non_differentiable(abs_min)
                outs = [depth, abs_min]
                if want_plane_eq:
                    ctx.mark_non_differentiable(pe)
                    outs.append(pe)
        ctx.stash = st
        ctx.save_for_backward(x, *ws)
        ctx.cfg = (float(max_depth), bool(is_final), r, bool(scaled), dims, [(tuple(t.shape), t.dtype) for t in weights])
        ctx.set_materialize_grads(False)
        return outs[0] if len(outs) == 1 else tuple(outs)

    @staticmethod
    def backward(ctx, grad_out, *_rest):
        x, *ws = ctx.saved_tensors
        max_depth, is_final, r, scaled, dims, wmeta = ctx.cfg
        B, _, h, w = x.shape
        if grad_out is None:                                       # the output was not used downstream
            shape = (B, 1 if is_final else 4, h, w) if r == 0 else ((B, 1, h * r, w * r) if scaled else (B, h * r, w * r))
            grad_out = torch.zeros(shape, dtype=torch.bfloat16, device=x.device)
        grad_out = _require_cuda_bf16(grad_out.to(torch.bfloat16), "grad_out")
        need_x = ctx.needs_input_grad[0]
        need_w = any(ctx.needs_input_grad[6:])
        gx = torch.empty_like(x) if need_x else None
        gws = _zero_grads(ws) if need_w else None
        wsb, wsp, wsn = _workspace_bf16(dims, B, h, w, r, x.device)
        W = _cabi.ptr_array([t.data_ptr() for t in ws])
        GW = _cabi.ptr_array([t.data_ptr() for t in gws]) if need_w else None
        st = ctx.stash
        stp, stn = (st.data_ptr(), st.numel() * 4) if st is not None else (None, 0)
        with torch.cuda.device(x.device):
            if r == 0:
                _cabi.check(_cabi.lib().bts_reduc_bwd_bf16(x.data_ptr(), W, _cabi.int_array(dims), len(ws), int(is_final), max_depth,
                                                           B, h, w, grad_out.data_ptr(), gx.data_ptr() if need_x else None, GW,
                                                           wsp, wsn, stp, stn, _stream()), "bts_reduc_bwd_bf16")
            else:
                _cabi.check(_cabi.lib().bts_reduc_lpg_bwd_bf16(grad_out.data_ptr(), max_depth if scaled else 1.0, x.data_ptr(), W,
                                                               _cabi.int_array(dims), len(ws), max_depth, B, h, w, r, EPS,
                                                               gx.data_ptr() if need_x else None, GW, wsp, wsn, stp, stn, _stream()),
                            "bts_reduc_lpg_bwd_bf16")
        gw_out = [g.view(sh).to(dt) for g, (sh, dt) in zip(gws, wmeta)] if need_w else [None] * len(ws)
        return (gx, None, None, None, None, None, *gw_out)


class _Silog(torch.autograd.Function):
    """silog_loss.forward (pytorch/bts.py:46-48) and its autograd backward, without the mask gather."""

    @staticmethod
    def forward(ctx, depth_est, depth_gt, mask, variance_focus):
        est = _require_cuda_f32(depth_est, "depth_est")
        gt = _require_cuda_f32(depth_gt, "depth_gt")
        if mask.dtype != torch.bool or not mask.is_cuda:
            raise RuntimeError("mask must be a CUDA torch.bool tensor")
        if est.shape != gt.shape or est.shape != mask.shape:
            raise RuntimeError("depth_est, depth_gt and mask must have the same shape")
        mask = mask.contiguous()
        loss = torch.empty((), dtype=torch.float32, device=est.device)
        nws = _cabi.lib().bts_silog_workspace_bytes()
        ws = torch.empty((nws + 7) // 8, dtype=torch.float64, device=est.device)
        with torch.cuda.device(est.device):
            _cabi.check(_cabi.lib().bts_silog_fwd(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), est.numel(),
                                                  float(variance_focus), loss.data_ptr(), ws.data_ptr(), nws, _stream()),
                        "bts_silog_fwd")
        ctx.save_for_backward(est, gt, mask, ws)
        ctx.vf = float(variance_focus)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        est, gt, mask, ws = ctx.saved_tensors
        grad_loss = _require_cuda_f32(grad_loss, "grad_loss")
        g = torch.empty_like(est)
        with torch.cuda.device(est.device):
            _cabi.check(_cabi.lib().bts_silog_bwd(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), est.numel(), ctx.vf,
                                                  grad_loss.data_ptr(), ws.data_ptr(), g.data_ptr(), _stream()),
                        "bts_silog_bwd")
        return g, None, None, None


def silog(depth_est, depth_gt, mask, variance_focus):
    """-> 0-dim loss. Drop-in math of silog_loss.forward (gradient flows to depth_est only, as in training)."""
    return _Silog.apply(depth_est, depth_gt, mask, variance_focus)


EVAL_METRICS = ("silog", "abs_rel", "log10", "rms", "sq_rel", "log_rms", "d1", "d2", "d3")   # bts_main.py:84


def eval_metrics(pred, gt, min_depth, max_depth, crop=None):
    """compute_errors (pytorch/bts_main.py:87-108) with the clamp and validity mask of online_eval (:233-252), per
    image, on the GPU. pred, gt: [B,H,W] (or [B,1,H,W] / [H,W]); crop = (top, bottom, left, right) or None.
    -> (metrics [B,9] in EVAL_METRICS order, valid-pixel counts [B])."""
    pred = _require_cuda_f32(pred, "pred")
    gt = _require_cuda_f32(gt, "gt")
    if pred.shape != gt.shape:
        raise RuntimeError("pred and gt must have the same shape")
    if pred.dim() == 2:
        pred, gt = pred[None], gt[None]
    if pred.dim() == 4:
        if pred.shape[1] != 1:
            raise RuntimeError("expected a single-channel depth map")
        pred, gt = pred[:, 0], gt[:, 0]
    B, H, W = pred.shape
    top, bottom, left, right = (0, H, 0, W) if crop is None else [int(v) for v in crop]
    out = torch.empty((B, 9), dtype=torch.float32, device=pred.device)
    count = torch.empty((B,), dtype=torch.int32, device=pred.device)
    nws = _cabi.lib().bts_eval_workspace_bytes(B, H, W)
    ws = torch.empty(((nws + 7) // 8,), dtype=torch.float64, device=pred.device)
    with torch.cuda.device(pred.device):
        _cabi.check(_cabi.lib().bts_eval_metrics(pred.data_ptr(), gt.data_ptr(), B, H, W, float(min_depth), float(max_depth),
                                                 top, bottom, left, right, out.data_ptr(), count.data_ptr(), ws.data_ptr(),
                                                 nws, _stream()), "bts_eval_metrics")
    return out, count


def backproject(depth, fx=518.8579, fy=518.8579, cx=None, cy=None, edge_thresh=0.3, hidden=1e6):
    """posFromDepth (pytorch/bts_live_3d.py:422-429) on the GPU: depth [B,H,W] or [H,W] -> points [B,H*W,3] with depth
    edges (Sobel magnitude > edge_thresh) pushed to `hidden`. Defaults are the viewer's constants (:415-417, 425)."""
    depth = _require_cuda_f32(depth, "depth")
    single = depth.dim() == 2
    if single:
        depth = depth[None]
    if depth.dim() != 3:
        raise RuntimeError("depth must be [B,H,W] or [H,W]")
    B, H, W = depth.shape
    pts = torch.empty((B, H * W, 3), dtype=torch.float32, device=depth.device)
    with torch.cuda.device(depth.device):
        _cabi.check(_cabi.lib().bts_backproject(depth.data_ptr(), pts.data_ptr(), B, H, W, float(fx), float(fy),
                                                float(W / 2 if cx is None else cx), float(H / 2 if cy is None else cy),
                                                float(edge_thresh), float(hidden), _stream()), "bts_backproject")
    return pts[0] if single else pts


def _upconv_fwd_raw(x, weight):
    B, cin, h, w = x.shape
    cout = weight.shape[0]
    y = torch.empty((B, cout, 2 * h, 2 * w), dtype=torch.float32, device=x.device)
    nws = _cabi.lib().bts_upconv_workspace_bytes(B, cin, cout, h, w)
    ws = torch.empty(((nws + 15) // 16 * 4,), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _cabi.check(_cabi.lib().bts_upconv_fwd(x.data_ptr(), weight.data_ptr(), B, cin, cout, h, w, y.data_ptr(), ws.data_ptr(),
                                               ws.numel() * 4, _stream()), "bts_upconv_fwd")
    return y


class _Upconv(torch.autograd.Function):
    """upconv.forward (pytorch/bts.py:90-94) and its autograd backward (bts_upconv_fwd / bts_upconv_bwd)."""

    @staticmethod
    def forward(ctx, x, weight):
        y = _upconv_fwd_raw(x, weight)
        ctx.save_for_backward(x, weight, y)
        return y

    @staticmethod
    def backward(ctx, grad_y):
        x, weight, y = ctx.saved_tensors
        grad_y = _require_cuda_f32(grad_y, "grad_y").contiguous()
        B, cin, h, w = x.shape
        cout = weight.shape[0]
        gx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        gw = torch.empty_like(weight) if ctx.needs_input_grad[1] else None
        nws = _cabi.lib().bts_upconv_bwd_workspace_bytes(B, cin, cout, h, w)
        ws = torch.empty(((nws + 15) // 16 * 4,), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            _cabi.check(_cabi.lib().bts_upconv_bwd(x.data_ptr(), weight.data_ptr(), y.data_ptr(), grad_y.data_ptr(), B, cin, cout, h, w,
                                                   gx.data_ptr() if gx is not None else None, gw.data_ptr() if gw is not None else None,
                                                   ws.data_ptr(), ws.numel() * 4, _stream()), "bts_upconv_bwd")
        return gx, gw


def upconv_supported(cin, cout, need_grad):
    """Channel counts the upconv kernels serve (every upconv of the BTS decoders qualifies)."""
    return cin % (32 if need_grad else 16) == 0 and cout % 32 == 0


def upconv_forward(x, weight):
    """upconv.forward (pytorch/bts.py:90-94): ELU(conv3x3(nearest x2 (x)), padding 1, no bias) without materialising the
    up-sampled tensor. x [B,Cin,h,w], weight [Cout,Cin,3,3] -> [B,Cout,2h,2w]; differentiable in x and
    weight. Cin % 16 == 0 and Cout % 32 == 0 (Cin % 32 == 0 when a gradient is required); raises otherwise."""
    x = _require_cuda_f32(x, "x")
    weight = _require_cuda_f32(weight, "weight")
    if x.dim() != 4 or weight.dim() != 4 or tuple(weight.shape[2:]) != (3, 3) or weight.shape[1] != x.shape[1]:
        raise RuntimeError("upconv: x [B,Cin,h,w] and weight [Cout,Cin,3,3] expected")
    need_grad = torch.is_grad_enabled() and (x.requires_grad or weight.requires_grad)
    if not upconv_supported(x.shape[1], weight.shape[0], need_grad):
        raise RuntimeError("bts_b200 upconv: Cin %% %d == 0 and Cout %% 32 == 0 required (Cin=%d, Cout=%d)" % (
            32 if need_grad else 16, x.shape[1], weight.shape[0]))
    x, weight = x.contiguous(), weight.contiguous()
    if need_grad:
        return _Upconv.apply(x, weight)
    return _upconv_fwd_raw(x, weight)


def _any_grad(x, weights):
    return torch.is_grad_enabled() and (x.requires_grad or any(w.requires_grad for w in weights))


def lpg(plane_eq, upratio):
    """-> (depth [B,h*r,w*r], abs_min 0-dim). Drop-in math of local_planar_guidance.forward."""
    if isinstance(plane_eq, torch.Tensor) and plane_eq.dtype == torch.bfloat16:
        if int(upratio) in (1, 2, 4, 8):
            return _LPGBF.apply(plane_eq, upratio)
        depth, abs_min = _LPG.apply(plane_eq.float(), upratio)
        return depth.to(torch.bfloat16), abs_min
    return _LPG.apply(plane_eq, upratio)


def reduc(x, weights, max_depth, is_final=False):
    """-> [B,4,h,w] plane parameters, or [B,1,h,w] sigmoid head. Drop-in math of reduction_1x1.forward."""
    if isinstance(x, torch.Tensor) and x.dtype == torch.bfloat16:
        dims = [x.shape[1]] + [w.shape[0] for w in weights]
        if x.is_cuda and bf16_native(dims, is_final, _any_grad(x, weights), x.shape[2], x.shape[3], x):
            return _ReducBF.apply(x, max_depth, is_final, 0, False, False, *weights)
        return _Reduc.apply(x.float(), max_depth, is_final, *[w.float() for w in weights]).to(torch.bfloat16)
    return _Reduc.apply(x, max_depth, is_final, *weights)


def reduc_lpg(x, weights, max_depth, upratio, return_plane_eq=False, scaled=False, downsample=False, out=None):
    """Fused stage -> (depth, abs_min[, depth_ds][, plane_eq]). scaled: depth is [B,1,H,W] / max_depth
    (bts.py:255); downsample: also its nearest down-sampling by upratio/2, [B,1,2h,2w] (bts.py:256); out: a
    caller-owned [B,1,H,W] view (channel slice of a concat buffer) that receives the scaled map in place."""
    if isinstance(x, torch.Tensor) and x.dtype == torch.bfloat16:
        if downsample or out is not None:
            raise RuntimeError("the bf16 twin has no down-sampled map and no strided destination (fp32 path only)")
        dims = [x.shape[1]] + [w.shape[0] for w in weights]
        r = int(upratio)
        grad = _any_grad(x, weights)
        if x.is_cuda and r in (1, 2, 4, 8) and bf16_native(dims, False, grad, x.shape[2], x.shape[3], x, r):
            return _ReducBF.apply(x, max_depth, False, r, scaled, return_plane_eq, *weights)
        res = _ReducLPG.apply(None, x.float(), max_depth, upratio, return_plane_eq, scaled, False, None, *[w.float() for w in weights])
        return tuple(t.to(torch.bfloat16) if i != 1 else t for i, t in enumerate(res))
    if out is None:
        return _ReducLPG.apply(None, x, max_depth, upratio, return_plane_eq, scaled, downsample, None, *weights)
    abs_min = torch.full((), float("inf"), dtype=torch.float32, device=x.device)
    depth = _ReducLPG.apply(out, x, max_depth, upratio, return_plane_eq, scaled, downsample, abs_min, *weights)
    return depth, abs_min


PATH_AUTO, PATH_GENERIC, PATH_TCGEN05 = -1, 0, 2


def force_path(path):
    """Test / profiling hook: pin the chain implementation for calls from this thread (PATH_AUTO resets)."""
    _cabi.lib().bts_force_path(int(path))


def chain_path(dims, is_final):
    """Which implementation serves this chain (PATH_GENERIC or PATH_TCGEN05)."""
    return _cabi.lib().bts_reduc_path(_cabi.int_array(list(dims)), len(dims) - 1, int(bool(is_final)))
