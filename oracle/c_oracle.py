"""ctypes/numpy front-end of the plain-C oracle (oracle/bts_oracle.c) -- TEST INFRASTRUCTURE ONLY.

Every function restates one piece of /root/reference/pytorch/bts.py (line numbers in
bts_oracle_impl.h) and exists in fp32 and fp64. Parity pin: tests/test_oracle_golden.py checks each
against tests/golden/ vectors produced by the unmodified reference (tests/golden/make_golden.py).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libbts_oracle.so")
_lib = None


def build(force=False):
    """Compile oracle/libbts_oracle.so with gcc (idempotent)."""
    src = [os.path.join(_HERE, f) for f in ("bts_oracle.c", "bts_oracle_impl.h")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _suf(dtype):
    dtype = np.dtype(dtype)
    if dtype == np.float32:
        return "f32", ctypes.c_float
    if dtype == np.float64:
        return "f64", ctypes.c_double
    raise TypeError(dtype)


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _wptrs(ws):
    arr = (ctypes.c_void_p * len(ws))(*[w.ctypes.data for w in ws])
    return arr


def _prep(x, weights, dtype):
    x = np.ascontiguousarray(x, dtype=dtype)
    ws = [np.ascontiguousarray(np.asarray(w).reshape(w.shape[0], w.shape[1]), dtype=dtype) for w in weights]
    dims = [x.shape[1]] + [w.shape[0] for w in ws]
    for l, w in enumerate(ws):
        assert w.shape[1] == dims[l], (l, w.shape, dims)
    return x, ws, (ctypes.c_int * len(dims))(*dims), dims


def reduc_fwd(x, weights, max_depth, is_final, dtype=np.float32):
    """reduction_1x1.forward (bts.py:124-136): x [B,C0,h,w] -> [B,4,h,w] or [B,1,h,w]."""
    suf, creal = _suf(dtype)
    x, ws, cdims, dims = _prep(x, weights, dtype)
    B, _, h, w = x.shape
    out = np.empty((B, 1 if is_final else 4, h, w), dtype=dtype)
    getattr(lib(), "oracle_reduc_fwd_" + suf)(_ptr(x), _wptrs(ws), cdims, len(ws), int(bool(is_final)),
                                              creal(max_depth), B, h * w, _ptr(out))
    return out


def reduc_bwd(x, weights, max_depth, is_final, grad_out, dtype=np.float32):
    """Backward of reduction_1x1.forward -> (grad_x, [grad_W...])."""
    suf, creal = _suf(dtype)
    x, ws, cdims, dims = _prep(x, weights, dtype)
    B, _, h, w = x.shape
    g = np.ascontiguousarray(grad_out, dtype=dtype)
    assert g.size == B * (1 if is_final else 4) * h * w
    gx = np.empty_like(x)
    gws = [np.zeros_like(wt) for wt in ws]
    getattr(lib(), "oracle_reduc_bwd_" + suf)(_ptr(x), _wptrs(ws), cdims, len(ws), int(bool(is_final)),
                                              creal(max_depth), B, h * w, _ptr(g), _ptr(gx), _wptrs(gws))
    return gx, gws


def glue_fwd(out4, dtype=np.float32):
    """bts.py:250-253: normalise the normal, keep n4."""
    suf, _ = _suf(dtype)
    o = np.ascontiguousarray(out4, dtype=dtype)
    B, _, h, w = o.shape
    pe = np.empty_like(o)
    getattr(lib(), "oracle_glue_fwd_" + suf)(_ptr(o), B, h * w, _ptr(pe))
    return pe


def glue_bwd(out4, grad_plane_eq, dtype=np.float32):
    suf, _ = _suf(dtype)
    o = np.ascontiguousarray(out4, dtype=dtype)
    g = np.ascontiguousarray(grad_plane_eq, dtype=dtype)
    B, _, h, w = o.shape
    go = np.empty_like(o)
    getattr(lib(), "oracle_glue_bwd_" + suf)(_ptr(o), _ptr(g), B, h * w, _ptr(go))
    return go


def lpg_fwd(plane_eq, upratio, dtype=np.float32):
    """local_planar_guidance.forward (bts.py:149-173) -> (depth [B,H,W], abs_min scalar)."""
    suf, creal = _suf(dtype)
    pe = np.ascontiguousarray(plane_eq, dtype=dtype)
    B, c, h, w = pe.shape
    assert c == 4
    r = int(upratio)
    depth = np.empty((B, h * r, w * r), dtype=dtype)
    amin = creal(0)
    getattr(lib(), "oracle_lpg_fwd_" + suf)(_ptr(pe), B, h, w, r, _ptr(depth), ctypes.byref(amin))
    return depth, np.dtype(dtype).type(amin.value)


def lpg_bwd(grad_depth, plane_eq, upratio, dtype=np.float32):
    suf, _ = _suf(dtype)
    pe = np.ascontiguousarray(plane_eq, dtype=dtype)
    g = np.ascontiguousarray(grad_depth, dtype=dtype)
    B, _, h, w = pe.shape
    r = int(upratio)
    assert g.shape == (B, h * r, w * r)
    gpe = np.empty_like(pe)
    getattr(lib(), "oracle_lpg_bwd_" + suf)(_ptr(g), _ptr(pe), B, h, w, r, _ptr(gpe))
    return gpe


def fused_fwd(x, weights, max_depth, upratio, dtype=np.float32):
    """reduc -> glue -> LPG (bts.py:249-254) -> dict(out, plane_eq, depth, abs_min)."""
    out = reduc_fwd(x, weights, max_depth, False, dtype)
    pe = glue_fwd(out, dtype)
    depth, amin = lpg_fwd(pe, upratio, dtype)
    return dict(out=out, plane_eq=pe, depth=depth, abs_min=amin)


def fused_bwd(x, weights, max_depth, upratio, grad_depth, dtype=np.float32):
    """Backward of fused_fwd -> (grad_x, [grad_W...])."""
    out = reduc_fwd(x, weights, max_depth, False, dtype)
    pe = glue_fwd(out, dtype)
    gpe = lpg_bwd(grad_depth, pe, upratio, dtype)
    gout = glue_bwd(out, gpe, dtype)
    return reduc_bwd(x, weights, max_depth, False, gout, dtype)
