"""ctypes binding of libbts_b200.so (include/bts_b200.h). No fallback: a missing library is an error."""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BTS_LIB") or os.path.join(_HERE, "lib", "libbts_b200.so")   # BTS_LIB: an A/B build of the same ABI
HEADER_PATH = os.path.join(_HERE, "..", "include", "bts_b200.h")

_c_float_p = ctypes.c_void_p  # raw device pointers travel as integers

_SIGNATURES = {
    "bts_abi_version": (ctypes.c_int, []),
    "bts_last_error": (ctypes.c_char_p, []),
    "bts_force_path": (None, [ctypes.c_int]),
    "bts_launch_count": (ctypes.c_ulonglong, []),
    "bts_reduc_path": (ctypes.c_int, [ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_int]),
    "bts_lpg_fwd": (ctypes.c_int, [_c_float_p, _c_float_p, _c_float_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                   ctypes.c_int, ctypes.c_float, ctypes.c_void_p]),
    "bts_lpg_bwd": (ctypes.c_int, [_c_float_p, _c_float_p, _c_float_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                   ctypes.c_int, ctypes.c_float, ctypes.c_void_p]),
    "bts_workspace_bytes": (ctypes.c_size_t, [ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "bts_reduc_fwd": (ctypes.c_int, [_c_float_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int),
                                     ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_int,
                                     ctypes.c_int, _c_float_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    "bts_reduc_bwd": (ctypes.c_int, [_c_float_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int),
                                     ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_int,
                                     ctypes.c_int, _c_float_p, _c_float_p, ctypes.POINTER(ctypes.c_void_p),
                                     ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    "bts_reduc_lpg_fwd": (ctypes.c_int, [_c_float_p, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int),
                                         ctypes.c_int, ctypes.c_float, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_float, _c_float_p, _c_float_p, _c_float_p,
                                         ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    "bts_reduc_lpg_bwd": (ctypes.c_int, [_c_float_p, _c_float_p, ctypes.POINTER(ctypes.c_void_p),
                                         ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_float, _c_float_p,
                                         ctypes.POINTER(ctypes.c_void_p), ctypes.c_void_p, ctypes.c_size_t,
                                         ctypes.c_void_p]),
}
_STASH_TAIL = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]   # stash, stash_bytes, stream
for _n in ("bts_reduc_fwd", "bts_reduc_bwd", "bts_reduc_lpg_fwd", "bts_reduc_lpg_bwd"):
    _res, _args = _SIGNATURES[_n]
    _SIGNATURES[_n + "_stash"] = (_res, _args[:-1] + _STASH_TAIL)
_SIGNATURES["bts_stash_bytes"] = (ctypes.c_size_t, [ctypes.POINTER(ctypes.c_int), ctypes.c_int, ctypes.c_int,
                                                    ctypes.c_int, ctypes.c_int])
_P, _I, _F, _LL, _SZ = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_longlong, ctypes.c_size_t
_SIGNATURES["bts_reduc_lpg_fwd_glue"] = (ctypes.c_int, [
    _P, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int), _I, _F, _I, _I, _I, _I, _F, _F,   # x .. eps, depth_div
    _P, _P, _LL, _P, _LL, _P,                                    # plane_eq_out, depth, depth_bstride, ds_out, ds_bstride, abs_min
    _P, _SZ, _P, _SZ, _P])                                       # ws, ws_bytes, stash, stash_bytes, stream
_SIGNATURES["bts_reduc_lpg_bwd_glue"] = (ctypes.c_int, [
    _P, _LL, _P, _LL, _F,                                        # grad_depth, grad_bstride, grad_ds, grad_ds_bstride, depth_div
    _P, ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int), _I, _F, _I, _I, _I, _I, _F,       # x .. eps
    _P, ctypes.POINTER(ctypes.c_void_p), _P, _SZ, _P, _SZ, _P])  # grad_x, grad_W, ws, ws_bytes, stash, stash_bytes, stream
_SIGNATURES["bts_silog_workspace_bytes"] = (ctypes.c_size_t, [])
_SIGNATURES["bts_silog_fwd"] = (ctypes.c_int, [_P, _P, _P, _LL, _F, _P, _P, _SZ, _P])
_SIGNATURES["bts_silog_bwd"] = (ctypes.c_int, [_P, _P, _P, _LL, _F, _P, _P, _P, _P])
_SIGNATURES["bts_bf16_supported"] = (ctypes.c_int, [_P, _I, _I, _I])
_SIGNATURES["bts_workspace_bytes_bf16"] = (ctypes.c_size_t, [_P, _I, _I, _I, _I, _I])
_SIGNATURES["bts_lpg_fwd_bf16"] = (ctypes.c_int, [_P, _P, _P, _I, _I, _I, _I, _F, _P])
_SIGNATURES["bts_lpg_bwd_bf16"] = (ctypes.c_int, [_P, _P, _P, _I, _I, _I, _I, _F, _P])
_SIGNATURES["bts_stash_bytes_bf16"] = (ctypes.c_size_t, [_P, _I, _I, _I, _I])
_SIGNATURES["bts_reduc_fwd_bf16"] = (ctypes.c_int, [_P, _P, _P, _I, _I, _F, _I, _I, _I, _P, _P, _SZ, _P, _SZ, _P])
_SIGNATURES["bts_reduc_bwd_bf16"] = (ctypes.c_int, [_P, _P, _P, _I, _I, _F, _I, _I, _I, _P, _P, _P, _P, _SZ, _P, _SZ, _P])
_SIGNATURES["bts_reduc_lpg_fwd_bf16"] = (ctypes.c_int, [_P, _P, _P, _I, _F, _I, _I, _I, _I, _F, _F, _P, _P, _P, _P, _SZ, _P, _SZ, _P])
_SIGNATURES["bts_reduc_lpg_bwd_bf16"] = (ctypes.c_int, [_P, _F, _P, _P, _P, _I, _F, _I, _I, _I, _I, _F, _P, _P, _P, _SZ, _P, _SZ, _P])
_SIGNATURES["bts_silog_fwd_static"] = (ctypes.c_int, [_P, _P, _P, _LL, _F, _P, _P, _P])
_SIGNATURES["bts_eval_workspace_bytes"] = (ctypes.c_size_t, [_I, _I, _I])
_SIGNATURES["bts_eval_metrics"] = (ctypes.c_int, [_P, _P, _I, _I, _I, _F, _F, _I, _I, _I, _I, _P, _P, _P, _SZ, _P])
_SIGNATURES["bts_backproject"] = (ctypes.c_int, [_P, _P, _I, _I, _I, _F, _F, _F, _F, _F, _F, _P])
_SIGNATURES["bts_upconv_workspace_bytes"] = (ctypes.c_size_t, [_I, _I, _I, _I, _I])
_SIGNATURES["bts_upconv_fwd"] = (ctypes.c_int, [_P, _P, _I, _I, _I, _I, _I, _P, _P, _SZ, _P])
_SIGNATURES["bts_upconv_bwd_workspace_bytes"] = (ctypes.c_size_t, [_I, _I, _I, _I, _I])
_SIGNATURES["bts_upconv_bwd"] = (ctypes.c_int, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _SZ, _P])
ABI_VERSION = 4

_lib = None


class BtsLibraryError(RuntimeError):
    pass


def declared_symbols(header=HEADER_PATH):
    """Names of every BTS_API function the public header declares."""
    with open(header) as f:
        return re.findall(r"BTS_API\s+[\w\s\*]+?\b(bts_\w+)\s*\(", f.read())


def lib():
    """The loaded C-ABI library. Raises loudly when it has not been built (no CPU / torch fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BtsLibraryError(
                "libbts_b200.so is missing (%s). Build it with `python -m bts_b200.build` "
                "(or __graft_entry__.build()). bts_b200 has no CPU or PyTorch fallback." % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        if l.bts_abi_version() != ABI_VERSION:
            raise BtsLibraryError("libbts_b200.so ABI version %d, expected %d" % (l.bts_abi_version(), ABI_VERSION))
        _lib = l
    return _lib


def check(code, what):
    if code != 0:
        msg = lib().bts_last_error().decode("utf-8", "replace")
        raise RuntimeError("%s failed (code %d): %s" % (what, code, msg))


def int_array(values):
    return (ctypes.c_int * len(values))(*values)


def ptr_array(ptrs):
    return (ctypes.c_void_p * len(ptrs))(*ptrs)
