"""CPU: the C-ABI library loads, exports every symbol include/bts_b200.h declares, validates its
arguments without touching a GPU, and the Python host mirrors the reference's module interface."""
import ctypes

import pytest
import torch

import bts_b200
from bts_b200 import _cabi
from oracle import torch_port as tp


def test_library_exports_every_declared_symbol():
    names = _cabi.declared_symbols()
    assert len(names) >= 11 and len(set(names)) == len(names)
    raw = ctypes.CDLL(_cabi.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "libbts_b200.so does not export " + n
    assert set(names) == set(_cabi._SIGNATURES), "ctypes table and header disagree"
    assert _cabi.lib().bts_abi_version() == _cabi.ABI_VERSION


def test_argument_validation_needs_no_gpu():
    lib = _cabi.lib()
    # r < 1
    assert lib.bts_lpg_fwd(None, None, None, 1, 2, 2, 0, 1e-3, None) == -1
    assert b"upratio" in lib.bts_last_error()
    # NULL pointers with non-empty shape
    assert lib.bts_lpg_fwd(None, None, None, 1, 2, 2, 8, 1e-3, None) == -1
    assert lib.bts_lpg_bwd(None, None, None, 1, 2, 2, 8, 1e-3, None) == -1
    # empty input is a successful no-op (reference: empty tensors flow through)
    assert lib.bts_lpg_fwd(None, None, None, 0, 2, 2, 8, 1e-3, None) == 0
    assert lib.bts_lpg_bwd(None, None, None, 3, 0, 2, 8, 1e-3, None) == 0
    # chain description errors
    dims = _cabi.int_array([32, 16, 8, 1])
    w = _cabi.ptr_array([16, 16, 16])          # fake non-NULL device pointers; never dereferenced
    assert lib.bts_reduc_fwd(None, w, dims, 3, 0, 10.0, 1, 2, 2, None, None, 0, None) == -1  # last != 3
    assert b"last layer" in lib.bts_last_error()
    assert lib.bts_reduc_fwd(None, w, dims, 9, 1, 10.0, 1, 2, 2, None, None, 0, None) == -1  # too many layers
    assert lib.bts_reduc_fwd(None, w, dims, 3, 1, 10.0, 1, 2, 2, None, None, 0, None) == -1  # x NULL
    assert lib.bts_reduc_fwd(None, w, dims, 3, 1, 10.0, 0, 2, 2, None, None, 0, None) == 0   # empty batch
    bad = _cabi.int_array([32, 0, 8, 1])
    assert lib.bts_reduc_fwd(None, w, bad, 3, 1, 10.0, 1, 2, 2, None, None, 0, None) == -1
    wnull = _cabi.ptr_array([16, None, 16])
    assert lib.bts_reduc_fwd(None, wnull, dims, 3, 1, 10.0, 1, 2, 2, None, None, 0, None) == -1
    # fused backward without workspace
    d3 = _cabi.int_array([32, 16, 8, 3])
    assert lib.bts_reduc_lpg_bwd(16, 16, w, d3, 3, 10.0, 1, 2, 2, 2, 1e-3, None, None, None, 0, None) in (-3, 0)
    assert lib.bts_workspace_bytes(d3, 3, 0, 4, 4, 2) == 0
    # upconv: channel counts the kernels do not serve, missing workspace, empty input
    assert lib.bts_upconv_fwd(16, 16, 1, 24, 32, 4, 4, 16, 16, 1 << 20, None) == -2
    assert b"cin" in lib.bts_last_error()
    assert lib.bts_upconv_fwd(16, 16, 1, 32, 32, 4, 4, 16, None, 0, None) == -1
    assert lib.bts_upconv_fwd(None, 16, 1, 32, 32, 4, 4, 16, 16, 1 << 20, None) == -1
    assert lib.bts_upconv_fwd(None, None, 0, 32, 32, 4, 4, None, None, 0, None) == 0
    assert lib.bts_upconv_workspace_bytes(2, 64, 32, 4, 6) == (16 * 64 * 32 + 2 * 64 * 4 * 6) * 8 and lib.bts_upconv_workspace_bytes(1, 0, 32, 4, 4) == 0
    assert lib.bts_upconv_bwd(16, 16, 16, 16, 1, 48, 32, 4, 4, 16, 16, 16, 1 << 20, None) == -2   # backward: cin % 32
    assert lib.bts_upconv_bwd(16, 16, 16, 16, 1, 32, 32, 4, 4, 16, 16, 16, 64, None) == -1       # workspace too small
    assert lib.bts_upconv_bwd_workspace_bytes(2, 64, 32, 4, 6) == 4 * (48 * 64 * 32 + 4 * 2 * 32 * 4 * 6)


@pytest.mark.parametrize("cfg", [(128, 128, False), (128, 64, False), (64, 32, False), (32, 16, True),
                                 (64, 64, False), (16, 8, True), (8, 4, True), (8, 4, False), (256, 256, False)])
def test_module_structure_matches_reference_chain(cfg):
    num_in, num_out, is_final = cfg
    m = bts_b200.reduction_1x1(num_in, num_out, 10.0, is_final)
    dims = tp.chain_dims(num_in, num_out, is_final)
    assert bts_b200.chain_channels(num_in, num_out, is_final) == dims
    ws = m.chain_weights()
    assert [tuple(w.shape) for w in ws] == [(o, i, 1, 1) for i, o in zip(dims[:-1], dims[1:])]
    keys = list(m.state_dict().keys())
    expect = ["reduc.inter_%d_%d.0.weight" % (i, o) for i, o in zip(dims[:-2], dims[1:-1])]
    expect.append("reduc.final.0.weight" if is_final else "reduc.plane_params.weight")
    assert keys == expect                      # checkpoints must keep loading
    assert m.max_depth == 10.0 and m.is_final == is_final and isinstance(m.sigmoid, torch.nn.Sigmoid)
    # weights_init_xavier (bts.py:34-38) matches on isinstance(m, nn.Conv2d): every weight must be reachable
    convs = [c for c in m.modules() if isinstance(c, torch.nn.Conv2d)]
    assert len(convs) == len(ws) and all(c.bias is None for c in convs)


def test_lpg_module_attributes():
    m = bts_b200.local_planar_guidance(8)
    assert m.upratio == 8.0 and isinstance(m.upratio, float) and m.abs_min is None
    assert tuple(m.u.shape) == (1, 1, 8) and tuple(m.v.shape) == (1, 8, 1)
    assert len(list(m.parameters())) == 0 and len(m.state_dict()) == 0


def test_cpu_tensors_fail_loudly():
    """No CPU / PyTorch fallback: the product refuses anything that is not a CUDA fp32 tensor."""
    red = bts_b200.reduction_1x1(32, 16, 10.0, True)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        red(torch.randn(1, 32, 4, 4))
    lpg = bts_b200.local_planar_guidance(2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lpg(torch.rand(1, 4, 2, 2), None)


def test_product_never_imports_oracle():
    import os
    import re
    root = os.path.dirname(os.path.abspath(bts_b200.__file__))
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|oracle/|libbts_oracle", src, re.M), f
