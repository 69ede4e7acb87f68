"""GPU (-m gpu), collected LAST on purpose (file name): the no-allocation entry point bts_silog_fwd_static.

It has no counterpart in the reference (added for callers that cannot call
cudaMalloc); its scratch is a file-scope `static __device__` array inside libbts_b200.so. Called through the C ABI and
checked against the existing bts_silog_fwd (caller workspace) on the same inputs: two calls in a row with different
sizes -- the second one smaller, so stale contents of the shared array would show -- and the backward from the state the
static variant hands out. Tolerance: the suite's 1e-5 (the two entry points run the same kernels on the same grid, so
the results are in fact bit-identical, which is asserted too).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _inputs(shape, frac, seed):
    gen = torch.Generator().manual_seed(seed)
    est = (torch.rand(shape, generator=gen) * 9.5 + 0.5).cuda()
    gt = (torch.rand(shape, generator=gen) * 9.5 + 0.5).cuda()
    mask = (torch.rand(shape, generator=gen) < frac).cuda()
    return est, gt, mask


def test_silog_fwd_static_matches_workspace_variant_twice_in_a_row():
    from bts_b200 import _cabi
    lib = _cabi.lib()
    st = torch.cuda.current_stream().cuda_stream
    vf = 0.85
    nws = lib.bts_silog_workspace_bytes()
    # everything the calls need exists BEFORE them: the entry point itself allocates nothing
    cases = [_inputs((8, 1, 480, 640), 0.4, 1), _inputs((3, 1, 37, 41), 0.8, 2)]
    bufs = [dict(ws=torch.zeros((nws + 7) // 8, dtype=torch.float64, device="cuda"),
                 loss_ref=torch.zeros((), device="cuda"), loss=torch.full((), -1.0, device="cuda"),
                 state=torch.zeros(8, dtype=torch.float64, device="cuda"),
                 g_ref=torch.empty_like(c[0]), g=torch.empty_like(c[0])) for c in cases]
    one = torch.ones((), device="cuda")
    torch.cuda.synchronize()
    mem0 = torch.cuda.memory_allocated()
    for (est, gt, mask), b in zip(cases, bufs):
        n = est.numel()
        _cabi.check(lib.bts_silog_fwd(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), n, vf, b["loss_ref"].data_ptr(),
                                      b["ws"].data_ptr(), nws, st), "bts_silog_fwd")
        _cabi.check(lib.bts_silog_fwd_static(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), n, vf, b["loss"].data_ptr(),
                                             b["state"].data_ptr(), st), "bts_silog_fwd_static")
        # backward from either state must agree too (state_out is what bts_silog_bwd takes as `ws`)
        _cabi.check(lib.bts_silog_bwd(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), n, vf, one.data_ptr(),
                                      b["ws"].data_ptr(), b["g_ref"].data_ptr(), st), "bts_silog_bwd")
        _cabi.check(lib.bts_silog_bwd(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), n, vf, one.data_ptr(),
                                      b["state"].data_ptr(), b["g"].data_ptr(), st), "bts_silog_bwd")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == mem0
    for b in bufs:
        ref, got = float(b["loss_ref"]), float(b["loss"])
        assert ref == ref and ref > 0
        assert abs(got - ref) <= RTOL * abs(ref), (got, ref)
        assert got == ref                                         # same kernels, same grid, same summation order
        assert torch.equal(b["g"], b["g_ref"])
    # state_out may be NULL (forward only)
    est, gt, mask = cases[1]
    loss = torch.full((), -1.0, device="cuda")
    _cabi.check(lib.bts_silog_fwd_static(est.data_ptr(), gt.data_ptr(), mask.data_ptr(), est.numel(), vf, loss.data_ptr(),
                                         None, st), "bts_silog_fwd_static")
    torch.cuda.synchronize()
    assert float(loss) == float(bufs[1]["loss_ref"])


def test_silog_fwd_static_rejects_sizes_beyond_its_scratch():
    from bts_b200 import _cabi
    lib = _cabi.lib()
    t = torch.ones(16, device="cuda")
    m = torch.ones(16, dtype=torch.bool, device="cuda")
    loss = torch.zeros((), device="cuda")
    rc = lib.bts_silog_fwd_static(t.data_ptr(), t.data_ptr(), m.data_ptr(), (1 << 30) + 1, 0.85, loss.data_ptr(), None,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.bts_last_error().decode()
    assert "exceeds" in msg and "static" in msg, msg
