#!/usr/bin/env python
"""bench.py -- LPG decoder Mpixels/s, forward + backward, on N B200s (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload nyu_b8]

A step = one forward + backward pass of the four hot-path stages of the BTS decoder
(reduc8x8+lpg8x8, reduc4x4+lpg4x4, reduc2x2+lpg2x2, reduc1x1 head; pytorch/bts.py:249-286) over one
batch of synthetic decoder features. Unit of work = one full-resolution pixel of one image
(SURVEY.md 8(d)). Workload at N=1: BASELINE.json configs[1] (NYU 480x640, batch 8, max_depth 10);
each extra GPU gets its own batch of 8 (weak scaling, batch sharding, no data-path collective).

Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for every field.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# name, H, W, per-GPU batch, max_depth, focal (train_test_inputs/*.txt col 3)
WORKLOADS = {
    "nyu_b8": dict(H=480, W=640, B=8, max_depth=10.0, focal=518.8579,
                   desc="BTS DenseNet161 NYU Depth V2 480x640 batch 8 (BASELINE.json configs[1]), LPG3+head stages"),
    "kitti_b16": dict(H=352, W=1216, B=16, max_depth=80.0, focal=721.5377,
                      desc="BTS ResNeXt101 KITTI Eigen 352x1216 batch 16 (configs[2]), LPG3+head stages"),
    "kitti_train_b4": dict(H=352, W=704, B=4, max_depth=80.0, focal=721.5377,
                           desc="BTS DenseNet161 KITTI 352x704 batch 4/GPU (configs[3]), LPG3+head stages"),
    "nyu_train_b4": dict(H=416, W=544, B=4, max_depth=10.0, focal=518.8579,
                         desc="NYU 416x544 batch 4 (configs[0]), LPG3+head stages"),
}
# decoder instances at bts_size=512 (pytorch/bts.py:198-217): name, C_in, num_out, upratio, is_final
STAGES = [("reduc8x8+lpg8x8", 128, 128, 8, False), ("reduc4x4+lpg4x4", 128, 64, 4, False),
          ("reduc2x2+lpg2x2", 64, 32, 2, False), ("reduc1x1_head", 32, 16, 1, True)]
# algorithmic bytes per FULL-RES pixel, fp32
ALGO_BYTES = {"reduc8x8+lpg8x8": (12.25, 20.0), "reduc4x4+lpg4x4": (37.0, 68.0),
              "reduc2x2+lpg2x2": (72.0, 132.0), "reduc1x1_head": (132.0, 260.0)}
HBM_FALLBACK_GBS = 6650.0   # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="nyu_b8", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU port timing (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg (profiling runs)")
    ap.add_argument("--kernel-only", action="store_true", help="only the timed device loop (for ncu)")
    ap.add_argument("--e2e-serial", action="store_true", help="end-to-end leg on one stream (no upload/download overlap)")
    ap.add_argument("--no-graph", action="store_true", help="launch every step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--e2e-d2h", default="all", choices=["all", "results"],
                    help="what the end-to-end leg reads back to the host every step: every output incl. grad_x (default, the "
                         "conservative reading) or only what the reference brings to the host -- depth maps / head output "
                         "(bts_test.py:134-138), weight gradients, abs_min (bts_main.py:484-486); grad_x stays on the GPU, "
                         "where the encoder's backward consumes it")
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"],
                    help="f32 (the reference's arithmetic: the headline) or the bf16 twin (bf16 storage, same arithmetic)")
    ap.add_argument("--grad-allreduce", action="store_true",
                    help="N > 1: every step ends with an NCCL all-reduce(avg) of the path's weight gradients as one "
                         "bucket, as DDP does in training (BASELINE.json configs[3]; pytorch/bts_main.py:314)")
    ap.add_argument("--allreduce-buckets", default="two", choices=["per-stage", "two", "one"],
                    help="--grad-allreduce: one collective per stage as soon as its backward is enqueued; two (the three LPG "
                         "stages under the head's kernels, then the head's 648 floats); or one at the end of the step")
    return ap.parse_args()


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# CPU baseline: the torch port of the reference modules (oracle/torch_port.py) on the host cores
# ------------------------------------------------------------------------------------------------
def make_cpu_inputs(wl, B, seed=0):
    import torch
    from oracle import torch_port as tp
    gen = torch.Generator().manual_seed(seed)
    H, W = wl["H"], wl["W"]
    items = []
    for name, cin, nout, r, final in STAGES:
        dims = tp.chain_dims(cin, nout, final)
        ws = [w.requires_grad_(True) for w in tp.xavier_weights(dims, final, gen)]
        x = torch.randn(B, cin, H // r, W // r, generator=gen).requires_grad_(True)
        g = torch.randn((B, 1, H, W) if final else (B, H, W), generator=gen)
        items.append((name, r, final, ws, x, g))
    return items


def cpu_step(wl, items, split=None):
    """One fwd+bwd pass of the four stages with the reference's own operator sequence. `split`: a 2-element list
    that receives the seconds spent in the forward and in the backward calls."""
    from oracle import torch_port as tp
    for name, r, final, ws, x, g in items:
        for t in ws + [x]:
            t.grad = None
        t0 = time.perf_counter()
        if final:
            out = tp.chain_forward(x, ws, wl["max_depth"], True)
        else:
            out, _ = tp.stage_forward(x, ws, wl["max_depth"], r)
        t1 = time.perf_counter()
        out.backward(g)
        if split is not None:
            split[0] += t1 - t0
            split[1] += time.perf_counter() - t1


def time_cpu(wl, B, steps, warmup, threads=None):
    import torch
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    items = make_cpu_inputs(wl, B)
    for _ in range(warmup):
        cpu_step(wl, items)
    ts, fb = [], []
    for _ in range(steps):
        sp = [0.0, 0.0]
        t0 = time.perf_counter()
        cpu_step(wl, items, sp)
        ts.append(time.perf_counter() - t0)
        fb.append(sp)
    med = statistics.median(ts)
    px = B * wl["H"] * wl["W"]
    return dict(value=px / med / 1e6, unit="Mpixels/s", cores=cores, kind="port",
                fwd_s=statistics.median(f for f, _ in fb), bwd_s=statistics.median(b for _, b in fb),
                sample="batch %d of %dx%d, all four stages fwd+bwd, median of %d steps (%.3f s/step), "
                       "oracle/torch_port.py (reference ATen op sequence) with %d torch threads"
                       % (B, wl["H"], wl["W"], steps, med, cores)), med, ts


# ------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe), sampled with NVML on a side thread
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop, self._t = [], set(), None, threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it lists integers
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis and all(v.strip().isdigit() for v in vis.split(",")):
                index = int(vis.split(",")[index])
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                fn = getattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or self.nv.nvmlDeviceGetCurrentClocksThrottleReasons
                bits = fn(self.h)
                for b, n in self.REASONS.items():
                    if bits & b and n != "gpu_idle":
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.002)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def build_stages(wl, device, seed, dtype="f32"):
    """Modules with the reference constructor signatures + synthetic features resident in HBM."""
    import torch
    import bts_b200
    torch.manual_seed(seed)
    td = torch.bfloat16 if dtype == "bf16" else torch.float32
    H, W, B = wl["H"], wl["W"], wl["B"]
    stages = []
    for name, cin, nout, r, final in STAGES:
        red = bts_b200.reduction_1x1(cin, nout, wl["max_depth"], is_final=final).to(device)
        for m in red.modules():                                   # weights_init_xavier, bts.py:34-38
            if isinstance(m, torch.nn.Conv2d):
                torch.nn.init.xavier_uniform_(m.weight)
        lpg = None if final else bts_b200.local_planar_guidance(r)
        x = torch.randn(B, cin, H // r, W // r, device=device).to(td).requires_grad_(True)
        g = torch.randn((B, 1, H, W) if final else (B, H, W), device=device).to(td)
        stages.append(dict(name=name, r=r, final=final, red=red, lpg=lpg, x=x, g=g))
    return stages


def stage_forward(st):
    import bts_b200
    if st["final"]:
        return st["red"](st["x"])
    return bts_b200.fused_stage(st["red"], st["lpg"], st["x"])


def _flat_grads(params):
    """One tensor over the weight gradients of a stage when they already share a buffer (the autograd Functions carve
    them out of ONE zero-filled allocation, functional._zero_grads, and AccumulateGrad keeps those views), else None."""
    import torch
    gs = [p.grad for p in params]
    st = gs[0].untyped_storage()
    if any(g.untyped_storage().data_ptr() != st.data_ptr() or not g.is_contiguous() for g in gs):
        return None
    lo = min(g.storage_offset() for g in gs)
    hi = max(g.storage_offset() + g.numel() for g in gs)
    return torch.empty(0, dtype=gs[0].dtype, device=gs[0].device).set_(st, lo, (hi - lo,))   # padding words are zeros


def gpu_step(stages):
    """fwd + bwd of the four stages through the public module API (autograd drives the backward)."""
    outs, works, pending = [], [], []
    allreduce = _DIST["on"] and _DIST.get("allreduce")
    policy = _DIST.get("buckets", "two")

    def reduce_now(param_groups):
        # what DDP's buckets do (pytorch/bts_main.py:314): weight gradients are averaged over the ranks on NCCL's own
        # stream, under the kernels that follow. A group whose gradients share one buffer is reduced in place.
        import torch
        import torch.distributed as dist
        flats = [(_flat_grads(ps), ps) for ps in param_groups]
        if len(flats) == 1 and flats[0][0] is not None:
            works.append((dist.all_reduce(flats[0][0], op=dist.ReduceOp.AVG, async_op=True), flats[0][0], None))
            return
        params = [p for ps in param_groups for p in ps]           # several stages (or a foreign layout): one copy each way
        flat = torch.cat([p.grad.reshape(-1) for p in params])
        works.append((dist.all_reduce(flat, op=dist.ReduceOp.AVG, async_op=True), flat, params))

    for k, st in enumerate(stages):
        st["x"].grad = None
        for p in st["red"].parameters():
            p.grad = None
        out = stage_forward(st)
        out.backward(st["g"])
        outs.append(out)
        if allreduce:
            pending.append(list(st["red"].parameters()))
            last = k + 1 == len(stages)
            if policy == "per-stage" or last or (policy == "two" and k + 2 == len(stages)):
                reduce_now(pending)
                pending = []
    for work, flat, params in works:
        work.wait()                                               # the compute stream waits for the collective
        if params is not None:
            off = 0
            for p in params:
                p.grad.copy_(flat[off:off + p.numel()].view_as(p.grad))
                off += p.numel()
    if works:
        _DIST["bucket"] = [w[1] for w in works]
    return outs


def stage_calls(st, wl, device):
    """(forward, backward) closures that call the C ABI of one stage directly on resident buffers."""
    import torch
    from bts_b200 import _cabi
    from bts_b200 import functional as Fn
    lib = _cabi.lib()
    cur = lambda: torch.cuda.current_stream().cuda_stream      # looked up per call: the timing loop captures into a graph
    x, g, r, final = st["x"].detach(), st["g"], st["r"], st["final"]
    ws = [w.detach() for w in st["red"].chain_weights()]
    dims = [x.shape[1]] + [w.shape[0] for w in ws]
    B, _, h, w = x.shape
    W = _cabi.ptr_array([t.data_ptr() for t in ws])
    D = _cabi.int_array(dims)
    nws = lib.bts_workspace_bytes(D, len(ws), B, h, w, 0 if final else r)
    wsb = torch.empty((nws + 3) // 4, dtype=torch.float32, device=device)
    gx = torch.empty_like(x)
    gws = [torch.zeros_like(t) for t in ws]
    GW = _cabi.ptr_array([t.data_ptr() for t in gws])
    md = float(wl["max_depth"])
    nst = lib.bts_stash_bytes(D, len(ws), B, h, w)            # activation stash, as the autograd Functions use it
    stb = torch.empty((nst + 3) // 4, dtype=torch.float32, device=device) if nst else None
    stp = stb.data_ptr() if nst else None
    keep = (x, g, ws, wsb, gx, gws, W, D, GW, stb)
    if final:
        out = torch.empty(B, 1, h, w, device=device)
        f = lambda: _cabi.check(lib.bts_reduc_fwd_stash(x.data_ptr(), W, D, len(ws), 1, md, B, h, w, out.data_ptr(),
                                                        wsb.data_ptr(), nws, stp, nst, cur()), "bts_reduc_fwd")
        b = lambda: _cabi.check(lib.bts_reduc_bwd_stash(x.data_ptr(), W, D, len(ws), 1, md, B, h, w, g.data_ptr(), gx.data_ptr(),
                                                        GW, wsb.data_ptr(), nws, stp, nst, cur()), "bts_reduc_bwd")
    else:
        out = torch.empty(B, h * r, w * r, device=device)
        amin = torch.full((), float("inf"), device=device)
        keep += (amin,)
        f = lambda: _cabi.check(lib.bts_reduc_lpg_fwd_stash(x.data_ptr(), W, D, len(ws), md, B, h, w, r, Fn.EPS, None,
                                                            out.data_ptr(), amin.data_ptr(), wsb.data_ptr(), nws, stp, nst,
                                                            cur()), "bts_reduc_lpg_fwd")
        b = lambda: _cabi.check(lib.bts_reduc_lpg_bwd_stash(g.data_ptr(), x.data_ptr(), W, D, len(ws), md, B, h, w, r, Fn.EPS,
                                                            gx.data_ptr(), GW, wsb.data_ptr(), nws, stp, nst, cur()),
                                "bts_reduc_lpg_bwd")
    f.keep = b.keep = keep + (out,)
    return f, b


def time_back_to_back(fn, iters):
    """Average DEVICE duration of one call: `iters` calls back to back, captured once into a CUDA graph (the C ABI
    only enqueues on the stream it is given) and replayed between two CUDA events on the launching stream, so no
    host launch cost is in the figure -- a 5 us kernel is otherwise timed at the rate Python can call ctypes.
    Inputs of a call (>= 79 MB for the dominant one, 1.2 GB across a step) do not survive in L2 between calls of
    the big stages."""
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(iters):
            fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def stage_calls_api(st):
    """(forward, forward + backward) closures through the module API -- the bf16 twin's per-call timing: its backward of the
    128-channel chains runs the fp32 kernels on upcast tensors, and that conversion belongs in the figure."""
    import torch

    def f():
        with torch.no_grad():
            return stage_forward(st)

    def fb():
        st["x"].grad = None
        for p in st["red"].parameters():
            p.grad = None
        stage_forward(st).backward(st["g"])
    return f, fb


def time_calls(stages, wl, device, iters=20, dtype="f32"):
    """Per-call device times of each stage's forward and backward C-ABI call."""
    res = {}
    px = wl["B"] * wl["H"] * wl["W"]
    scale = 0.5 if dtype == "bf16" else 1.0                       # bf16 storage: half the algorithmic bytes
    for st in stages:
        if dtype == "bf16":
            f, fbk = stage_calls_api(st)
            tf = time_back_to_back(f, iters)
            tb = max(time_back_to_back(fbk, iters) - tf, 1e-6)
        else:
            f, b = stage_calls(st, wl, device)
            tf, tb = time_back_to_back(f, iters), time_back_to_back(b, iters)
        fb, bb = (scale * v for v in ALGO_BYTES[st["name"]])
        res[st["name"]] = dict(fwd_ms=tf, bwd_ms=tb, fwd_gbs=fb * px / tf / 1e6, bwd_gbs=bb * px / tb / 1e6)
    return res


def time_standalone_lpg(wl, device, iters=20):
    """Strict drop-in local_planar_guidance alone (SURVEY.md 8(d) report 4): plane_eq [B,4,h,w] -> depth
    [B,H,W] and back, per upratio; algorithmic bytes fwd 16/r^2 + 4, bwd 4 + 32/r^2 per full-res pixel."""
    import torch
    from bts_b200 import _cabi
    from bts_b200 import functional as Fn
    lib = _cabi.lib()
    cur = lambda: torch.cuda.current_stream().cuda_stream
    B, H, W = wl["B"], wl["H"], wl["W"]
    px = B * H * W
    res = {}
    for r in (8, 4, 2):
        h, w = H // r, W // r
        th = torch.rand(B, h, w, device=device) * 1.0
        ph = torch.rand(B, h, w, device=device) * 6.28
        pe = torch.stack([th.sin() * ph.cos(), th.sin() * ph.sin(), th.cos(), torch.rand(B, h, w, device=device) * 10], 1).contiguous()
        depth = torch.empty(B, H, W, device=device)
        g = torch.randn(B, H, W, device=device)
        gpe = torch.empty_like(pe)
        amin = torch.full((), float("inf"), device=device)
        f = lambda: _cabi.check(lib.bts_lpg_fwd(pe.data_ptr(), depth.data_ptr(), amin.data_ptr(), B, h, w, r, Fn.EPS, cur()), "bts_lpg_fwd")
        b = lambda: _cabi.check(lib.bts_lpg_bwd(g.data_ptr(), pe.data_ptr(), gpe.data_ptr(), B, h, w, r, Fn.EPS, cur()), "bts_lpg_bwd")
        tf, tb = time_back_to_back(f, iters), time_back_to_back(b, iters)
        fb, bb = 16.0 / (r * r) + 4.0, 4.0 + 32.0 / (r * r)
        res["lpg%dx%d_standalone" % (r, r)] = dict(fwd_ms=tf, bwd_ms=tb, fwd_gbs=fb * px / tf / 1e6, bwd_gbs=bb * px / tb / 1e6)
    return res


def measured_traffic(kernel):
    """DRAM bytes per launch of the dominant call from the committed ncu --set full capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return t.get(kernel, {}).get("dram_bytes_per_launch")
    except Exception:
        return None


def e2e_leg(stages, wl, steps, warmup, device, pipelined=True, d2h="all"):
    """Same step through the public module API with HOST buffers: every step copies the features and
    upstream gradients from pinned host memory, and reads every result (depth maps / head output,
    grad_x, grad_W, abs_min) back to pinned host memory, inside the timed region.

    pipelined: three streams (upload / compute / download) and two sets of device input and host output
    buffers, so step i+1's upload and step i-1's download overlap step i (PCIe is full duplex); every
    dependency is an event, the timed region ends with a full device synchronisation. serial: one stream."""
    import torch
    host_in = []
    for st in stages:
        host_in.append((st["x"].detach().cpu().pin_memory(), st["g"].cpu().pin_memory()))
    h2d = sum(hx.numel() * hx.element_size() + hg.numel() * hg.element_size() for hx, hg in host_in)
    nset = 2 if pipelined else 1
    dev_in = [[(torch.empty_like(st["x"]).requires_grad_(True), torch.empty_like(st["g"])) for st in stages]
              for _ in range(nset)]
    host_out = [None] * nset
    d2h_bytes = [0]
    cur = torch.cuda.current_stream()
    s_in, s_comp, s_out = (torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()) if pipelined else (cur, cur, cur)
    ev_free = [None] * nset          # compute of the step that last used input set j has finished

    def step(i):
        j = i % nset
        ev_in, ev_comp, results = [], [], []
        with torch.cuda.stream(s_in):
            if ev_free[j] is not None:
                s_in.wait_event(ev_free[j])
            for (hx, hg), (dx, dg) in zip(host_in, dev_in[j]):
                with torch.no_grad():
                    dx.copy_(hx, non_blocking=True)
                    dg.copy_(hg, non_blocking=True)
                e = torch.cuda.Event()
                e.record(s_in)
                ev_in.append(e)
        with torch.cuda.stream(s_comp):
            for k, (st, (dx, dg)) in enumerate(zip(stages, dev_in[j])):
                s_comp.wait_event(ev_in[k])
                dx.grad = None
                for p in st["red"].parameters():
                    p.grad = None
                saved = st["x"]
                st["x"] = dx
                out = stage_forward(st)
                out.backward(dg)
                st["x"] = saved
                res = [out.detach()] + ([dx.grad] if d2h == "all" else []) + [p.grad for p in st["red"].parameters()]
                if st["lpg"] is not None:
                    res.append(st["lpg"].abs_min)
                if pipelined:
                    for t in res:
                        t.record_stream(s_out)
                results.append(res)
                e = torch.cuda.Event()
                e.record(s_comp)
                ev_comp.append(e)
            ev_free[j] = ev_comp[-1]
        if host_out[j] is None:
            host_out[j] = [[torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in res] for res in results]
            d2h_bytes[0] = sum(t.numel() * t.element_size() for res in results for t in res)
        with torch.cuda.stream(s_out):
            for k, (res, hres) in enumerate(zip(results, host_out[j])):
                s_out.wait_event(ev_comp[k])
                for t, ht in zip(res, hres):
                    ht.copy_(t, non_blocking=True)

    for i in range(max(2, warmup)):
        step(i)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return dt, h2d, d2h_bytes[0], host_out


_DIST = {"on": False}


def barrier():
    if _DIST["on"]:
        import torch.distributed as dist
        dist.barrier()


def max_over_ranks(value, device):
    if not _DIST["on"]:
        return value
    import torch
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_reference_arm(args, wl, rank):
    """--impl reference: the reference's own CPU implementation of the path (torch port of
    pytorch/bts.py:97-173, pinned bit-exact to it by tests/golden -- the reference tree itself does not travel to the
    GPU box) on all host cores, SAME per-step batch as our arm (a step is 0.4 s at nyu_b8). Rank 0 only."""
    if rank != 0:
        return
    B = wl["B"]
    cb, med, ts = time_cpu(wl, B, args.steps, args.warmup)
    one, _, _ = time_cpu(wl, 1, 3, 1, threads=1)                   # the 1-thread figure (batch 1)
    cb["one_thread_mpix_s"] = one["value"]
    line = {
        "impl": "reference", "metric": "LPG decoder Mpixels/s fwd+bwd", "value": cb["value"], "unit": "Mpixels/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": med * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "per_gpu_batch": B, "H": wl["H"], "W": wl["W"], "max_depth": wl["max_depth"],
                   "stages": [s[0] for s in STAGES],
                   "note": "one CPU process whatever --gpus says: compare with our arm at N=1 only"},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "Mpixels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    wl = WORKLOADS[args.workload]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1 and "RANK" not in os.environ:
        # convenience: `python bench.py --gpus N` re-launches itself one process per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))

    # stdout carries ONE JSON line and nothing else: whatever a library writes to file descriptor 1 (NCCL's version
    # banner, for one) is sent to stderr; print() below goes through the saved descriptor.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout

    if args.impl == "reference":
        run_reference_arm(args, wl, rank)
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: bts_b200 has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    from bts_b200 import _cabi
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG", "WARN")    # stdout is already redirected; NCCL_DEBUG=INFO from the env is kept
        dist.init_process_group("nccl", device_id=device)
        _DIST["on"] = True
        _DIST["allreduce"] = bool(args.grad_allreduce)
        _DIST["buckets"] = args.allreduce_buckets
    lib = _cabi.lib()

    stages = build_stages(wl, device, seed=rank, dtype=args.dtype)
    bscale = 0.5 if args.dtype == "bf16" else 1.0
    px_per_rank = wl["B"] * wl["H"] * wl["W"]
    in_bytes = sum((st["x"].numel() + st["g"].numel()) * st["x"].element_size() for st in stages)

    for _ in range(max(3, args.warmup)):
        gpu_step(stages)
    torch.cuda.synchronize()
    # The step is ~30 short launches behind Python/autograd dispatch: capture it once (the C ABI only
    # enqueues on the given stream) and replay the graph, so the timed loop is device-bound.
    graph, launch_mode, per_step = None, "eager", None
    if not args.no_graph:
        try:
            l0 = lib.bts_launch_count()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                gpu_step(stages)
            per_step = lib.bts_launch_count() - l0
            for _ in range(2):
                graph.replay()
            torch.cuda.synchronize()
            launch_mode = "cuda_graph"
        except Exception as exc:  # noqa: BLE001 -- report and fall back to eager launches (same kernels)
            sys.stderr.write("bench: CUDA graph capture failed (%s); timing eager launches\n" % exc)
            graph = None
            torch.cuda.synchronize()
    barrier()
    launches0 = lib.bts_launch_count()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with ClockSampler(local_rank) as clocks:
        torch.cuda.synchronize()
        evs[0].record()
        for i in range(args.steps):
            if graph is not None:
                graph.replay()
            else:
                gpu_step(stages)
            evs[i + 1].record()                                   # an event record costs nothing on the device
        torch.cuda.synchronize()
    barrier()
    launches = per_step * args.steps if graph is not None else lib.bts_launch_count() - launches0
    ms_total = max_over_ranks(evs[0].elapsed_time(evs[-1]), device)
    ms_step = ms_total / args.steps
    # per-step times beside the total: one rank's hiccup inside a 15 ms region shows up as max >> median
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    ms_median = max_over_ranks(statistics.median(per), device)
    ms_worst = max_over_ranks(max(per), device)
    value = world * px_per_rank / (ms_step * 1e-3) / 1e6

    if args.kernel_only:
        if rank == 0:
            print(json.dumps({"metric": "LPG decoder Mpixels/s fwd+bwd", "value": value, "ms_per_step": ms_step,
                              "gpu_launches": int(launches), "launch": launch_mode}), flush=True)
        return

    # per-call device times -> roofline of the dominant call (the one that moves the most bytes)
    calls = time_calls(stages, wl, device, dtype=args.dtype)
    peak, peak_src = hbm_peak()
    dom = max(((n, d, k) for n, d in calls.items() for k in ("fwd", "bwd")),
              key=lambda t: ALGO_BYTES[t[0]][0 if t[2] == "fwd" else 1])
    dn, dd, dk = dom
    algo_b = bscale * ALGO_BYTES[dn][0 if dk == "fwd" else 1] * px_per_rank
    achieved = dd[dk + "_gbs"]
    step_bytes = bscale * sum(sum(ALGO_BYTES[n]) for n in calls) * px_per_rank
    roofline = {"bound": "hbm", "kernel": "%s %s" % (dn, dk), "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": measured_traffic("%s %s" % (dn, dk)), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": algo_b, "launch_ms": dd[dk + "_ms"],
                "whole_step": {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms_step * 1e-3) / 1e9,
                               "frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak},
                "per_call": calls, "standalone_lpg": time_standalone_lpg(wl, device) if args.dtype == "f32" else None}
    # SURVEY.md 8(d) reports 2 and 3 from the per-call device times: the three LPG stages, and all four
    for key, names in (("lpg3", [n for n in calls if "lpg" in n]), ("lpg3_head", list(calls))):
        ms = sum(calls[n]["fwd_ms"] + calls[n]["bwd_ms"] for n in names)
        by = bscale * sum(sum(ALGO_BYTES[n]) for n in names) * px_per_rank
        roofline[key] = {"ms": ms, "mpixels_per_s": px_per_rank / ms / 1e3, "achieved": by / ms / 1e6, "frac": by / ms / 1e6 / peak}

    e2e = None
    if not args.no_e2e:
        e_steps = max(2, min(args.steps, 10))
        dt, h2d, d2h, _ = e2e_leg(stages, wl, e_steps, 2, device, pipelined=not args.e2e_serial, d2h=args.e2e_d2h)
        dt = max_over_ranks(dt, device)
        e2e = {"value": world * px_per_rank * e_steps / dt / 1e6, "unit": "Mpixels/s",
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e_steps,
               "ms_per_step": dt / e_steps * 1e3,
               "mode": "serial (one stream)" if args.e2e_serial else
                       "pipelined (upload / compute / download streams, double-buffered inputs and host outputs)",
               "what": "module API with pinned HOST buffers: H2D of features + upstream grads, D2H of " +
                       ("every output (depths, grad_x, grad_W, abs_min)" if args.e2e_d2h == "all" else
                        "the results the reference brings to the host (depths, grad_W, abs_min; grad_x stays on the GPU)") +
                       " inside the timed region"}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_baseline, _, _ = time_cpu(wl, wl["B"], 10, 3)          # same batch as the GPU arm, median of 10
        one, _, _ = time_cpu(wl, 1, 3, 1, threads=1)
        cpu_baseline["one_thread_mpix_s"] = one["value"]

    if rank == 0:
        line = {
            "metric": "LPG decoder Mpixels/s fwd+bwd", "value": value, "unit": "Mpixels/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step,
            "ms_per_step_median": ms_median, "ms_per_step_max": ms_worst, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": wl["desc"], "per_gpu_batch": wl["B"], "H": wl["H"], "W": wl["W"],
                       "max_depth": wl["max_depth"], "stages": [s[0] for s in STAGES], "parallelism": "batch-sharded x%d" % world + (
                           " + NCCL all-reduce(avg) of the weight gradients every step (%s)" % _DIST.get("buckets") if _DIST.get("allreduce") else ""),
                       "l2": "inputs %.0f MB per step > 126 MB L2 (no flush needed)" % (in_bytes / 1e6),
                       "api": "bts_b200.reduction_1x1 / fused_stage via libbts_b200.so C ABI",
                       "launch": launch_mode},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(launches) * world,
            "clocks": clocks.summary(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        # a captured graph that holds NCCL work (--grad-allreduce) must go before its communicator does
        graph = None
        _DIST.pop("bucket", None)
        torch.cuda.synchronize()
        barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
