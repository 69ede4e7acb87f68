"""CPU, world size 2 over gloo (127.0.0.1): the host-side logic of the N > 1 bench path -- batch sharding with no
data-path collective, max-over-ranks timing, and the per-stage weight-gradient all-reduce of `--grad-allreduce`
(BASELINE.json configs[3]; DDP's role in pytorch/bts_main.py:314), which reduces the gradients IN PLACE through one
flat view when they share the buffer functional._zero_grads carves them from. The kernels themselves need a GPU and are
covered by the -m gpu tests; here the "gradients" are filled by hand."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        from bts_b200 import functional as Fn
        bench._DIST["on"] = True
        cpu = torch.device("cpu")
        # max over ranks: the slowest rank's time is the step time
        assert bench.max_over_ranks(1.0 + rank, cpu) == float(world)
        # weight gradients of one stage as the autograd Functions produce them: views of ONE zero-filled buffer
        shapes = [(16, 32, 1, 1), (8, 16, 1, 1), (1, 8, 1, 1)]
        params = [torch.nn.Parameter(torch.zeros(s)) for s in shapes]
        views = Fn._zero_grads([p.data for p in params])
        for i, (p, v) in enumerate(zip(params, views)):
            v.fill_(float((rank + 1) * (i + 1)))
            p.grad = v
        flat = bench._flat_grads(params)
        assert flat is not None and flat.numel() >= sum(p.numel() for p in params)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)                # gloo has no AVG; NCCL's AVG = SUM / world
        flat.div_(world)
        mean = sum(r + 1 for r in range(world)) / world
        for i, p in enumerate(params):
            assert torch.all(p.grad == mean * (i + 1)), (rank, i)  # reduced in place, through the views
        # gradients that do not share a buffer take the copying route
        for p in params:
            p.grad = torch.ones_like(p)
        assert bench._flat_grads(params) is None
        # batch sharding: every rank draws its own batch (seed = rank), whole-job pixels = world x per-rank pixels
        wl = bench.WORKLOADS["nyu_train_b4"]
        a = bench.make_cpu_inputs(dict(wl, H=32, W=64), 1, seed=rank)[0][4]
        gathered = [torch.empty_like(a) for _ in range(world)]
        dist.all_gather(gathered, a.detach())
        assert not torch.equal(gathered[0], gathered[1])
        dist.barrier()
        out[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_host_logic():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: "ok", 1: "ok"}


def test_reference_arm_runs_on_rank_0_only(capsys):
    sys.path.insert(0, ROOT)
    import bench

    class A:
        steps, warmup, gpus = 1, 0, 2
    bench.run_reference_arm(A, dict(bench.WORKLOADS["nyu_train_b4"], B=1, H=32, W=64), rank=1)
    assert capsys.readouterr().out == ""
