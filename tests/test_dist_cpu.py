"""N>1 host logic on CPU: world_size-2 gloo run of the gradient bucketing (wav2vec2forbrain_b200.dist)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class Toy(nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(3)
        self.a = nn.Linear(6, 5)
        self.unused = nn.Linear(3, 3)      # never receives a gradient (inpLayer* / hidden_start analogue)
        self.b = nn.Linear(5, 2)
        self.skipped = nn.Linear(5, 5)     # LayerDrop analogue: used on some steps only

    def forward(self, x, use_skipped):
        h = torch.tanh(self.a(x))
        if use_skipped:
            h = h + self.skipped(h)
        return self.b(h).pow(2).mean()


def _worker(rank, world, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from wav2vec2forbrain_b200 import dist as bdist
    bdist.init_process_group("gloo")
    m = Toy()
    with torch.no_grad():
        for p in m.parameters():
            p.add_(rank)                   # ranks start different; DataParallel must broadcast rank 0's weights
    dp = bdist.DataParallel(m, bucket_bytes=64, overlap=False)
    opt = torch.optim.SGD(m.parameters(), lr=0.1)
    dp.attach_optimizer(opt)
    g = torch.Generator().manual_seed(100)
    data = torch.randn(2, 8, 6, generator=g)
    res = []
    for step, use in enumerate([True, False]):
        opt.zero_grad()
        loss = m(data[step][bdist.shard_indices(8, rank, world)], use)
        loss.backward()
        dp.finish()
        dp.finish()                        # idempotent within a step
        res.append({k: (None if p.grad is None else p.grad.clone()) for k, p in m.named_parameters()})
        opt.step()
    torch.save((rank, res, {k: v.detach().clone() for k, v in m.named_parameters()}), os.path.join(outdir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_matches_global_batch_gradient(tmp_path):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    outs = [torch.load(os.path.join(str(tmp_path), f"r{r}.pt"), weights_only=False) for r in range(world)]
    # single-process reference on the full batch with rank 0's initial weights
    ref = Toy()
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    g = torch.Generator().manual_seed(100)
    data = torch.randn(2, 8, 6, generator=g)
    for step, use in enumerate([True, False]):
        opt.zero_grad()
        ref(data[step], use).backward()
        for r in range(world):
            grads = outs[r][1][step]
            for k, p in ref.named_parameters():
                if p.grad is None:
                    assert grads[k] is None or grads[k].abs().sum() == 0, k
                else:
                    assert torch.allclose(grads[k], p.grad, atol=1e-6), (k, step)
        opt.step()
    for k, p in ref.named_parameters():
        assert torch.allclose(outs[0][2][k], p.detach(), atol=1e-6) and torch.allclose(outs[1][2][k], p.detach(), atol=1e-6)


class _NodeFn(torch.autograd.Function):
    """One autograd node that returns the gradients of several parameters at once and hands them to the sink while
    its backward is still running (the shape of the engine's encoder node)."""

    @staticmethod
    def forward(ctx, buckets, x, wa, wb):
        ctx.buckets = buckets
        ctx.save_for_backward(x, wa, wb)
        return ((x @ wa.t()) @ wb.t()).pow(2).mean()

    @staticmethod
    def backward(ctx, dl):
        x, wa, wb = ctx.saved_tensors
        with torch.enable_grad():
            wa_, wb_ = wa.detach().requires_grad_(), wb.detach().requires_grad_()
            l = ((x @ wa_.t()) @ wb_.t()).pow(2).mean()
            gb, = torch.autograd.grad(l, wb_, retain_graph=True)
            gb = gb * dl
            ctx.buckets.sink([("wb", gb)])             # "layer above" finished first
            ga, = torch.autograd.grad(l, wa_)
            ga = ga * dl
            ctx.buckets.sink([("wa", ga)])
        ctx.buckets.sync_early()
        return None, None, ga, gb


def _worker_sink(rank, world, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from wav2vec2forbrain_b200 import dist as bdist
    bdist.init_process_group("gloo")
    torch.manual_seed(5)
    wa, wb = nn.Parameter(torch.randn(5, 6)), nn.Parameter(torch.randn(2, 5))
    other = nn.Parameter(torch.randn(3))           # reduced by the ordinary hook path in the same step
    gb = bdist.GradientBuckets([wa, wb, other], bucket_bytes=16, process_group=None, overlap=False)
    gb.param_of_name = {"wa": wa, "wb": wb}
    g = torch.Generator().manual_seed(100)
    data = torch.randn(8, 6, generator=g)
    x = data[bdist.shard_indices(8, rank, world)]
    loss = _NodeFn.apply(gb, x, wa, wb) + (other * (rank + 1)).sum()
    loss.backward()
    gb.finish()
    torch.save((wa.grad.clone(), wb.grad.clone(), other.grad.clone()), os.path.join(outdir, f"s{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_early_sink_reduces_inside_the_node_and_hooks_skip_those_parameters(tmp_path):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker_sink, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    outs = [torch.load(os.path.join(str(tmp_path), f"s{r}.pt"), weights_only=False) for r in range(world)]
    torch.manual_seed(5)
    wa, wb = torch.randn(5, 6, requires_grad=True), torch.randn(2, 5, requires_grad=True)
    g = torch.Generator().manual_seed(100)
    data = torch.randn(8, 6, generator=g)
    ((data @ wa.t()) @ wb.t()).pow(2).mean().backward()       # equal shards: mean of per-rank means = global mean
    for r in range(world):
        assert torch.allclose(outs[r][0], wa.grad, atol=1e-6)
        assert torch.allclose(outs[r][1], wb.grad, atol=1e-6)   # averaged once, not twice
        assert torch.allclose(outs[r][2], torch.full((3,), 1.5), atol=1e-6)
