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
    dp.attach_optimizer(opt, lazy=True)    # lazy needs FusedAdam + the arena buckets: anything else keeps the full wait
    assert not getattr(dp.buckets, "lazy", False)
    bdist.tune_gc()
    g = torch.Generator().manual_seed(100)
    data = torch.randn(2, 8, 6, generator=g)
    res = []
    for step, use in enumerate([True, False]):
        opt.zero_grad()
        loss = m(data[step][bdist.shard_indices(8, rank, world)], use)
        loss.backward()                    # the end-of-backward callback finishes the reduction (no finish() call:
        #                                    clip_grad_norm_ of the unmodified Trainer, train_loop.py:72-75, sees averages)
        tn = torch.nn.utils.clip_grad_norm_(m.parameters(), 1e9)
        dp.finish()                        # idempotent within a step
        assert abs(tn.item() - torch.nn.utils.clip_grad_norm_(m.parameters(), 1e9).item()) < 1e-6
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


def _product_model(rank, world, port):
    """The product model classes on the ABI emulator (CPU), weights offset per rank (DataParallel must broadcast rank 0's)."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from tests import abi_emulator
    from tests.golden_util import load_case
    from wav2vec2forbrain_b200 import dist as bdist, engine, models
    from wav2vec2forbrain_b200.config import W2VConfig
    engine.ops = abi_emulator
    models._EMULATED = True
    bdist.init_process_group("gloo")
    rec, _, _, sd, batch = load_case("tiny_base")
    be, w = rec["be"], rec["w2v"]
    ck = "golden/tiny_base_dp"
    models.register_architecture(ck, lambda **kw: W2VConfig(**dict(dict(
        kind="base", hidden=w["hidden"], layers=w["layers"], heads=w["heads"], ffn=w["ffn"], vocab=w["vocab"],
        pos_conv_k=w["pos_conv_k"], pos_conv_groups=w["pos_conv_groups"], hidden_dropout=0.0, attention_dropout=0.0,
        activation_dropout=0.0, final_dropout=0.0, layerdrop=0.5), **kw)), be["out_size"])
    args = models.BrainEncoderArgs(encoder_gru_hidden_size=be["gru_hidden"], encoder_num_gru_layers=be["gru_layers"],
                                   encoder_fc_hidden_sizes=tuple(be["fc_hidden"]), gaussian_smooth_width=be["smooth_sigma"])
    brain = models.B2P2TModel(args, models.B2TBrainFeatureExtractor(args, ck, 256 * 32), wav2vec_checkpoint=ck)
    m = models.W2VBrainEncoderModel(args, brain, ck, None, True)
    m.load_state_dict(sd, strict=True)
    with torch.no_grad():
        for p in m.parameters():
            p.add_(0.01 * rank)                      # DataParallel must broadcast rank 0's weights
    B = batch["input"].shape[0]
    sel = [rank % B]                                 # every rank gets a shard of the same size (B = 3 in the golden)
    mb = models.B2tSampleBatch(batch["input"][sel], batch["target"][sel])
    mb.day_idxs, mb.input_lens, mb.target_lens = batch["day_idxs"][sel], batch["input_lens"][sel], batch["target_lens"][sel]
    return m, mb, bdist


def _worker_model(rank, world, port, outdir):
    """flat gradient arena + in-place bucket all-reduce over gloo"""
    m, mb, bdist = _product_model(rank, world, port)
    dp = bdist.DataParallel(m, bucket_bytes=4096, overlap=False)
    assert isinstance(dp.buckets, bdist.ArenaBuckets) and len(dp.buckets.buckets) > 4
    m.train()
    res = []
    for step in range(3):
        m.zero_grad(set_to_none=True)
        # local gradients of this shard, no collective
        dp.buckets.enabled = False
        st = m.runtime.layerdrop_gen.get_state()
        seed_step = m.runtime.step
        m.forward(mb).loss.backward()
        local = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
        # the same step again (same LayerDrop draw, same seeds) with the in-place all-reduce
        m.zero_grad(set_to_none=True)
        dp.buckets.enabled = True
        m.runtime.layerdrop_gen.set_state(st)
        m.runtime.step = seed_step
        m.forward(mb).loss.backward()
        tn = torch.nn.utils.clip_grad_norm_(m.parameters(), 1e9)       # no finish() call before it
        red = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
        assert all(m.runtime.grad_arena().owns(k, p.grad) for k, p in m.named_parameters() if p.grad is not None)
        res.append((local, red, tn.item()))
    torch.save((rank, res, m.runtime.engine.arena_copies), os.path.join(outdir, f"m{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_product_model_arena_buckets_average_in_place(tmp_path):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker_model, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    outs = [torch.load(os.path.join(str(tmp_path), f"m{r}.pt"), weights_only=False) for r in range(world)]
    saw_skip = False
    for step in range(3):
        l0, r0, n0 = outs[0][1][step]
        l1, r1, n1 = outs[1][1][step]
        assert set(l0) == set(l1) == set(r0) == set(r1)           # LayerDrop decisions identical on both ranks
        saw_skip = saw_skip or not any(".layers.0." in k for k in r0) or not any(".layers.1." in k for k in r0)
        for k in l0:
            want = 0.5 * (l0[k] + l1[k])
            assert torch.allclose(r0[k], want, rtol=1e-5, atol=1e-8), (step, k)
            assert torch.equal(r0[k], r1[k]), (step, k)
        total = torch.sqrt(sum((v.double() ** 2).sum() for v in r0.values())).item()
        assert abs(n0 - total) < 1e-5 * total and abs(n1 - total) < 1e-5 * total   # clip saw the averaged gradients
    assert saw_skip, "layerdrop=0.5 over 3 steps should have skipped a layer at least once (seeded)"


class _GroupedSGD(torch.optim.Optimizer):
    """Stand-in for FusedAdam's data-parallel protocol (optim.py `group_wait` / `group_order`, CUDA only): one update per
    parameter group, in group_order, each after group_wait(group index)."""

    def __init__(self, groups):
        super().__init__(groups, dict(lr=0.1))
        self.group_wait = None
        self.group_order = None
        self.waited = []

    @torch.no_grad()
    def step(self, closure=None):
        for gi in (self.group_order or range(len(self.param_groups))):
            self.group_wait(gi)
            self.waited.append(gi)
            for p in self.param_groups[gi]["params"]:
                if p.grad is not None:
                    p.add_(p.grad, alpha=-self.param_groups[gi]["lr"])


def _worker_lazy(rank, world, port, outdir):
    """lazy waits (bench.py's default under data parallel): the end of backward() leaves the buckets in flight, the optimizer
    waits per parameter group (wav2vec2 encoder first), its post-hook closes the step; a backward pass that starts without an
    optimizer step in between closes the previous reduction itself"""
    m, mb, bdist = _product_model(rank, world, port)
    dp = bdist.DataParallel(m, bucket_bytes=4096, overlap=False)
    opt = _GroupedSGD([{"params": list(m.brain_encoder.parameters()), "lr": 0.05},          # the reference's two groups
                       {"params": list(m.w2v_encoder.parameters()), "lr": 0.01}])           # (b2t_gru_w2v_experiment.py:111-128)
    dp.attach_optimizer(opt, lazy=True)
    bk = dp.buckets
    assert bk.lazy and opt.group_order == [1, 0]
    m.train()
    res = []
    for step in range(3):
        m.zero_grad(set_to_none=True)
        m.forward(mb).loss.backward()
        assert bk._ended and bk._active and all(bk._launched[bi] for bi in range(len(bk.buckets)) if bk._seen[bi])
        if step == 1:
            # no optimizer step: the next backward pass has to finish this one first (sink), and must not see
            # "arrived twice"
            m.zero_grad(set_to_none=True)
            m.forward(mb).loss.backward()
            assert bk._ended and bk._active
        grads = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}
        opt.step()
        assert not bk._ended and not bk._active and bk._work == []
        res.append(grads)
    assert opt.waited == [1, 0] * 3
    torch.save((rank, res, {k: v.detach().clone() for k, v in m.named_parameters()}), os.path.join(outdir, f"z{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_lazy_group_waits_keep_ranks_identical(tmp_path):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker_lazy, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
        assert p.exitcode == 0
    outs = [torch.load(os.path.join(str(tmp_path), f"z{r}.pt"), weights_only=False) for r in range(world)]
    for step in range(3):
        g0, g1 = outs[0][1][step], outs[1][1][step]
        assert set(g0) == set(g1)
        for k in g0:
            assert torch.equal(g0[k], g1[k]), (step, k)             # both ranks hold the averaged gradient
    for k, v in outs[0][2].items():
        assert torch.equal(v, outs[1][2][k]), k                     # and the same weights after three steps
