"""Data-parallel training across the GPUs of one box: one process per GPU, torch.distributed (NCCL
over NVLink 5 / NVSwitch), bucketed gradient all-reduce launched as soon as a bucket's gradients
exist, on a side stream, overlapped with the rest of backward.

The reference is single-process (no torch.distributed anywhere); the semantics kept
are those of SURVEY.md §5.4 / §8(e): each rank runs the reference step on its shard of the batch,
gradients are averaged (CTC reduction is a batch mean, so equal shards reproduce the global-batch
gradient), parameters that never receive a gradient (unused inpLayer*, hidden_start, conformer
pos_conv_embed, LayerDrop-skipped layers) are tolerated, LayerDrop decisions are identical on all
ranks, dropout masks differ per rank.

Our backward produces gradients in two bursts (w2v encoder + head first, brain encoder second: the
two autograd nodes of models.py), so the w2v buckets travel while the brain-encoder backward
(GRU recurrence, ~30 % of the step) still computes.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist
from torch import nn


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def init_process_group(backend: Optional[str] = None) -> None:
    """Rendezvous from torchrun's environment (MASTER_ADDR/PORT, RANK, WORLD_SIZE)."""
    if dist.is_initialized():
        return
    rank, local_rank, world = env_rank_world()
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    dist.init_process_group(backend=backend, rank=rank, world_size=world)


def shard_indices(n: int, rank: int, world: int) -> List[int]:
    """rank r takes indices r::world of the (identically shuffled) epoch order."""
    return list(range(rank, n, world))


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class GradientBuckets:
    """Bucketed asynchronous all-reduce(mean) of parameter gradients.

    Buckets are formed over the parameters that may receive a gradient, in reverse registration
    order (≈ reverse execution order: lm_head, encoder layers L-1..0, FC, GRU, day weights), capped
    at `bucket_bytes`.  A bucket is launched from the post-accumulate-grad hook of its last arriving
    parameter; parameters whose gradient never arrives (LayerDrop, unused) are flushed as zeros at
    `finish()`, which every rank calls before optimizer.step / clip_grad_norm_."""

    def __init__(self, params: List[nn.Parameter], bucket_bytes: int = 48 << 20, process_group=None,
                 overlap: bool = True):
        self.group = process_group
        self.world = dist.get_world_size(process_group)
        self.params = [p for p in params if p.requires_grad]
        self.overlap = overlap and torch.cuda.is_available()
        self.stream = torch.cuda.Stream() if self.overlap else None
        self.bucket_bytes = bucket_bytes
        # early path (sink / sync_early): gradients handed over by the engine while the backward schedule of the
        # same autograd node is still being issued; keyed by parameter name
        self.param_of_name: Dict[str, nn.Parameter] = {}
        self._early: List[torch.Tensor] = []
        self._early_bytes = 0
        self._early_work = []
        self._early_done = set()
        self.buckets: List[List[nn.Parameter]] = []
        cur, size = [], 0
        for p in reversed(self.params):
            nb = p.numel() * p.element_size()
            if cur and size + nb > bucket_bytes:
                self.buckets.append(cur)
                cur, size = [], 0
            cur.append(p)
            size += nb
        if cur:
            self.buckets.append(cur)
        self._bucket_of: Dict[int, int] = {}
        for bi, b in enumerate(self.buckets):
            for p in b:
                self._bucket_of[id(p)] = bi
        self._pending = [0] * len(self.buckets)
        self._launched = [False] * len(self.buckets)
        self._work = []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.reset()

    def reset(self) -> None:
        self._pending = [len(b) for b in self.buckets]
        self._launched = [False] * len(self.buckets)
        self._work = []
        self._active = False
        self._early_done = set()

    # ---- early path ---------------------------------------------------------------------------------
    def sink(self, named) -> None:
        """named: [(parameter name, finished gradient tensor)].  The three autograd nodes of the model return all
        their gradients at once, so post-accumulate hooks alone cannot overlap anything with the node's own
        backward; the engine calls this after every encoder layer instead.  Full buckets start their all-reduce on
        the side stream right away (behind everything issued so far on the compute stream)."""
        for name, g in named:
            p = self.param_of_name.get(name)
            if p is None or not p.requires_grad:
                continue
            self._active = True
            self._early.append(g)
            self._early_done.add(id(p))
            self._early_bytes += g.numel() * g.element_size()
        if self._early_bytes >= self.bucket_bytes:
            self._flush_early()

    def _flush_early(self) -> None:
        grads, self._early, self._early_bytes = self._early, [], 0
        if not grads:
            return
        if self.overlap:
            self.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.stream):
                flat = torch.cat([g.reshape(-1) for g in grads])
                flat.div_(self.world)
                work = dist.all_reduce(flat, group=self.group, async_op=True)
        else:
            flat = torch.cat([g.reshape(-1) for g in grads])
            flat.div_(self.world)
            dist.all_reduce(flat, group=self.group)
            work = None
        self._early_work.append((work, flat, grads))

    def sync_early(self) -> None:
        """End of an autograd node: reduce what is left and write every averaged value back into the tensors
        the node is about to return, ordered before anything autograd does with them."""
        self._flush_early()
        for work, flat, grads in self._early_work:
            ctx = torch.cuda.stream(self.stream) if self.overlap else _null()
            with ctx:
                if work is not None:
                    work.wait()
                off = 0
                for g in grads:
                    n = g.numel()
                    g.copy_(flat[off:off + n].view_as(g))
                    off += n
        if self.overlap and self._early_work:
            torch.cuda.current_stream().wait_stream(self.stream)
        self._early_work = []

    def _on_grad(self, p: nn.Parameter) -> None:
        self._active = True
        bi = self._bucket_of[id(p)]
        self._pending[bi] -= 1
        if self._pending[bi] == 0 and not self._launched[bi]:
            self._launch(bi)

    def _launch(self, bi: int) -> None:
        self._launched[bi] = True
        grads = [p.grad for p in self.buckets[bi] if p.grad is not None and id(p) not in self._early_done]
        if not grads:
            return
        if self.overlap:
            self.stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.stream):
                flat = torch.cat([g.reshape(-1) for g in grads])
                flat.div_(self.world)
                work = dist.all_reduce(flat, group=self.group, async_op=True)
            self._work.append((work, flat, grads))
        else:
            flat = torch.cat([g.reshape(-1) for g in grads])
            flat.div_(self.world)
            dist.all_reduce(flat, group=self.group)
            self._work.append((None, flat, grads))

    def finish(self) -> None:
        """Launch what has not been launched (identical bucket composition on every rank is guaranteed by
        rank-consistent LayerDrop), wait, scatter the averaged values back into .grad."""
        if not self._active:
            return   # nothing arrived since the last finish(): second call of the same step
        for bi in range(len(self.buckets)):
            if not self._launched[bi]:
                self._launch(bi)
        for work, flat, grads in self._work:
            if self.overlap:
                with torch.cuda.stream(self.stream):
                    if work is not None:
                        work.wait()
                    off = 0
                    for g in grads:
                        n = g.numel()
                        g.copy_(flat[off:off + n].view_as(g))
                        off += n
            else:
                off = 0
                for g in grads:
                    n = g.numel()
                    g.copy_(flat[off:off + n].view_as(g))
                    off += n
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.stream)
        self.reset()

    def remove(self) -> None:
        for h in self._hooks:
            h.remove()


class DataParallel:
    """Wraps a model for the unmodified reference Trainer loop (train_loop.py:44-78): broadcasts the
    initial weights/buffers from rank 0, installs the bucket hooks, and waits for the all-reduce in an
    optimizer step pre-hook (call `finish()` yourself before clip_grad_norm_)."""

    def __init__(self, model: nn.Module, bucket_bytes: int = 48 << 20, process_group=None, overlap: bool = True,
                 early_reduce: Optional[bool] = None):
        self.model = model
        self.group = process_group
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t.data, src=0, group=process_group)
        rt = getattr(model, "runtime", None)
        if rt is not None:
            rt.rank = dist.get_rank(process_group)   # per-rank dropout masks; LayerDrop generator stays shared
        self.buckets = GradientBuckets(list(model.parameters()), bucket_bytes, process_group, overlap)
        # early_reduce: start each bucket's all-reduce from inside the encoder's backward schedule (engine.emit_grads)
        # instead of after the whole autograd node.  Off by default: measured on 2 B200 (conformer-large, B = 32 per
        # GPU) the early path is SLOWER (963 vs 991 trials/s) - NCCL's CTAs take SMs from the persistent one-CTA-per-SM
        # GEMM grid, whose statically strided tiles then need a second wave.  B2T_DP_EARLY=1 turns it on.
        if early_reduce is None:
            early_reduce = os.environ.get("B2T_DP_EARLY", "0") == "1"
        if early_reduce and rt is not None and hasattr(rt, "engine"):
            self.buckets.param_of_name = dict(model.named_parameters())
            rt.engine.grad_sink = self.buckets.sink
            rt.engine.grad_sync = self.buckets.sync_early

    def attach_optimizer(self, optimizer: torch.optim.Optimizer) -> None:
        optimizer.register_step_pre_hook(lambda opt, args, kwargs: self.finish())

    def finish(self) -> None:
        self.buckets.finish()

    def __call__(self, *a, **kw):
        return self.model(*a, **kw)
