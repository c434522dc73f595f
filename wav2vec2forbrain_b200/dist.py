"""Data-parallel training across the GPUs of one box: one process per GPU, torch.distributed (NCCL over NVLink 5 /
NVSwitch), gradient buckets all-reduced IN PLACE on a side stream as soon as their last gradient exists, overlapped with
the rest of backward.

The reference is single-process (no torch.distributed anywhere); the semantics kept are those of
SURVEY.md §5.4 / §8(e): each rank runs the reference step on its shard of the batch, gradients are averaged (CTC
reduction is a batch mean, so equal shards reproduce the global-batch gradient), parameters that never receive a
gradient (unused inpLayer*, hidden_start, conformer pos_conv_embed, LayerDrop-skipped layers) are tolerated, LayerDrop
decisions are identical on all ranks, dropout masks differ per rank, BatchNorm statistics stay per rank.

Two bucket implementations:
  * `ArenaBuckets` (the product path): the model's gradients live in one flat buffer (flat.GradArena); a bucket is a
    contiguous range of it, so `all_reduce` runs on the gradients themselves — no flatten copy, no copy back.  The
    engine reports finished gradients after every encoder layer (`Engine.emit_grads`), so buckets leave while the
    layers below are still being back-propagated, not only when the whole autograd node returns.
  * `GradientBuckets` (any nn.Module): classic post-accumulate-grad hooks + flatten / all-reduce / copy back.
Both finish (wait for the collectives and make the averaged values visible to the compute stream) in a callback the
autograd engine runs when `loss.backward()` ends, so `clip_grad_norm_` (src/train/train_loop.py:72-75) and
`optimizer.step()` (:78) of the unmodified Trainer see reduced gradients without any extra call.
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
    if backend == "nccl" and os.environ.get("B2T_NCCL_HIPRI", "0") == "1":
        # opt-in: NCCL's kernels on a high-priority stream (the compute stream runs persistent one-CTA-per-SM GEMMs back to
        # back, so at normal priority a collective's CTAs may wait for SMs).  Measured: no effect on the intermittent
        # exposed all-reduce it was meant to cure, hence not the default.
        try:
            opts = dist.ProcessGroupNCCL.Options()
            opts.is_high_priority_stream = True
            dist.init_process_group(backend=backend, rank=rank, world_size=world, pg_options=opts)
            return
        except Exception:        # an older / different torch: plain initialisation
            if dist.is_initialized():
                return
    dist.init_process_group(backend=backend, rank=rank, world_size=world)


def tune_gc() -> None:
    """Host-side hygiene for a loop that issues ~1300 kernel launches per step from Python: everything alive now (model,
    optimizer state, cached plans) is moved out of the garbage collector's reach and full collections are made rare.  A
    generation-2 collection in ONE rank stalls that rank's launch stream for 100s of milliseconds, and every other rank
    waits for it inside the next collective (round 2: 2-GPU runs whose median step was 55 ms contained single steps of
    410 ms; B2T_GC_TUNE=0 switches this off)."""
    if os.environ.get("B2T_GC_TUNE", "1") != "1":
        return
    import gc
    gc.collect()
    gc.freeze()
    gc.set_threshold(200000, 100, 1000)


def shard_indices(n: int, rank: int, world: int) -> List[int]:
    """rank r takes indices r::world of the (identically shuffled) epoch order."""
    return list(range(rank, n, world))


def _queue_backward_callback(fn) -> bool:
    """Run `fn` when the current loss.backward() finishes (what DDP does); False outside a backward pass."""
    try:
        torch.autograd.Variable._execution_engine.queue_callback(fn)
        return True
    except RuntimeError:
        return False


def _side_streams():
    from . import ops as _ops
    return _ops.SIDE_STREAMS if torch.cuda.is_available() else ()


class _BucketBase:
    def __init__(self, process_group, overlap: bool):
        self.group = process_group
        self.world = dist.get_world_size(process_group)
        self.overlap = overlap and torch.cuda.is_available()
        self.stream = torch.cuda.Stream() if self.overlap else None
        self._avg = dist.get_backend(process_group) == "nccl"      # gloo has no ReduceOp.AVG
        self.enabled = True          # bench.py switches the collective off to measure the exposed communication
        self.bytes_per_step = 0
        self._bytes_now = 0
        self._cb_queued = False
        self._active = False

    def _all_reduce(self, flat: torch.Tensor):
        """mean over ranks, in place, asynchronously on the side stream when overlapping"""
        self._bytes_now += flat.numel() * flat.element_size()
        if not self.enabled:
            return None
        if self.overlap:
            self.stream.wait_stream(torch.cuda.current_stream())
            for st in _side_streams():      # weight gradients are produced on the engine's second stream
                self.stream.wait_stream(st)
            with torch.cuda.stream(self.stream):
                if self._avg:
                    return dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group, async_op=True)
                flat.div_(self.world)
                return dist.all_reduce(flat, group=self.group, async_op=True)
        for st in _side_streams():
            torch.cuda.current_stream().wait_stream(st)
        if self._avg:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group)
        else:
            flat.div_(self.world)
            dist.all_reduce(flat, group=self.group)
        return None

    def _mark_active(self) -> None:
        if not self._active:
            self._active = True
            self._cb_queued = _queue_backward_callback(self._backward_done)

    def _backward_done(self) -> None:
        self._cb_queued = False
        self.finish()

    def finish(self) -> None:       # pragma: no cover - interface
        raise NotImplementedError


class ArenaBuckets(_BucketBase):
    """In-place bucketed all-reduce over a flat gradient arena (see module docstring)."""

    def __init__(self, arena, bucket_bytes: int = 48 << 20, process_group=None, overlap: bool = True,
                 boundaries=("brain_encoder.", "w2v_encoder."), early: bool = True):
        super().__init__(process_group, overlap)
        self.arena = arena
        self.early = early
        self.boundaries = tuple(boundaries)
        raw = arena.make_buckets(bucket_bytes, self.boundaries)
        self.buckets = [arena.buf[s:e] for s, e, _ in raw]
        self.bucket_names = [names for _, _, names in raw]
        self.bucket_tag = [self._tag(names[0]) for names in self.bucket_names]
        self.bucket_of: Dict[str, int] = {n: bi for bi, names in enumerate(self.bucket_names) for n in names}
        self._pending: List[int] = []
        self._seen: List[int] = []
        self._launched: List[bool] = []
        self._work: list = []
        self._reset()

    def _tag(self, name: str) -> str:
        for b in self.boundaries:
            if name.startswith(b):
                return b
        return ""

    def _reset(self) -> None:
        self._pending = [len(n) for n in self.bucket_names]
        self._seen = [0] * len(self.buckets)
        self._launched = [False] * len(self.buckets)
        self._work = []
        self._active = False
        self.bytes_per_step, self._bytes_now = self._bytes_now or self.bytes_per_step, 0

    # ---- called by the engine -------------------------------------------------------------------------------
    def sink(self, named) -> None:
        """named: [(parameter name, finished gradient = its arena slot)].  A bucket whose every gradient has arrived
        starts its all-reduce right away, behind everything issued so far on the compute stream."""
        if self._ended:               # lazy mode, and no optimizer step since the last backward pass: finish that one now
            self.finish(force=True)
        self._mark_active()
        for name, _ in named:
            bi = self.bucket_of.get(name)
            if bi is None:
                continue
            self._pending[bi] -= 1
            self._seen[bi] += 1
            if self._pending[bi] < 0:
                raise RuntimeError(f"gradient of {name} arrived twice before finish(): one backward pass per "
                                   "optimizer step (gradient accumulation is not supported by ArenaBuckets)")
            if self.early and self._pending[bi] == 0 and not self._launched[bi]:
                self._launch(bi)

    def node_end(self, prefix: str = "") -> None:
        """End of an autograd node: everything of that node which received at least one gradient leaves now (LayerDrop
        or unused parameters keep some buckets from ever completing; their slots hold zeros on every rank, and the
        launch order is the same on every rank because LayerDrop decisions are shared)."""
        for bi in range(len(self.buckets) - 1, -1, -1):       # backward fills the arena from its end
            if not self._launched[bi] and self._seen[bi] > 0 and (not prefix or self.bucket_tag[bi] == prefix):
                self._launch(bi)

    def _launch(self, bi: int) -> None:
        self._launched[bi] = True
        work = self._all_reduce(self.buckets[bi])
        self._work.append((bi, work))

    # ---- end of backward ------------------------------------------------------------------------------------
    lazy = False      # DataParallel.attach_optimizer(..., lazy=True): see wait_prefix
    _ended = False

    def finish(self, force: bool = False) -> None:
        """Everything that has not left yet is launched; the compute stream then waits for all of it - unless `lazy`, where
        the waiting is left to wait_prefix() calls of the optimizer (and to the finish(force=True) after its step)."""
        if not self._active:
            return
        self.node_end("")
        if self.lazy and not force:
            self._ended = True        # a backward pass that starts before the optimizer ran must wait and reset first (sink)
            return
        self._ended = False
        if self.overlap:
            for _, w in self._work:
                if w is not None:
                    w.wait()          # the current (compute) stream waits for that collective
        self._reset()

    def wait_prefix(self, prefix: str) -> None:
        """The compute stream waits for the buckets of one autograd node only ("brain_encoder." / "w2v_encoder."): the
        optimizer updates the encoder's parameters while the last buckets - the brain encoder's, whose backward runs
        last - are still being reduced."""
        if not self._active:
            return
        self.node_end("")
        if self.overlap:
            rest = []
            for bi, w in self._work:
                if self.bucket_tag[bi] == prefix:
                    if w is not None:
                        w.wait()
                else:
                    rest.append((bi, w))
            self._work = rest


class GradientBuckets(_BucketBase):
    """Bucketed asynchronous all-reduce(mean) of parameter gradients of ANY module (hook driven).

    Buckets are formed over the parameters that may receive a gradient, in reverse registration order (≈ reverse
    execution order), capped at `bucket_bytes`.  A bucket is launched from the post-accumulate-grad hook of its last
    arriving parameter; parameters whose gradient never arrives (LayerDrop, unused) are left out at `finish()`, which
    runs when loss.backward() ends.  Gradients are flattened into a bucket copy and copied back (the arena path avoids
    both)."""

    def __init__(self, params: List[nn.Parameter], bucket_bytes: int = 48 << 20, process_group=None, overlap: bool = True):
        super().__init__(process_group, overlap)
        self.params = [p for p in params if p.requires_grad]
        self.bucket_bytes = bucket_bytes
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
        self.bytes_per_step, self._bytes_now = self._bytes_now or self.bytes_per_step, 0

    def _on_grad(self, p: nn.Parameter) -> None:
        self._mark_active()
        bi = self._bucket_of[id(p)]
        self._pending[bi] -= 1
        if self._pending[bi] < 0 or self._launched[bi]:
            raise RuntimeError("a gradient arrived for a bucket that was already reduced: call finish() (it runs by itself "
                               "at the end of loss.backward()) before the next backward pass; gradient accumulation "
                               "over several backward passes is not supported")
        if self._pending[bi] == 0:
            self._launch(bi)

    def _launch(self, bi: int) -> None:
        self._launched[bi] = True
        grads = [p.grad for p in self.buckets[bi] if p.grad is not None]
        if not grads:
            return
        if self.overlap:
            self.stream.wait_stream(torch.cuda.current_stream())
            for st in _side_streams():      # weight gradients are produced on the engine's second stream
                self.stream.wait_stream(st)
            with torch.cuda.stream(self.stream):
                flat = torch.cat([g.reshape(-1) for g in grads])
        else:
            flat = torch.cat([g.reshape(-1) for g in grads])
        work = self._all_reduce(flat)
        self._work.append((work, flat, grads))

    def finish(self) -> None:
        """Launch what has not been launched (identical bucket composition on every rank is guaranteed by
        rank-consistent LayerDrop), wait, scatter the averaged values back into .grad."""
        if not self._active:
            return   # nothing arrived since the last finish(): second call of the same step
        for bi in range(len(self.buckets)):
            if not self._launched[bi]:
                self._launch(bi)
        for work, flat, grads in self._work:
            ctx = torch.cuda.stream(self.stream) if self.overlap else _null()
            with ctx:
                if work is not None:
                    work.wait()
                if self.enabled:
                    torch._foreach_copy_(grads, [v.view_as(g) for v, g in zip(flat.split([g.numel() for g in grads]), grads)])
        if self.overlap:
            torch.cuda.current_stream().wait_stream(self.stream)
        self.reset()

    def remove(self) -> None:
        for h in self._hooks:
            h.remove()


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class DataParallel:
    """Wraps a model for the unmodified reference Trainer loop (train_loop.py:44-78): broadcasts the initial
    weights / buffers from rank 0, installs the gradient buckets, and finishes the reduction when loss.backward()
    ends (so clip_grad_norm_ and optimizer.step see averaged gradients)."""

    def __init__(self, model: nn.Module, bucket_bytes: int = 48 << 20, process_group=None, overlap: bool = True,
                 early_reduce: Optional[bool] = None):
        self.model = model
        self.group = process_group
        with torch.no_grad():
            for t in list(model.parameters()) + list(model.buffers()):
                dist.broadcast(t, src=0, group=process_group)
        rt = getattr(model, "runtime", None)
        if rt is not None:
            rt.rank = dist.get_rank(process_group)   # per-rank dropout masks; LayerDrop generator stays shared
            # the engine caches bf16 operand copies per parameter version: whatever ran before the broadcast is stale
            rt.engine.weights._cache.clear()
        # (B2T_GEMM_CLC=1 switches the GEMMs to the dynamic cluster-launch-control schedule and B2T_NCCL_HIPRI=1 puts NCCL on
        # a high-priority stream; both were tried against an intermittent 20-55 ms of exposed all-reduce per step seen at
        # 2 and 8 GPUs and neither removed it, see DESIGN.md section 6 - they stay opt-in.)
        if early_reduce is None:
            early_reduce = os.environ.get("B2T_DP_EARLY", "1") == "1"
        arena = rt.grad_arena() if (rt is not None and hasattr(rt, "grad_arena")) else None
        if arena is not None:
            self.buckets = ArenaBuckets(arena, bucket_bytes, process_group, overlap, early=early_reduce)
            rt.engine.grad_sink = self.buckets.sink
            rt.engine.grad_sync = self.buckets.node_end
        else:
            self.buckets = GradientBuckets(list(model.parameters()), bucket_bytes, process_group, overlap)

    def attach_optimizer(self, optimizer: torch.optim.Optimizer, lazy: bool = False) -> None:
        """Belt and braces: finish() is idempotent, and backward passes driven outside autograd's engine (where the
        end-of-backward callback cannot be queued) are still reduced before the step.

        lazy=True (only with this package's FusedAdam and the arena buckets, and only when NOTHING reads gradients between
        backward() and step() - no clip_grad_norm_): the end of backward no longer makes the compute stream wait for
        every bucket; the optimizer updates one parameter group per launch and waits, per group, for that group's
        buckets only.  The brain encoder's gradients are produced last, so their all-reduce then runs under the update
        of the (much larger) wav2vec2 encoder instead of in front of the optimizer."""
        lazy = lazy and isinstance(self.buckets, ArenaBuckets) and hasattr(optimizer, "group_wait")
        if not lazy:
            optimizer.register_step_pre_hook(lambda opt, args, kwargs: self.finish())
            return
        name_of = {id(p): n for n, p in self.model.named_parameters()}
        prefixes = []
        for g in optimizer.param_groups:
            tags = {self.buckets._tag(name_of.get(id(p), "")) for p in g["params"] if p.requires_grad}
            prefixes.append(tags.pop() if len(tags) == 1 else None)
        if any(t is None or t == "" for t in prefixes):
            optimizer.register_step_pre_hook(lambda opt, args, kwargs: self.finish())
            return
        self.buckets.lazy = True
        # groups whose gradients are complete first go first: the arena is filled from its end, the brain encoder last
        order = sorted(range(len(prefixes)), key=lambda gi: 0 if prefixes[gi] == "w2v_encoder." else 1)
        optimizer.group_order = order
        optimizer.group_wait = lambda gi: self.buckets.wait_prefix(prefixes[gi])
        optimizer.register_step_post_hook(lambda opt, args, kwargs: self.buckets.finish(force=True))

    def finish(self) -> None:
        """explicit synchronisation point (e.g. before reading gradients in lazy mode): waits for every bucket"""
        if isinstance(self.buckets, ArenaBuckets):
            self.buckets.finish(force=True)
        else:
            self.buckets.finish()

    def __call__(self, *a, **kw):
        return self.model(*a, **kw)
