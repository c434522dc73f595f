"""Fused multi-tensor Adam and global-norm gradient clipping on the library's own kernels
(csrc/train_aux.cu).

Drop-in for what the reference builds in `create_optimizer`
(src/experiments/b2t_gru_w2v_experiment.py:109-145, src/experiments/experiment.py:25-28):
`torch.optim.Adam(param_groups, lr=..., weight_decay=..., eps=...)` — L2 weight-decay form (not AdamW), two
learning-rate groups — and for `torch.nn.utils.clip_grad_norm_` (src/train/train_loop.py:72-75).  `FusedAdam` is a
`torch.optim.Optimizer`, so the reference's LambdaLR warm-up scheduler (src/util/warmup_scheduler.py), `zero_grad`,
`param_groups[i]["lr"]` edits and `state_dict()` keep working; parameters whose `.grad` is None are skipped exactly as
torch does (unused / LayerDrop-skipped parameters must not be moved by weight decay).

One launch updates every parameter of every group: a device table of (param, grad, exp_avg, exp_avg_sq) pointers plus
a (tensor, chunk) list, one CTA per 64K elements.  The chunk list is static over all parameters; the 48-byte-per-tensor
table is refreshed every step through pinned staging and a non-blocking copy (class _Plan), with n = 0 for the
parameters that have no gradient in that step.
"""
from __future__ import annotations

import ctypes
import math
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import c_int, check, lib, ptr, stream

MT_CHUNK = 65536          # B2T_MT_CHUNK
MAX_GROUPS = 32           # B2T_ADAM_MAX_GROUPS

_TENSOR_DT = np.dtype([("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("n", "<i8"), ("group", "<i4"), ("pad", "<i4")])


class AdamGroup(ctypes.Structure):
    """Mirror of b2t_adam_group (include/b2t_sm100.h)."""
    _fields_ = [("beta2", ctypes.c_float), ("eps", ctypes.c_float), ("weight_decay", ctypes.c_float),
                ("one_minus_beta1", ctypes.c_float), ("one_minus_beta2", ctypes.c_float), ("step_size", ctypes.c_float),
                ("bias_c2_sqrt", ctypes.c_float), ("pad_", ctypes.c_int32)]


class _Table:
    """Device-side (tensor table, chunk list) of one multi-tensor launch."""

    def __init__(self, rows: List[Tuple[int, int, int, int, int, int]], device):
        arr = np.zeros(len(rows), dtype=_TENSOR_DT)
        chunks = []
        for i, (p, g, m, v, n, grp) in enumerate(rows):
            arr[i] = (p, g, m, v, n, grp, 0)
            nch = (n + MT_CHUNK - 1) // MT_CHUNK
            chunks.append(np.stack([np.full(nch, i, dtype=np.int32), np.arange(nch, dtype=np.int32)], axis=1))
        ch = np.concatenate(chunks, axis=0) if chunks else np.zeros((0, 2), dtype=np.int32)
        self.n_chunks = int(ch.shape[0])
        self.tensors = torch.from_numpy(arr.view(np.uint8).reshape(-1).copy()).to(device)
        self.chunks = torch.from_numpy(np.ascontiguousarray(ch).reshape(-1)).to(device)


class _TableCache:
    def __init__(self, cap: int = 64):
        self.cap = cap
        self.d: Dict[tuple, _Table] = {}

    def get(self, sig: tuple, rows, device) -> _Table:
        t = self.d.get(sig)
        if t is None:
            if len(self.d) >= self.cap:
                self.d.pop(next(iter(self.d)))
            t = self.d[sig] = _Table(rows, device)
        return t


class _Plan:
    """Static part of the one-launch update of a fixed parameter list: the (tensor, chunk) list over ALL parameters and
    the device table; per step only the table (34 KB for the conformer's 716 tensors) is refreshed, through a ring of
    pinned staging buffers and a non-blocking copy.  A parameter without gradient in this step (LayerDrop-skipped
    layer, unused parameter) gets n = 0: its chunks find nothing to do, exactly torch's "skip parameters whose grad is
    None".  (The first version rebuilt and synchronously uploaded a table whenever the set of parameters with a
    gradient changed - with LayerDrop almost every step, 6 ms of host time in a step whose issue time is what bounds it.)"""
    RING = 8

    def __init__(self, params: List[torch.Tensor], device):
        self.params = params
        n = len(params)
        self.numel = np.array([p.numel() for p in params], dtype=np.int64)
        chunks = []
        for i, ne in enumerate(self.numel):
            nch = (int(ne) + MT_CHUNK - 1) // MT_CHUNK
            chunks.append(np.stack([np.full(nch, i, dtype=np.int32), np.arange(nch, dtype=np.int32)], axis=1))
        ch = np.concatenate(chunks, axis=0) if chunks else np.zeros((0, 2), dtype=np.int32)
        self.n_chunks = int(ch.shape[0])
        self.chunks = torch.from_numpy(np.ascontiguousarray(ch).reshape(-1)).to(device)
        nbytes = max(n, 1) * _TENSOR_DT.itemsize
        self.dev = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        self.host = [torch.zeros(nbytes, dtype=torch.uint8).pin_memory() for _ in range(self.RING)]
        self.views = [h.numpy().view(_TENSOR_DT) for h in self.host]
        self.events = [None] * self.RING
        self.slot = 0

    def upload(self, fill) -> None:
        """fill(view) writes this step's table into a pinned staging buffer; it is then copied to the device table"""
        i = self.slot
        self.slot = (i + 1) % self.RING
        if self.events[i] is not None:
            self.events[i].synchronize()        # the copy that last used this staging buffer has completed
        fill(self.views[i])
        self.dev.copy_(self.host[i], non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self.events[i] = ev


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (amsgrad=False, maximize=False, L2 weight decay) in one kernel launch per step."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        if lr < 0 or eps < 0 or weight_decay < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1):
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self._plan: Optional[_Plan] = None
        self._plan_key = None
        self.grad_scale: Optional[torch.Tensor] = None     # optional device scalar multiplied into every gradient
        self.launches = 0
        # data parallel (dist.DataParallel.attach_optimizer(lazy=True)): one launch per parameter group, in `group_order`,
        # each after group_wait(group index) has made the stream wait for that group's gradient buckets
        self.group_wait = None
        self.group_order = None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        _lib.ensure_device()
        # candidate parameters (fixed list -> static plan); the set that has a gradient varies per step
        cand: List[torch.Tensor] = []
        cand_group: List[int] = []
        for gi, group in enumerate(self.param_groups):
            for p in group["params"]:
                if p.requires_grad:
                    cand.append(p)
                    cand_group.append(gi)
        if not cand:
            return loss
        key = tuple(id(p) for p in cand)
        if self._plan is None or self._plan_key != key or self._plan.dev.device != cand[0].device:
            for p in cand:
                if p.dtype != torch.float32 or not p.is_cuda:
                    raise RuntimeError("FusedAdam: dense fp32 CUDA parameters only")
                if not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous")
            self._plan, self._plan_key = _Plan(cand, cand[0].device), key
        plan = self._plan
        n = len(cand)
        p_ptr = np.zeros(n, dtype=np.uint64)
        g_ptr = np.zeros(n, dtype=np.uint64)
        m_ptr = np.zeros(n, dtype=np.uint64)
        v_ptr = np.zeros(n, dtype=np.uint64)
        hyper_idx = np.zeros(n, dtype=np.int32)
        active = np.zeros(n, dtype=bool)
        hyper: Dict[tuple, int] = {}
        hlist: List[AdamGroup] = []
        keep = []
        state = self.state
        groups = self.param_groups
        for i, p in enumerate(cand):
            g = p.grad
            if g is None:
                continue
            if g.is_sparse or g.dtype != torch.float32:
                raise RuntimeError("FusedAdam: dense fp32 CUDA parameters only")
            if not g.is_contiguous():
                g = p.grad = g.contiguous()
            st = state[p]
            if not st:
                st["step"] = 0
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
            t = st["step"] = st["step"] + 1
            gi = cand_group[i]
            group = groups[gi]
            hk = (gi, t, float(group["lr"]))
            hi = hyper.get(hk)
            if hi is None:
                hi = hyper[hk] = len(hlist)
                b1, b2 = group["betas"]
                # derived scalars in double (Python floats), rounded to fp32 once — torch's _single_tensor_adam
                hlist.append(AdamGroup(b2, group["eps"], group["weight_decay"], 1.0 - b1, 1.0 - b2,
                                       group["lr"] / (1.0 - b1 ** t), math.sqrt(1.0 - b2 ** t), 0))
            p_ptr[i], g_ptr[i] = p.data_ptr(), g.data_ptr()
            m_ptr[i], v_ptr[i] = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
            hyper_idx[i] = hi
            active[i] = True
            keep.append(g)
        if not hlist:
            return loss
        cg_arr = np.asarray(cand_group, dtype=np.int32)

        def launch(mask) -> None:
            # more hyper-parameter classes than one launch carries (LayerDrop lets the per-parameter step counts drift
            # apart): one launch per slice of MAX_GROUPS classes, tensors of the other slices masked out (n = 0)
            for lo in range(0, len(hlist), MAX_GROUPS):
                sel = mask & (hyper_idx >= lo) & (hyper_idx < lo + MAX_GROUPS)
                if not sel.any():
                    continue

                def fill(view, sel=sel, lo=lo):
                    view["p"], view["g"], view["m"], view["v"] = p_ptr, g_ptr, m_ptr, v_ptr
                    view["n"] = np.where(sel, plan.numel, 0)
                    view["group"] = np.where(sel, hyper_idx - lo, 0)
                    view["pad"] = 0
                plan.upload(fill)
                harr = (AdamGroup * len(hlist[lo:lo + MAX_GROUPS]))(*hlist[lo:lo + MAX_GROUPS])
                check(lib().b2t_adam_multi(ptr(plan.dev), ptr(plan.chunks), c_int(plan.n_chunks), harr, c_int(len(harr)),
                                           ptr(self.grad_scale), stream()), "b2t_adam_multi")
                self.launches += 1

        if self.group_wait is None:
            launch(active)
        else:
            for gi in (self.group_order or range(len(self.param_groups))):
                self.group_wait(gi)          # the stream waits for this group's gradient buckets only
                launch(active & (cg_arr == gi))
        return loss


_CLIP_CACHE = _TableCache()
_CLIP_WS: Dict[int, torch.Tensor] = {}


@torch.no_grad()
def clip_grad_norm_(parameters: Iterable[torch.Tensor], max_norm: float, norm_type: float = 2.0,
                    error_if_nonfinite: bool = False, foreach=None) -> torch.Tensor:
    """torch.nn.utils.clip_grad_norm_ (L2) as two launches over all gradients, without a host synchronisation:
    sum of squares into a device double, then every gradient scaled by min(1, max_norm / (norm + 1e-6)).
    Returns the total norm as a 0-dim device tensor, like torch."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    grads = [p.grad for p in parameters if p.grad is not None]
    if float(norm_type) != 2.0:
        raise NotImplementedError("fused clip_grad_norm_: L2 norm only (what train_loop.py:72-75 uses)")
    if not grads:
        return torch.tensor(0.0)
    _lib.ensure_device()
    device = grads[0].device
    rows, sig = [], []
    for i, g in enumerate(grads):
        if g.dtype != torch.float32 or not g.is_cuda:
            raise RuntimeError("fused clip_grad_norm_: dense fp32 CUDA gradients only")
        if not g.is_contiguous():
            raise RuntimeError("fused clip_grad_norm_: gradients must be contiguous")
        rows.append((0, g.data_ptr(), 0, 0, g.numel(), 0))
        sig.append((g.data_ptr(), g.numel()))
    table = _CLIP_CACHE.get(("clip",) + tuple(sig), rows, device)
    ws = torch.zeros(1, dtype=torch.float64, device=device)
    total = torch.empty((), dtype=torch.float32, device=device)
    check(lib().b2t_grad_sumsq_multi(ptr(table.tensors), ptr(table.chunks), c_int(table.n_chunks), ptr(ws), stream()),
          "b2t_grad_sumsq_multi")
    check(lib().b2t_grad_clip_multi(ptr(table.tensors), ptr(table.chunks), c_int(table.n_chunks), ptr(ws),
                                    ctypes.c_float(float(max_norm)), ptr(total), stream()), "b2t_grad_clip_multi")
    if error_if_nonfinite and not bool(torch.isfinite(total)):
        raise RuntimeError("The total norm for gradients is non-finite, so it cannot be clipped")
    return total
