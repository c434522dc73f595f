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
a (tensor, chunk) list, one CTA per 64K elements.  Tables are cached per (pointer set, hyper-parameter class)
signature, so with the flat gradient arena (stable gradient addresses) the host does no per-step table work.
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


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (amsgrad=False, maximize=False, L2 weight decay) in one kernel launch per step."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        if lr < 0 or eps < 0 or weight_decay < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1):
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self._cache = _TableCache()
        self.grad_scale: Optional[torch.Tensor] = None     # optional device scalar multiplied into every gradient
        self.launches = 0

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        _lib.ensure_device()
        rows, sig = [], []
        hyper: Dict[tuple, int] = {}
        hlist: List[AdamGroup] = []
        device = None
        keep = []
        for gi, group in enumerate(self.param_groups):
            b1, b2 = group["betas"]
            for p in group["params"]:
                g = p.grad
                if g is None:
                    continue
                if g.is_sparse or p.dtype != torch.float32 or g.dtype != torch.float32 or not p.is_cuda:
                    raise RuntimeError("FusedAdam: dense fp32 CUDA parameters only")
                if not p.is_contiguous():
                    raise RuntimeError("FusedAdam: parameters must be contiguous")
                if not g.is_contiguous():
                    g = p.grad = g.contiguous()
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] += 1
                t = st["step"]
                hk = (gi, t, float(group["lr"]))
                hi = hyper.get(hk)
                if hi is None:
                    hi = hyper[hk] = len(hlist)
                    # derived scalars in double (Python floats), rounded to fp32 once — torch's _single_tensor_adam
                    hlist.append(AdamGroup(b2, group["eps"], group["weight_decay"], 1.0 - b1, 1.0 - b2,
                                           group["lr"] / (1.0 - b1 ** t), math.sqrt(1.0 - b2 ** t), 0))
                device = p.device
                rows.append((p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel(), hi))
                sig.append((rows[-1][0], rows[-1][1], hi))
                keep.append(g)
        if not rows:
            return loss
        # more hyper-parameter classes than one launch carries (LayerDrop lets the per-parameter step counts drift
        # apart): one launch per slice of MAX_GROUPS classes
        for lo in range(0, len(hlist), MAX_GROUPS):
            sel = [(r, s) for r, s in zip(rows, sig) if lo <= r[5] < lo + MAX_GROUPS]
            r2 = [r[:5] + (r[5] - lo,) for r, _ in sel]
            table = self._cache.get(("adam", lo) + tuple(s for _, s in sel), r2, device)
            harr = (AdamGroup * len(hlist[lo:lo + MAX_GROUPS]))(*hlist[lo:lo + MAX_GROUPS])
            check(lib().b2t_adam_multi(ptr(table.tensors), ptr(table.chunks), c_int(table.n_chunks), harr, c_int(len(harr)),
                                       ptr(self.grad_scale), stream()), "b2t_adam_multi")
            self.launches += 1
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
