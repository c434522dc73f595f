"""Thin Python wrappers over the C ABI (include/b2t_sm100.h).

Every function allocates its outputs with torch (device memory + current stream are the only
things torch provides here) and enqueues exactly the kernels of libb2t_sm100.so.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from ._lib import GemmArgs, c_int, c_int64, c_float, check, lib, ptr, stream

Tensor = torch.Tensor
ACT = {None: 0, "none": 0, "gelu": 1, "silu": 2, "swish": 2, "softsign": 3}

# number of kernels of this library launched so far (bench.py reports it as gpu_launches)
LAUNCHES = 0


def _count(n=1):
    global LAUNCHES
    LAUNCHES += n


@dataclass
class Operand:
    """A bf16 (hi[, lo]) GEMM operand described as a 3-D strided view, d0 contiguous."""
    hi: Tensor
    lo: Optional[Tensor]
    d0: int
    d1: int
    d2: int
    s1: int
    s2: int
    mn_major: bool = False

    def as_mn(self) -> "Operand":
        return Operand(self.hi, self.lo, self.d0, self.d1, self.d2, self.s1, self.s2, True)


def kmajor(hi: Tensor, lo: Optional[Tensor] = None) -> Operand:
    """[rows, K] row-major (optionally [batch, rows, K]) used with K as the contraction."""
    if hi.dim() == 2:
        r, k = hi.shape
        return Operand(hi, lo, k, r, 1, hi.stride(0), r * hi.stride(0))
    b, r, k = hi.shape
    return Operand(hi, lo, k, r, b, hi.stride(1), hi.stride(0))


def mnmajor(hi: Tensor, lo: Optional[Tensor] = None) -> Operand:
    """[Krows, MN] row-major (optionally [batch, Krows, MN]) contracted over its ROWS."""
    op = kmajor(hi, lo)
    op.mn_major = True
    return op


def unfold_view(hi: Tensor, lo: Optional[Tensor], klen: int, stride: int, mn_major: bool = False) -> Operand:
    """[B, T, C] seen as [B, T', klen*C] with row j starting at frame stride*j — the im2col of
    nn.Unfold((klen,1), stride) (b2p2t_model.py:162-167) without materialising it; feature order
    is k*C + c."""
    b, t, c = hi.shape
    tp = (t - klen) // stride + 1
    return Operand(hi, lo, klen * c, tp, b, stride * c, t * c, mn_major)


def split(x: Tensor, three: bool = True, transpose: bool = False):
    """fp32 [rows, cols] -> bf16 (hi, lo) (lo None in 1-term mode), optionally transposed."""
    _lib.ensure_device()
    assert x.dtype == torch.float32 and x.is_cuda
    x2 = x.reshape(-1, x.shape[-1])
    assert x2.stride(1) == 1
    rows, cols = x2.shape
    shape = (cols, rows) if transpose else tuple(x.shape)
    hi = torch.empty(shape, dtype=torch.bfloat16, device=x.device)
    lo = torch.empty(shape, dtype=torch.bfloat16, device=x.device) if three else None
    check(lib().b2t_split_bf16(ptr(x2), ptr(hi), ptr(lo), c_int64(rows), c_int64(cols), c_int64(x2.stride(0)),
                               c_int(1 if transpose else 0), stream()), "b2t_split_bf16")
    _count()
    return hi, lo


def permute_unfold_cols(w: Tensor, C: int, KL: int, three: bool = True):
    """W_ih_l0 [rows, C*KL] (column c*KL+k) -> bf16 hi/lo with column k*C+c."""
    rows = w.shape[0]
    hi = torch.empty_like(w, dtype=torch.bfloat16)
    lo = torch.empty_like(w, dtype=torch.bfloat16) if three else None
    check(lib().b2t_permute_unfold_cols(ptr(w), ptr(hi), ptr(lo), ptr(None), c_int64(rows), c_int(C), c_int(KL), c_int(0),
                                        stream()), "b2t_permute_unfold_cols")
    _count()
    return hi, lo


def unpermute_unfold_cols(g: Tensor, C: int, KL: int) -> Tensor:
    """gradient in k*C+c column order -> reference order c*KL+k (fp32)."""
    out = torch.empty_like(g)
    check(lib().b2t_permute_unfold_cols(ptr(g), ptr(None), ptr(None), ptr(out), c_int64(g.shape[0]), c_int(C), c_int(KL),
                                        c_int(1), stream()), "b2t_permute_unfold_cols")
    _count()
    return out


def gemm(a: Operand, b: Operand, *, m_per_batch: int, n: int, k: int, batches: int = 1,
         out_f32: Optional[Tensor] = None, out_pre: Optional[Tensor] = None,
         out_hi: Optional[Tensor] = None, out_lo: Optional[Tensor] = None,
         ldc: Optional[int] = None, c_batch_stride: Optional[int] = None,
         bias: Optional[Tensor] = None, residual: Optional[Tensor] = None, aux: Optional[Tensor] = None,
         colsum: Optional[Tensor] = None, act=None, aux_mode=None, atomic: bool = False, split_k: int = 1,
         alpha: float = 1.0, k_spans_batches: bool = False, a_batched: bool = False, b_batched: bool = False,
         b_batch_idx: Optional[Tensor] = None, c_batch_idx: Optional[Tensor] = None, block_n: int = 0) -> None:
    """D = epilogue(alpha * A . B^T) on the tcgen05 GEMM (see b2t_gemm in include/b2t_sm100.h)."""
    _lib.ensure_device()
    g = GemmArgs()
    g.a_hi, g.a_lo, g.b_hi, g.b_lo = ptr(a.hi), ptr(a.lo), ptr(b.hi), ptr(b.lo)
    g.a_d0, g.a_d1, g.a_d2, g.a_s1, g.a_s2 = a.d0, a.d1, a.d2, a.s1, a.s2
    g.b_d0, g.b_d1, g.b_d2, g.b_s1, g.b_s2 = b.d0, b.d1, b.d2, b.s1, b.s2
    g.a_mn_major, g.b_mn_major = int(a.mn_major), int(b.mn_major)
    g.k_spans_batches = int(k_spans_batches)
    g.m_per_batch, g.batches, g.n, g.k = m_per_batch, batches, n, k
    g.a_batched, g.b_batched = int(a_batched), int(b_batched)
    g.b_batch_idx, g.c_batch_idx = ptr(b_batch_idx), ptr(c_batch_idx)
    g.out_f32, g.out_pre, g.out_hi, g.out_lo = ptr(out_f32), ptr(out_pre), ptr(out_hi), ptr(out_lo)
    g.ldc = n if ldc is None else ldc
    g.c_batch_stride = m_per_batch * g.ldc if c_batch_stride is None else c_batch_stride
    g.bias, g.residual, g.aux, g.colsum = ptr(bias), ptr(residual), ptr(aux), ptr(colsum)
    g.act = ACT[act]
    g.aux_mode = ACT[aux_mode]
    g.atomic = int(atomic)
    g.split_k = split_k
    g.alpha = alpha
    g.block_n = block_n
    check(lib().b2t_gemm(ctypes.byref(g), stream()), "b2t_gemm")
    _count()


def gemm_simt(a: Tensor, b: Tensor, bias: Optional[Tensor] = None) -> Tensor:
    """fp32 debug GEMM: a [m,k] . b [n,k]^T (any strides)."""
    _lib.ensure_device()
    m, k = a.shape
    n = b.shape[0]
    c = torch.empty(m, n, dtype=torch.float32, device=a.device)
    check(lib().b2t_gemm_simt(ptr(a), ptr(b), ptr(c), ptr(bias), c_int(m), c_int(n), c_int(k), c_int64(a.stride(0)),
                              c_int64(a.stride(1)), c_int64(b.stride(0)), c_int64(b.stride(1)), c_int64(n), stream()),
          "b2t_gemm_simt")
    _count()
    return c
