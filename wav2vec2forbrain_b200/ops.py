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
ACT = {None: 0, "none": 0, "gelu": 1, "silu": 2, "swish": 2, "softsign": 3, "mul": 4}

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


def unpermute_unfold_cols(g: Tensor, C: int, KL: int, out: Optional[Tensor] = None) -> Tensor:
    """gradient in k*C+c column order -> reference order c*KL+k (fp32)."""
    out = torch.empty_like(g) if out is None else out
    assert out.is_contiguous() and out.shape == g.shape
    check(lib().b2t_permute_unfold_cols(ptr(g), ptr(None), ptr(None), ptr(out), c_int64(g.shape[0]), c_int(C), c_int(KL),
                                        c_int(1), stream()), "b2t_permute_unfold_cols")
    _count()
    return out


class _ZeroArena:
    """One memset per backward pass instead of ~500 small fill launches: inside `zero_arena(key, device)` every
    ops.zeros() is a slice of a single zero-filled fp32 buffer whose size was learned from the previous pass of
    the same key.  A fresh buffer is allocated per pass (slices may become .grad tensors and must not be reused)."""

    def __init__(self):
        self.need = {}
        self.buf = None
        self.off = 0
        self.used = 0
        self.key = None

    def begin(self, key, device):
        self.key, self.off, self.used = key, 0, 0
        n = self.need.get(key, 0)
        self.buf = torch.zeros(n, dtype=torch.float32, device=device) if n else None

    def end(self):
        if self.key is not None:
            self.need[self.key] = self.used
        self.buf, self.key = None, None

    def take(self, shape, device):
        n = 1
        for d in shape:
            n *= int(d)
        n64 = (n + 63) // 64 * 64      # 256-byte aligned slices (vector kernels want 16)
        if self.key is None:
            return torch.zeros(shape, dtype=torch.float32, device=device)
        self.used += n64
        if self.buf is not None and self.off + n64 <= self.buf.numel() and self.buf.device == torch.device(device):
            v = self.buf[self.off:self.off + n].view(shape)
            self.off += n64
            return v
        return torch.zeros(shape, dtype=torch.float32, device=device)


_ARENA = _ZeroArena()


class zero_arena:
    def __init__(self, key, device):
        self.key, self.device = key, device

    def __enter__(self):
        _ARENA.begin(self.key, self.device)
        return self

    def __exit__(self, *exc):
        _ARENA.end()
        return False


def zeros(*shape, device) -> Tensor:
    """fp32 zeros (from the active zero arena if there is one)."""
    if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
        shape = tuple(shape[0])
    return _ARENA.take(tuple(shape), device)


SIDE_STREAMS = []        # streams other than the launching one that write gradients (engine weight-gradient stream): dist waits on them
STREAM_K = False         # stream-K schedule switch: off by default (round-1 measurements: slower than whole-tile waves at K <= 4096)
_SK_WS = {}


def _sk_workspace(device, raw=None) -> Tensor:
    """Per (device, stream) stream-K scratch, zeroed once (b2t_gemm leaves its flags at zero)."""
    key = (device.index, _lib.raw_stream() if raw is None else raw)
    ws = _SK_WS.get(key)
    if ws is None:
        ws = torch.zeros(int(lib().b2t_gemm_workspace_bytes()), dtype=torch.uint8, device=device)
        _SK_WS[key] = ws
    return ws


# b2t_gemm_args is filled with ONE struct.pack_into call into a reused buffer: ~45 ctypes attribute assignments and a
# dozen c_void_p objects per call were ~10 us of the ~20 us host cost of a GEMM launch, in a training step whose issue
# time (1300 launches) had become what bounds it.  The format string is derived from the ctypes mirror, whose layout
# tests/test_abi_cpu.py checks against the C compiler.
import struct as _struct

_GEMM_FIELDS = [f[0] for f in GemmArgs._fields_]
_CT2FMT = {ctypes.c_void_p: "Q", ctypes.c_int64: "q", ctypes.c_int32: "i", ctypes.c_float: "f", ctypes.c_uint64: "Q"}
_GEMM_FMT = "".join(_CT2FMT[f[1]] for f in GemmArgs._fields_)
_GEMM_STRUCT = _struct.Struct("@" + _GEMM_FMT)
# native alignment gives the packed values the offsets of the ctypes Structure (= the C compiler's)
assert _GEMM_STRUCT.size <= ctypes.sizeof(GemmArgs)
assert all(getattr(GemmArgs, name).offset == _struct.calcsize("@" + _GEMM_FMT[:i] + "0" + _GEMM_FMT[i])
           for i, name in enumerate(_GEMM_FIELDS))
assert _GEMM_FIELDS == [
    "a_hi", "a_lo", "b_hi", "b_lo", "a_d0", "a_d1", "a_d2", "a_s1", "a_s2", "b_d0", "b_d1", "b_d2", "b_s1", "b_s2",
    "a_mn_major", "b_mn_major", "k_spans_batches", "m_per_batch", "batches", "n", "k", "a_batched", "b_batched",
    "b_batch_idx", "c_batch_idx", "out_f32", "out_pre", "out_hi", "out_lo", "ldc", "c_batch_stride", "bias", "residual",
    "aux", "colsum", "act", "aux_mode", "atomic", "split_k", "alpha", "block_n", "drop_p", "drop_seed",
    "bias_batch_stride", "out_scale", "sk_workspace", "sk_workspace_bytes", "stream_k", "cta_group", "pre_mode",
    "reserved_"]   # order gemm() packs in
_GEMM_BUF = ctypes.create_string_buffer(ctypes.sizeof(GemmArgs))
_GEMM_BUF_PTR = ctypes.c_void_p(ctypes.addressof(_GEMM_BUF))
_B2T_GEMM = None


def _dp(t) -> int:
    return 0 if t is None else t.data_ptr()


def gemm(a: Operand, b: Operand, *, m_per_batch: int, n: int, k: int, batches: int = 1,
         out_f32: Optional[Tensor] = None, out_pre: Optional[Tensor] = None,
         out_hi: Optional[Tensor] = None, out_lo: Optional[Tensor] = None,
         ldc: Optional[int] = None, c_batch_stride: Optional[int] = None,
         bias: Optional[Tensor] = None, residual: Optional[Tensor] = None, aux: Optional[Tensor] = None,
         colsum: Optional[Tensor] = None, act=None, aux_mode=None, atomic: bool = False, split_k: int = 1,
         alpha: float = 1.0, k_spans_batches: bool = False, a_batched: bool = False, b_batched: bool = False,
         b_batch_idx: Optional[Tensor] = None, c_batch_idx: Optional[Tensor] = None, block_n: int = 0,
         drop_p: float = 0.0, drop_seed: int = 0, bias_batch_stride: int = 0, out_scale: float = 1.0,
         stream_k: int = 0, cta_group: int = 0, on_stream: Optional[int] = None, pre_mode: int = 0) -> None:
    """D = epilogue(alpha * A . B^T) on the tcgen05 GEMM (see b2t_gemm in include/b2t_sm100.h).
    on_stream: raw cudaStream_t to launch on instead of torch's current stream (the engine's weight-gradient stream)."""
    global _B2T_GEMM
    _lib.ensure_device()
    if _B2T_GEMM is None:
        _B2T_GEMM = lib().b2t_gemm
    sk_ws, sk_bytes, sk = 0, 0, -1
    if STREAM_K or stream_k > 0:
        ws = _sk_workspace(a.hi.device, on_stream)
        sk_ws, sk_bytes, sk = ws.data_ptr(), ws.numel(), stream_k
    ldc_ = n if ldc is None else ldc
    # positional, in the field order of b2t_gemm_args (_GEMM_FIELDS; asserted at import below the definition)
    _GEMM_STRUCT.pack_into(
        _GEMM_BUF, 0,
        a.hi.data_ptr(), _dp(a.lo), b.hi.data_ptr(), _dp(b.lo),
        a.d0, a.d1, a.d2, a.s1, a.s2, b.d0, b.d1, b.d2, b.s1, b.s2,
        int(a.mn_major), int(b.mn_major), int(k_spans_batches), m_per_batch, batches, n, k,
        int(a_batched), int(b_batched), _dp(b_batch_idx), _dp(c_batch_idx),
        _dp(out_f32), _dp(out_pre), _dp(out_hi), _dp(out_lo),
        ldc_, m_per_batch * ldc_ if c_batch_stride is None else c_batch_stride,
        _dp(bias), _dp(residual), _dp(aux), _dp(colsum),
        ACT[act], ACT[aux_mode], int(atomic), split_k, alpha, block_n, drop_p, drop_seed & 0xFFFFFFFFFFFFFFFF,
        bias_batch_stride, out_scale, sk_ws, sk_bytes, sk, cta_group, pre_mode, 0)
    rc = _B2T_GEMM(_GEMM_BUF_PTR, ctypes.c_void_p(_lib.raw_stream() if on_stream is None else on_stream))
    if rc != 0:
        check(rc, "b2t_gemm")
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


# ---------------------------------------------------------------------------------------------
# non-GEMM kernels
# ---------------------------------------------------------------------------------------------
c_uint64 = ctypes.c_uint64


def _bf16_like(x: Tensor, enable: bool = True):
    return torch.empty(x.shape, dtype=torch.bfloat16, device=x.device) if enable else None


def smooth_fwd(x: Tensor, taps: Tensor, pad_left: int, three: bool = True, want_f32: bool = False,
               white_sd: float = 0.0, offset_sd: float = 0.0, seed: int = 0):
    """x [B,T,C] -> smoothed (y_f32 | None, y_hi, y_lo | None)."""
    _lib.ensure_device()
    B, T, C = x.shape
    taps2 = taps.reshape(C, -1).contiguous()
    y = torch.empty_like(x) if want_f32 else None
    hi, lo = _bf16_like(x), _bf16_like(x, three)
    check(lib().b2t_smooth_fwd(ptr(x), ptr(taps2), ptr(y), ptr(hi), ptr(lo), c_int(B), c_int(T), c_int(C),
                               c_int(taps2.shape[1]), c_int(pad_left), c_float(white_sd), c_float(offset_sd),
                               c_uint64(seed & 0xFFFFFFFFFFFFFFFF), stream()), "b2t_smooth_fwd")
    _count()
    return y, hi, lo


def softsign_bwd(ds: Tensor, z: Tensor, day: Tensor, dbias: Tensor, three: bool = True):
    B, T, C = ds.shape
    hi, lo = _bf16_like(ds), _bf16_like(ds, three)
    check(lib().b2t_softsign_bwd(ptr(ds), ptr(z), ptr(day), ptr(hi), ptr(lo), ptr(dbias), c_int(B), c_int(T), c_int(C),
                                 stream()), "b2t_softsign_bwd")
    _count()
    return hi, lo


def layernorm_fwd(x: Tensor, gamma: Tensor, beta: Tensor, eps: float, three: bool = True, want_f32: bool = True,
                  want_split: bool = True, drop_p: float = 0.0, seed: int = 0):
    """x [rows, n] -> (y | None, y_hi | None, y_lo | None, mean, rstd)."""
    _lib.ensure_device()
    rows, n = x.shape
    y = torch.empty_like(x) if want_f32 else None
    hi = _bf16_like(x, want_split)
    lo = _bf16_like(x, want_split and three)
    mean = torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    check(lib().b2t_layernorm_fwd(ptr(x), ptr(gamma), ptr(beta), c_float(eps), ptr(y), ptr(hi), ptr(lo), ptr(mean),
                                  ptr(rstd), c_int(rows), c_int(n), c_float(drop_p), c_uint64(seed), stream()),
          "b2t_layernorm_fwd")
    _count()
    return y, hi, lo, mean, rstd


def layernorm_bwd(dy: Tensor, x: Tensor, gamma: Tensor, mean: Tensor, rstd: Tensor, dgamma: Tensor, dbeta: Tensor,
                  dcol: Optional[Tensor] = None, three: bool = True, want_f32: bool = True, want_split: bool = True,
                  out_drop_p: float = 0.0, out_seed: int = 0, br_drop_p: float = 0.0, br_seed: int = 0,
                  dres: Optional[Tensor] = None):
    """-> (dx | None, dx_hi | None, dx_lo | None); dgamma/dbeta/dcol are accumulated into."""
    rows, n = x.shape
    dx = torch.empty_like(x) if want_f32 else None
    hi = _bf16_like(x, want_split)
    lo = _bf16_like(x, want_split and three)
    check(lib().b2t_layernorm_bwd(ptr(dy), ptr(x), ptr(gamma), ptr(mean), ptr(rstd), ptr(dx), ptr(hi), ptr(lo),
                                  ptr(dgamma), ptr(dbeta), ptr(dcol), c_int(rows), c_int(n), c_float(out_drop_p),
                                  c_uint64(out_seed), c_float(br_drop_p), c_uint64(br_seed), ptr(dres), stream()),
          "b2t_layernorm_bwd")
    _count()
    return dx, hi, lo


def dropout(x: Tensor, p: float, seed: int, three: bool = True, want_f32: bool = True, want_split: bool = False):
    """y = dropout(x) (p == 0: identity / plain split) -> (y | None, hi | None, lo | None)."""
    _lib.ensure_device()
    assert x.is_contiguous()
    y = torch.empty_like(x) if want_f32 else None
    hi = _bf16_like(x, want_split)
    lo = _bf16_like(x, want_split and three)
    check(lib().b2t_dropout(ptr(x), ptr(y), ptr(hi), ptr(lo), c_int64(x.numel()), c_float(p), c_uint64(seed), stream()),
          "b2t_dropout")
    _count()
    return y, hi, lo


def colsum(x: Tensor, out: Tensor, n: Optional[int] = None, ld: Optional[int] = None, rows: Optional[int] = None) -> None:
    """out[c] += sum_r x[r, c] over a [rows, n] view with row stride ld."""
    x2 = x.reshape(-1, x.shape[-1])
    rows = x2.shape[0] if rows is None else rows
    n = x2.shape[1] if n is None else n
    ld = x2.stride(0) if ld is None else ld
    check(lib().b2t_colsum(ptr(x), ptr(out), c_int64(rows), c_int(n), c_int64(ld), stream()), "b2t_colsum")
    _count()


def axpby(a: Tensor, b: Optional[Tensor], alpha: float = 1.0, beta: float = 1.0, out: Optional[Tensor] = None) -> Tensor:
    out = torch.empty_like(a) if out is None else out
    check(lib().b2t_axpby(ptr(a), ptr(b), ptr(out), c_int64(a.numel()), c_float(alpha), c_float(beta), stream()), "b2t_axpby")
    _count()
    return out


def attention_fwd(qkv: Tensor, B: int, T: int, H: int, d: int, three: bool = True, save_probs: bool = True,
                  drop_p: float = 0.0, seed: int = 0, kv_lens: Optional[Tensor] = None, want_f32: bool = False):
    """qkv [B*T, 3*H*d] (q | k | v) -> (ctx | None, ctx_hi, ctx_lo | None, probs | None)."""
    _lib.ensure_device()
    hd = H * d
    probs = torch.empty(B, H, T, T, dtype=torch.float32, device=qkv.device) if save_probs else None
    ctx = torch.empty(B * T, hd, dtype=torch.float32, device=qkv.device) if want_f32 else None
    hi = torch.empty(B * T, hd, dtype=torch.bfloat16, device=qkv.device)
    lo = torch.empty(B * T, hd, dtype=torch.bfloat16, device=qkv.device) if three else None
    check(lib().b2t_attention_fwd(ptr(qkv), c_int64(qkv.stride(0)), c_int(0), c_int(hd), c_int(2 * hd), ptr(kv_lens),
                                  ptr(probs), ptr(ctx), ptr(hi), ptr(lo), c_int64(hd), c_int(B), c_int(T), c_int(H),
                                  c_int(d), c_float(d ** -0.5), c_float(drop_p), c_uint64(seed), stream()),
          "b2t_attention_fwd")
    _count()
    return ctx, hi, lo, probs


def attention_bwd(qkv: Tensor, probs: Tensor, dctx: Tensor, ds_ws: Tensor, B: int, T: int, H: int, d: int,
                  three: bool = True, drop_p: float = 0.0, seed: int = 0):
    """-> (dqkv fp32, dqkv_hi, dqkv_lo | None) packed like qkv."""
    hd = H * d
    dqkv = torch.empty_like(qkv)
    hi, lo = _bf16_like(qkv), _bf16_like(qkv, three)
    check(lib().b2t_attention_bwd(ptr(qkv), c_int64(qkv.stride(0)), c_int(0), c_int(hd), c_int(2 * hd), ptr(probs),
                                  ptr(dctx), c_int64(dctx.stride(0)), ptr(ds_ws), ptr(dqkv), ptr(hi), ptr(lo), c_int(B),
                                  c_int(T), c_int(H), c_int(d), c_float(d ** -0.5), c_float(drop_p), c_uint64(seed),
                                  stream()), "b2t_attention_bwd")
    _count()
    return dqkv, hi, lo


def gru_supported_hidden(H: int) -> bool:
    return bool(lib().b2t_gru_supported_hidden(c_int(H)))


def gru_fwd(gi: Tensor, w_hh: Tensor, b_hh: Tensor, h0: Optional[Tensor], B: int, T: int, H: int, ndir: int,
            three: bool = True, training: bool = True, drop_p: float = 0.0, seed: int = 0):
    """gi [B*T, ndir*3H]; w_hh [ndir,3H,H]; b_hh [ndir,3H]; h0 [ndir,H] | None.
    -> dict(out, out_hi, out_lo, r, z, n, ghn, hprev_hi, hprev_lo)"""
    _lib.ensure_device()
    dev = gi.device
    M = B * T
    f32 = dict(dtype=torch.float32, device=dev)
    b16 = dict(dtype=torch.bfloat16, device=dev)
    res = {"out": torch.empty(M, ndir * H, **f32), "out_hi": torch.empty(M, ndir * H, **b16),
           "out_lo": torch.empty(M, ndir * H, **b16) if three else None}
    if training:
        for k in ("r", "z", "n", "ghn"):
            res[k] = torch.empty(ndir, M, H, **f32)
        res["hprev_hi"] = torch.empty(ndir, M, H, **b16)
        res["hprev_lo"] = torch.empty(ndir, M, H, **b16) if three else None
    else:
        for k in ("r", "z", "n", "ghn", "hprev_hi", "hprev_lo"):
            res[k] = None
    check(lib().b2t_gru_fwd(ptr(gi), ptr(w_hh), ptr(b_hh), ptr(h0), ptr(res["out"]), ptr(res["out_hi"]), ptr(res["out_lo"]),
                            ptr(res["r"]), ptr(res["z"]), ptr(res["n"]), ptr(res["ghn"]), ptr(res["hprev_hi"]),
                            ptr(res["hprev_lo"]), c_int(B), c_int(T), c_int(H), c_int(ndir), c_float(drop_p), c_uint64(seed),
                            stream()), "b2t_gru_fwd")
    _count()
    return res


def gru_bwd(dout: Tensor, fw: dict, w_hh: Tensor, h0: Optional[Tensor], dh0: Optional[Tensor], B: int, T: int, H: int,
            ndir: int, three: bool = True, drop_p: float = 0.0, seed: int = 0):
    """-> dict(dgi, dgi_hi, dgi_lo, dgh_hi, dgh_lo, dghn)"""
    dev = dout.device
    M = B * T
    f32 = dict(dtype=torch.float32, device=dev)
    b16 = dict(dtype=torch.bfloat16, device=dev)
    res = {"dgi": torch.empty(M, ndir * 3 * H, **f32), "dgi_hi": torch.empty(M, ndir * 3 * H, **b16),
           "dgi_lo": torch.empty(M, ndir * 3 * H, **b16) if three else None,
           "dgh_hi": torch.empty(ndir, M, 3 * H, **b16), "dgh_lo": torch.empty(ndir, M, 3 * H, **b16) if three else None,
           "dghn": torch.empty(ndir, M, H, **f32)}
    check(lib().b2t_gru_bwd(ptr(dout), ptr(fw["out"]), ptr(w_hh), ptr(h0), ptr(fw["r"]), ptr(fw["z"]), ptr(fw["n"]),
                            ptr(fw["ghn"]), ptr(res["dgi"]), ptr(res["dgi_hi"]), ptr(res["dgi_lo"]), ptr(res["dgh_hi"]),
                            ptr(res["dgh_lo"]), ptr(res["dghn"]), ptr(dh0), c_int(B), c_int(T), c_int(H), c_int(ndir),
                            c_float(drop_p), c_uint64(seed), stream()), "b2t_gru_bwd")
    _count()
    return res


def ctc_fwd_bwd(logits: Tensor, targets: Tensor, input_lens: Tensor, target_lens: Tensor, want_grad: bool = True):
    """logits [B,T,V] fp32, targets [B,S] int64, input_lens [B] int32, target_lens [B] int64.
    -> (loss 0-dim, nll [B], dlogits [B,T,V] | None)"""
    _lib.ensure_device()
    B, T, V = logits.shape
    S = targets.shape[1]
    lib().b2t_ctc_workspace_floats.restype = ctypes.c_int64
    nws = lib().b2t_ctc_workspace_floats(c_int(B), c_int(T), c_int(S))
    ws = torch.empty(nws, dtype=torch.float32, device=logits.device)
    nll = torch.empty(B, dtype=torch.float32, device=logits.device)
    loss = torch.zeros((), dtype=torch.float32, device=logits.device)
    grad = torch.empty_like(logits) if want_grad else None
    assert targets.dtype == torch.int64 and target_lens.dtype == torch.int64 and input_lens.dtype == torch.int32
    assert targets.is_contiguous() and logits.stride(2) == 1
    assert V <= 32, "the CTC kernel keeps one vocabulary entry per lane (V <= 32)"
    check(lib().b2t_ctc_fwd_bwd(ptr(logits), c_int64(logits.stride(1)), c_int64(logits.stride(0)), ptr(targets), c_int(S),
                                ptr(input_lens), ptr(target_lens), ptr(ws), ptr(nll), ptr(loss), ptr(grad), c_int(B),
                                c_int(T), c_int(V), c_int(1), stream()), "b2t_ctc_fwd_bwd")
    _count()
    return loss, nll, grad


def greedy_decode(logits: Tensor):
    """-> (argmax [B,T] int32, ids [B,T] int32 collapsed/no blanks/0-padded, lens [B] int32)"""
    _lib.ensure_device()
    B, T, V = logits.shape
    am = torch.empty(B, T, dtype=torch.int32, device=logits.device)
    ids = torch.empty(B, T, dtype=torch.int32, device=logits.device)
    lens = torch.empty(B, dtype=torch.int32, device=logits.device)
    check(lib().b2t_greedy_decode(ptr(logits), c_int64(logits.stride(1)), c_int64(logits.stride(0)), ptr(am), ptr(ids),
                                  ptr(lens), c_int(B), c_int(T), c_int(V), stream()), "b2t_greedy_decode")
    _count()
    return am, ids, lens


POSCONV_CP = 64   # channels per group after padding (one 128-byte swizzle row)


def posconv_pack_input(x: Tensor, B: int, T: int, G: int, pad_left: int, Tp: int, three: bool = True):
    """x [B*T, D] -> xg (hi, lo) [G, B, Tp, 64]."""
    _lib.ensure_device()
    D = x.shape[-1]
    hi = torch.empty(G, B, Tp, POSCONV_CP, dtype=torch.bfloat16, device=x.device)
    lo = torch.empty_like(hi) if three else None
    check(lib().b2t_posconv_pack_input(ptr(x), ptr(hi), ptr(lo), c_int(B), c_int(T), c_int(D), c_int(G), c_int(pad_left),
                                       c_int(Tp), stream()), "b2t_posconv_pack_input")
    _count()
    return hi, lo


def posconv_pack_weight(g: Tensor, v: Tensor, G: int, three: bool = True):
    """weight_norm(g [1,1,K], v [D,Cg,K]) -> (wg_hi, wg_lo [D, K*64], norm [K])."""
    _lib.ensure_device()
    D, Cg, K = v.shape
    norm = torch.empty(K, dtype=torch.float32, device=v.device)
    hi = torch.empty(D, K * POSCONV_CP, dtype=torch.bfloat16, device=v.device)
    lo = torch.empty_like(hi) if three else None
    check(lib().b2t_posconv_pack_weight(ptr(g), ptr(v), ptr(norm), ptr(hi), ptr(lo), c_int(D), c_int(Cg), c_int(K), c_int(G),
                                        stream()), "b2t_posconv_pack_weight")
    _count(2)
    return hi, lo, norm


def posconv_unpack_dx(dxg: Tensor, dx: Tensor, B: int, T: int, G: int, pad_left: int, Tp: int) -> None:
    D = dx.shape[-1]
    check(lib().b2t_posconv_unpack_dx(ptr(dxg), ptr(dx), c_int(B), c_int(T), c_int(D), c_int(G), c_int(pad_left), c_int(Tp),
                                      stream()), "b2t_posconv_unpack_dx")
    _count()


def posconv_weight_bwd(dwg: Tensor, g: Tensor, v: Tensor, norm: Tensor, dg_out: Optional[Tensor] = None,
                       dv_out: Optional[Tensor] = None):
    D, Cg, K = v.shape
    dg = torch.empty_like(g) if dg_out is None else dg_out
    dv = torch.empty_like(v) if dv_out is None else dv_out
    check(lib().b2t_posconv_weight_bwd(ptr(dwg), ptr(g), ptr(v), ptr(norm), ptr(dg), ptr(dv), c_int(D), c_int(Cg), c_int(K),
                                       stream()), "b2t_posconv_weight_bwd")
    _count()
    return dg, dv


def act_bwd(dy: Tensor, pre: Tensor, mode, three: bool = True, want_f32: bool = False, want_split: bool = True):
    out = torch.empty_like(dy) if want_f32 else None
    hi = _bf16_like(dy, want_split)
    lo = _bf16_like(dy, want_split and three)
    check(lib().b2t_act_bwd(ptr(dy), ptr(pre), ptr(out), ptr(hi), ptr(lo), c_int64(dy.numel()), c_int(ACT[mode]), stream()),
          "b2t_act_bwd")
    _count()
    return out, hi, lo


# ---------------------------------------------------------------------------------------------
# conformer kernels (csrc/conformer.cu)
# ---------------------------------------------------------------------------------------------
def rope_table(inv_freq: Tensor, T: int, head_dim: int):
    """-> (cos, sin) [T, head_dim] of Wav2Vec2ConformerRotaryPositionalEmbedding."""
    _lib.ensure_device()
    cos = torch.empty(T, head_dim, dtype=torch.float32, device=inv_freq.device)
    sin = torch.empty_like(cos)
    check(lib().b2t_rope_table(ptr(inv_freq), ptr(cos), ptr(sin), c_int(T), c_int(head_dim), stream()), "b2t_rope_table")
    _count()
    return cos, sin


def layernorm_rope_fwd(x: Tensor, gamma: Tensor, beta: Tensor, eps: float, cos: Tensor, sin: Tensor, T: int, head_dim: int,
                       three: bool = True):
    """x [rows, n] -> (y_hi, y_lo, rot_hi, rot_lo, mean, rstd): LayerNorm output plain and rotated."""
    _lib.ensure_device()
    rows, n = x.shape
    y_hi, y_lo = _bf16_like(x), _bf16_like(x, three)
    r_hi, r_lo = _bf16_like(x), _bf16_like(x, three)
    mean = torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    check(lib().b2t_layernorm_rope_fwd(ptr(x), ptr(gamma), ptr(beta), c_float(eps), ptr(y_hi), ptr(y_lo), ptr(r_hi), ptr(r_lo),
                                       ptr(mean), ptr(rstd), ptr(cos), ptr(sin), c_int(rows), c_int(n), c_int(T),
                                       c_int(head_dim), stream()), "b2t_layernorm_rope_fwd")
    _count()
    return y_hi, y_lo, r_hi, r_lo, mean, rstd


def rope_bwd_add(d_rot: Tensor, d_plain: Optional[Tensor], cos: Tensor, sin: Tensor, T: int, head_dim: int) -> Tensor:
    """gradient w.r.t. the un-rotated state: R^T d_rot (+ d_plain).  Inputs may be column slices (row stride != n)."""
    rows, n = d_rot.shape
    assert d_rot.stride(1) == 1 and (d_plain is None or d_plain.stride(1) == 1)
    out = torch.empty(rows, n, dtype=torch.float32, device=d_rot.device)
    check(lib().b2t_rope_bwd_add(ptr(d_rot), c_int64(d_rot.stride(0)), ptr(d_plain),
                                 c_int64(0 if d_plain is None else d_plain.stride(0)), ptr(out), ptr(cos), ptr(sin),
                                 c_int64(rows), c_int(n), c_int(T), c_int(head_dim), stream()), "b2t_rope_bwd_add")
    _count()
    return out


def glu_dwconv_fwd(g: Tensor, w: Tensor, B: int, T: int, want_stats: bool = True):
    """g [B*T, 2D] (value | gate), w [D, 1, K] -> (out [B*T, D], stats [2, D] float64 (sum, sum of squares) | None)."""
    _lib.ensure_device()
    D = g.shape[1] // 2
    K = w.shape[-1]
    out = torch.empty(B * T, D, dtype=torch.float32, device=g.device)
    stats = torch.zeros(2, D, dtype=torch.float64, device=g.device) if want_stats else None
    check(lib().b2t_glu_dwconv_fwd(ptr(g), ptr(w), ptr(out), ptr(stats), c_int(B), c_int(T), c_int(D), c_int(K), stream()),
          "b2t_glu_dwconv_fwd")
    _count()
    return out, stats


def bn_silu_fwd(x: Tensor, stats: Optional[Tensor], gamma: Tensor, beta: Tensor, eps: float, momentum: float,
                running_mean: Tensor, running_var: Tensor, num_batches_tracked: Optional[Tensor], training: bool,
                three: bool = True):
    """BatchNorm1d + SiLU.  -> (y_hi, y_lo, mean [D], rstd [D]); running statistics are updated in place when training."""
    M, D = x.shape
    y_hi, y_lo = _bf16_like(x), _bf16_like(x, three)
    mean = torch.empty(D, dtype=torch.float32, device=x.device)
    rstd = torch.empty(D, dtype=torch.float32, device=x.device)
    check(lib().b2t_bn_silu_fwd(ptr(x), ptr(stats), ptr(gamma), ptr(beta), c_float(eps), c_float(momentum), ptr(running_mean),
                                ptr(running_var), ptr(num_batches_tracked), ptr(mean), ptr(rstd), ptr(y_hi), ptr(y_lo),
                                c_int64(M), c_int(D), c_int(1 if training else 0), stream()), "b2t_bn_silu_fwd")
    _count()
    return y_hi, y_lo, mean, rstd


def bn_silu_bwd_stats(ds: Tensor, x: Tensor, mean: Tensor, rstd: Tensor, gamma: Tensor, beta: Tensor,
                      dgamma_out: Optional[Tensor] = None, dbeta_out: Optional[Tensor] = None):
    """-> (dgamma [D], dbeta [D]) of BatchNorm behind the SiLU (accumulated into the zero-initialised *_out if given)."""
    M, D = x.shape
    dgamma = zeros(D, device=x.device) if dgamma_out is None else dgamma_out
    dbeta = zeros(D, device=x.device) if dbeta_out is None else dbeta_out
    check(lib().b2t_bn_silu_bwd_stats(ptr(ds), ptr(x), ptr(mean), ptr(rstd), ptr(gamma), ptr(beta), ptr(dgamma), ptr(dbeta),
                                      c_int64(M), c_int(D), stream()), "b2t_bn_silu_bwd_stats")
    _count()
    return dgamma, dbeta


def glu_dwconv_bwd(ds: Tensor, x: Tensor, g: Tensor, w: Tensor, mean: Tensor, rstd: Tensor, gamma: Tensor, beta: Tensor,
                   dgamma: Tensor, dbeta: Tensor, B: int, T: int, training: bool, three: bool = True,
                   dw_out: Optional[Tensor] = None):
    """-> (dg_hi, dg_lo [B*T, 2D], dw [D, 1, K]): gradient w.r.t. the pointwise_conv1 output and the depthwise weight."""
    D = x.shape[1]
    K = w.shape[-1]
    dg_hi, dg_lo = _bf16_like(g), _bf16_like(g, three)
    dw = zeros(tuple(w.shape), device=w.device) if dw_out is None else dw_out
    check(lib().b2t_glu_dwconv_bwd(ptr(ds), ptr(x), ptr(g), ptr(w), ptr(mean), ptr(rstd), ptr(gamma), ptr(beta), ptr(dgamma),
                                   ptr(dbeta), ptr(dg_hi), ptr(dg_lo), ptr(dw), c_int(B), c_int(T), c_int(D), c_int(K),
                                   c_int(1 if training else 0), stream()), "b2t_glu_dwconv_bwd")
    _count()
    return dg_hi, dg_lo, dw
