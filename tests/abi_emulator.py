"""CPU emulator of the C ABI (include/b2t_sm100.h) with the call signatures of
wav2vec2forbrain_b200.ops.  TEST INFRASTRUCTURE ONLY: it lets the `-m "not gpu"` suite run the
engine's forward/backward schedule (operand views, epilogue wiring, gradient bookkeeping) against
the oracle without a GPU.  The product never imports it; on a GPU box the same engine code calls
the real kernels and is checked against the same oracle.

Semantics follow the kernels bit-for-bit where it matters for the schedule: bf16 hi/lo operand
storage, 3-D strided operand views (including overlapping unfold views), output addressing
slot*c_batch_stride + row*ldc + col with red.add accumulation, and the counter-based dropout hash.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from wav2vec2forbrain_b200.ops import ACT, POSCONV_CP, Operand, kmajor, mnmajor, unfold_view  # noqa: F401 (pure python)

LAUNCHES = 0
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def zeros(*shape, device):
    if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
        shape = tuple(shape[0])
    return torch.zeros(shape, dtype=torch.float32, device=device)


def hash_uniform(seed: int, idx: torch.Tensor) -> torch.Tensor:
    """same 32-bit mixer as common.cuh::hash_uniform"""
    with np.errstate(over="ignore"):
        i = idx.detach().cpu().numpy().astype(np.uint64)
        s = seed & 0xFFFFFFFFFFFFFFFF
        m32 = np.uint64(0xFFFFFFFF)
        h = (((i & m32) + (i >> np.uint64(32)) * np.uint64(0x85EBCA77)) & m32) ^ np.uint64(s & 0xFFFFFFFF)
        h = (h * np.uint64(0x9E3779B1)) & m32
        h = h ^ np.uint64(s >> 32)
        h = h ^ (h >> np.uint64(16))
        h = (h * np.uint64(0x85EBCA6B)) & m32
        h = h ^ (h >> np.uint64(13))
        h = (h * np.uint64(0xC2B2AE35)) & m32
        h = h ^ (h >> np.uint64(16))
        u = (h >> np.uint64(8)).astype(np.float64) / 16777216.0
    return torch.from_numpy(u)


def _keep(p: float, seed: int, offs: torch.Tensor) -> torch.Tensor:
    if p <= 0.0:
        return torch.ones(offs.shape, dtype=torch.float64)
    u = hash_uniform(seed, offs)
    return torch.where(u < p, torch.zeros_like(u), torch.full_like(u, 1.0 / (1.0 - p)))


def _split_t(x: torch.Tensor, three: bool):
    hi = x.to(torch.bfloat16)
    lo = (x - hi.float()).to(torch.bfloat16) if three else None
    return hi, lo


def _full(op_hi, op_lo):
    return op_hi.double() + (op_lo.double() if op_lo is not None else 0.0)


def _base(t: torch.Tensor) -> torch.Tensor:
    n = t.untyped_storage().nbytes() // t.element_size()
    return torch.as_strided(t, (n,), (1,), 0)


def _view3(op: Operand) -> torch.Tensor:
    def v(t):
        return torch.as_strided(t, (op.d2, op.d1, op.d0), (op.s2, op.s1, 1), t.storage_offset()).double()
    x = v(op.hi)
    if op.lo is not None:
        x = x + v(op.lo)
    return x   # [d2, d1, d0]


def split(x, three=True, transpose=False):
    x2 = x.reshape(-1, x.shape[-1])
    if transpose:
        return _split_t(x2.t().contiguous(), three)
    return _split_t(x.contiguous(), three)


def permute_unfold_cols(w, C, KL, three=True):
    rows = w.shape[0]
    wp = w.reshape(rows, C, KL).permute(0, 2, 1).reshape(rows, KL * C).contiguous()
    return _split_t(wp, three)


def unpermute_unfold_cols(g, C, KL, out=None):
    rows = g.shape[0]
    r = g.reshape(rows, KL, C).permute(0, 2, 1).reshape(rows, C * KL).contiguous()
    if out is not None:
        out.copy_(r)
        return out
    return r


def _act(x, mode):
    m = ACT[mode]
    if m == 1:
        return F.gelu(x)
    if m == 2:
        return F.silu(x)
    if m == 3:
        return F.softsign(x)
    return x


def _dact(a, mode):
    with torch.enable_grad():      # the engine's backward runs inside autograd.Function.backward (grad mode off)
        a = a.detach().clone().requires_grad_(True)
        g, = torch.autograd.grad(_act(a, mode).sum(), a)
    return g


def gemm(a, b, *, m_per_batch, n, k, batches=1, out_f32=None, out_pre=None, out_hi=None, out_lo=None, ldc=None,
         c_batch_stride=None, bias=None, residual=None, aux=None, colsum=None, act=None, aux_mode=None, atomic=False,
         split_k=1, alpha=1.0, k_spans_batches=False, a_batched=False, b_batched=False, b_batch_idx=None,
         c_batch_idx=None, block_n=0, drop_p=0.0, drop_seed=0, bias_batch_stride=0, out_scale=1.0, stream_k=0, cta_group=0,
         on_stream=None, pre_mode=0):
    global LAUNCHES
    LAUNCHES += 1
    ldc = n if ldc is None else ldc
    cbs = m_per_batch * ldc if c_batch_stride is None else c_batch_stride
    assert split_k == 1 or (atomic and ACT[act] == 0 and out_hi is None and out_pre is None)
    A3, B3 = _view3(a), _view3(b)
    if k_spans_batches:
        assert a.mn_major and b.mn_major and batches == 1
    rows = torch.arange(m_per_batch).view(-1, 1)
    cols = torch.arange(n).view(1, -1)
    for bt in range(batches):
        bslot = int(b_batch_idx[bt]) if b_batch_idx is not None else (bt if b_batched else 0)
        if k_spans_batches:
            acc = torch.einsum("bkm,bkn->mn", A3[:, :k, :m_per_batch], B3[:, :k, :n])
        else:
            Ab = A3[bt if a_batched else 0]
            Ab = Ab[:k, :m_per_batch].t() if a.mn_major else Ab[:m_per_batch, :k]
            Bb = B3[bslot]
            Bb = Bb[:k, :n].t() if b.mn_major else Bb[:n, :k]
            assert Ab.shape == (m_per_batch, k) and Bb.shape == (n, k), (Ab.shape, Bb.shape, m_per_batch, n, k)
            acc = Ab @ Bb.t()
        x = acc * alpha
        if bias is not None:
            x = x + _base(bias)[bias.storage_offset() + bslot * bias_batch_stride + cols.flatten()].double().view(1, -1)
        cslot = int(c_batch_idx[bt]) if c_batch_idx is not None else bt
        offs = cslot * cbs + rows * ldc + cols          # relative to each tensor's own base pointer

        def rd(t):
            return _base(t)[t.storage_offset() + offs].double()

        def wr(t, val, add=False):
            bt_ = _base(t)
            idx = (t.storage_offset() + offs).flatten()
            if add:
                bt_.index_add_(0, idx, val.flatten().to(t.dtype))
            else:
                bt_[idx] = val.flatten().to(t.dtype)
        if out_pre is not None and pre_mode == 0:
            wr(out_pre, x)
        dfac = _dact(x, act) if (pre_mode == 1 and ACT[act]) else torch.ones_like(x)
        x = _act(x, act)
        if ACT[aux_mode] == 4:
            x = x * rd(aux)
        elif ACT[aux_mode]:
            x = x * _dact(rd(aux), aux_mode)
        if drop_p > 0:
            keep = _keep(drop_p, drop_seed, offs)
            x = x * keep
            dfac = dfac * keep
        x = x * out_scale
        if out_pre is not None and pre_mode == 1:
            wr(out_pre, dfac * out_scale)
        if residual is not None:
            x = x + rd(residual)
        if colsum is not None:
            colsum += x.sum(0).to(colsum.dtype)
        if out_f32 is not None:
            wr(out_f32, x, add=atomic)
        if out_hi is not None:
            h = x.float().to(torch.bfloat16)
            wr(out_hi, h)
            if out_lo is not None:
                wr(out_lo, (x.float() - h.float()).to(torch.bfloat16))


def smooth_fwd(x, taps, pad_left, three=True, want_f32=False, white_sd=0.0, offset_sd=0.0, seed=0):
    assert white_sd == 0.0 and offset_sd == 0.0, "augmentation noise is checked on the GPU only"
    B, T, C = x.shape
    w = taps.reshape(C, -1).double()
    K = w.shape[1]
    xp = F.pad(x.double(), (0, 0, pad_left, K - 1 - pad_left))
    y = torch.zeros(B, T, C, dtype=torch.float64)
    for kk in range(K):
        y = y + w[:, kk].view(1, 1, C) * xp[:, kk:kk + T, :]
    y = y.float()
    hi, lo = _split_t(y, three)
    return (y if want_f32 else None), hi, lo


def softsign_bwd(ds, z, day, dbias, three=True):
    dz = ds / (1 + z.abs()) ** 2
    for b in range(ds.shape[0]):
        dbias[int(day[b])] += dz[b].sum(0).view_as(dbias[0])
    return _split_t(dz, three)


def layernorm_fwd(x, gamma, beta, eps, three=True, want_f32=True, want_split=True, drop_p=0.0, seed=0):
    xd = x.double()
    mean = xd.mean(-1)
    var = xd.var(-1, unbiased=False)
    rstd = (var + eps).rsqrt()
    y = (xd - mean[:, None]) * rstd[:, None] * gamma.double() + beta.double()
    if drop_p > 0:
        y = y * _keep(drop_p, seed, torch.arange(x.numel()).view_as(x))
    y = y.float()
    hi, lo = _split_t(y, three) if want_split else (None, None)
    return (y if want_f32 else None), hi, lo, mean.float(), rstd.float()


def layernorm_bwd(dy, x, gamma, mean, rstd, dgamma, dbeta, dcol=None, three=True, want_f32=True, want_split=True,
                  out_drop_p=0.0, out_seed=0, br_drop_p=0.0, br_seed=0, dres=None):
    offs = torch.arange(x.numel()).view_as(x)
    d = dy.double() * _keep(out_drop_p, out_seed, offs)
    xh = (x.double() - mean.double()[:, None]) * rstd.double()[:, None]
    dgamma += (d * xh).sum(0).float()
    dbeta += d.sum(0).float()
    g = d * gamma.double()
    n = x.shape[1]
    dx = rstd.double()[:, None] * (g - g.mean(-1, keepdim=True) - xh * (g * xh).mean(-1, keepdim=True))
    if dres is not None:
        dx = dx + dres.double()
    br = dx * _keep(br_drop_p, br_seed, offs)
    if dcol is not None:
        dcol += br.sum(0).float()
    hi, lo = _split_t(br.float(), three) if want_split else (None, None)
    return (dx.float() if want_f32 else None), hi, lo


def dropout(x, p, seed, three=True, want_f32=True, want_split=False):
    y = (x.double() * _keep(p, seed, torch.arange(x.numel()).view_as(x))).float()
    hi, lo = _split_t(y, three) if want_split else (None, None)
    return (y if want_f32 else None), hi, lo


def colsum(x, out, n=None, ld=None, rows=None):
    x2 = x.reshape(-1, x.shape[-1])
    rows = x2.shape[0] if rows is None else rows
    out += x2[:rows].double().sum(0).float()


def axpby(a, b, alpha=1.0, beta=1.0, out=None):
    r = alpha * a + (beta * b if b is not None else 0.0)
    if out is not None:
        out.copy_(r)
        return out
    return r


def _attn_mask(p, seed, B, H, T):
    return _keep(p, seed, torch.arange(B * H * T * T).view(B, H, T, T))


def attention_fwd(qkv, B, T, H, d, three=True, save_probs=True, drop_p=0.0, seed=0, kv_lens=None, want_f32=False):
    hd = H * d
    q, k, v = [t.double().reshape(B, T, H, d).transpose(1, 2) for t in qkv.split(hd, dim=1)]
    s = q @ k.transpose(-1, -2) * d ** -0.5
    if kv_lens is not None:
        for b in range(B):
            s[b, :, :, int(kv_lens[b]):] = float("-inf")
    p = torch.softmax(s, -1)
    pd = p * _attn_mask(drop_p, seed, B, H, T)
    ctx = (pd @ v).transpose(1, 2).reshape(B * T, hd).float()
    hi, lo = _split_t(ctx, three)
    return (ctx if want_f32 else None), hi, lo, (p.float() if save_probs else None)


def attention_bwd(qkv, probs, dctx, ds_ws, B, T, H, d, three=True, drop_p=0.0, seed=0):
    hd = H * d
    q, k, v = [t.double().reshape(B, T, H, d).transpose(1, 2) for t in qkv.split(hd, dim=1)]
    do = dctx.double().reshape(B, T, H, d).transpose(1, 2)
    p = probs.double()
    mask = _attn_mask(drop_p, seed, B, H, T)
    dv = (p * mask).transpose(-1, -2) @ do
    dp = (do @ v.transpose(-1, -2)) * mask
    ds = p * (dp - (dp * p).sum(-1, keepdim=True)) * d ** -0.5
    dq = ds @ k
    dk = ds.transpose(-1, -2) @ q
    out = torch.cat([t.transpose(1, 2).reshape(B * T, hd) for t in (dq, dk, dv)], 1).float()
    hi, lo = _split_t(out, three)
    return out, hi, lo


def gru_supported_hidden(H):
    return H in (32, 64, 128, 256, 512)


def gru_fwd(gi, w_hh, b_hh, h0, B, T, H, ndir, three=True, training=True, drop_p=0.0, seed=0):
    M = B * T
    g4 = gi.double().reshape(B, T, ndir, 3 * H)
    out = torch.zeros(B, T, ndir * H, dtype=torch.float64)
    sv = {k_: torch.zeros(ndir, B, T, H, dtype=torch.float64) for k_ in ("r", "z", "n", "ghn", "hprev")}
    for dr in range(ndir):
        h = h0[dr].double().expand(B, H) if h0 is not None else torch.zeros(B, H, dtype=torch.float64)
        for s in range(T):
            t = s if dr == 0 else T - 1 - s
            gh = h @ w_hh[dr].double().t() + b_hh[dr].double()
            g = g4[:, t, dr]
            r = torch.sigmoid(g[:, :H] + gh[:, :H])
            z = torch.sigmoid(g[:, H:2 * H] + gh[:, H:2 * H])
            n = torch.tanh(g[:, 2 * H:] + r * gh[:, 2 * H:])
            sv["r"][dr, :, t], sv["z"][dr, :, t], sv["n"][dr, :, t] = r, z, n
            sv["ghn"][dr, :, t], sv["hprev"][dr, :, t] = gh[:, 2 * H:], h
            h = (1 - z) * n + z * h
            out[:, t, dr * H:(dr + 1) * H] = h
    out = out.reshape(M, ndir * H).float()
    od = (out.double() * _keep(drop_p, seed, torch.arange(out.numel()).view_as(out))).float()
    res = {"out": out}
    res["out_hi"], res["out_lo"] = _split_t(od, three)
    if training:
        for k_ in ("r", "z", "n", "ghn"):
            res[k_] = sv[k_].reshape(ndir, M, H).float()
        res["hprev_hi"], res["hprev_lo"] = _split_t(sv["hprev"].reshape(ndir, M, H).float(), three)
    else:
        for k_ in ("r", "z", "n", "ghn", "hprev_hi", "hprev_lo"):
            res[k_] = None
    return res


def gru_bwd(dout, fw, w_hh, h0, dh0, B, T, H, ndir, three=True, drop_p=0.0, seed=0):
    M = B * T
    d = (dout.double() * _keep(drop_p, seed, torch.arange(dout.numel()).view_as(dout))).reshape(B, T, ndir, H)
    out = fw["out"].double().reshape(B, T, ndir, H)
    R, Z, N, GHN = [fw[k_].double().reshape(ndir, B, T, H) for k_ in ("r", "z", "n", "ghn")]
    dgi = torch.zeros(B, T, ndir, 3 * H, dtype=torch.float64)
    dgh = torch.zeros(ndir, B, T, 3 * H, dtype=torch.float64)
    for dr in range(ndir):
        W = w_hh[dr].double()
        rec = torch.zeros(B, H, dtype=torch.float64)
        for s in range(T):
            t = T - 1 - s if dr == 0 else s
            tp = t - 1 if dr == 0 else t + 1
            if 0 <= tp < T:
                hp = out[:, tp, dr]
            else:
                hp = h0[dr].double().expand(B, H) if h0 is not None else torch.zeros(B, H, dtype=torch.float64)
            r, z, n, ghn = R[dr, :, t], Z[dr, :, t], N[dr, :, t], GHN[dr, :, t]
            dh = rec + d[:, t, dr]
            dan = dh * (1 - z) * (1 - n * n)
            dar = dan * ghn * r * (1 - r)
            daz = dh * (hp - n) * z * (1 - z)
            dgi[:, t, dr] = torch.cat([dar, daz, dan], -1)
            gh_ = torch.cat([dar, daz, dan * r], -1)
            dgh[dr, :, t] = gh_
            rec = dh * z + gh_ @ W
        if dh0 is not None:
            dh0[dr] += rec.sum(0).float()
    res = {"dgi": dgi.reshape(M, ndir * 3 * H).float()}
    res["dgi_hi"], res["dgi_lo"] = _split_t(res["dgi"], three)
    g2 = dgh.reshape(ndir, M, 3 * H).float()
    res["dgh_hi"], res["dgh_lo"] = _split_t(g2, three)
    res["dghn"] = g2[:, :, 2 * H:].contiguous()
    return res


def ctc_fwd_bwd(logits, targets, input_lens, target_lens, want_grad=True):
    with torch.enable_grad():     # also called from inside autograd.Function.forward (models._CTCFn)
        x = logits.detach().double().requires_grad_(True)
        lp = F.log_softmax(x, -1).transpose(0, 1)
        loss = F.ctc_loss(lp, targets, input_lens.long(), target_lens, blank=0, reduction="mean", zero_infinity=True)
        nll = F.ctc_loss(lp, targets, input_lens.long(), target_lens, blank=0, reduction="none", zero_infinity=True)
        grad = None
        if want_grad:
            grad, = torch.autograd.grad(loss, x)
            grad = grad.float()
    return loss.detach().float(), nll.detach().float(), grad


def greedy_decode(logits):
    am = logits.argmax(-1).int()
    B, T = am.shape
    ids = torch.zeros(B, T, dtype=torch.int32)
    lens = torch.zeros(B, dtype=torch.int32)
    for b in range(B):
        prev, c = -1, 0
        for t in range(T):
            v = int(am[b, t])
            if v != prev and v != 0:
                ids[b, c] = v
                c += 1
            prev = v
        lens[b] = c
    return am, ids, lens


def posconv_pack_input(x, B, T, G, pad_left, Tp, three=True):
    D = x.shape[-1]
    Cg = D // G
    xg = torch.zeros(G, B, Tp, POSCONV_CP)
    x4 = x.reshape(B, T, G, Cg)
    for g in range(G):
        xg[g, :, pad_left:pad_left + T, :Cg] = x4[:, :, g]
    return _split_t(xg, three)


def posconv_pack_weight(g, v, G, three=True):
    D, Cg, K = v.shape
    norm = v.double().pow(2).sum((0, 1)).sqrt()
    wn = v.double() * (g.double().reshape(1, 1, K) / norm.view(1, 1, K))
    wg = torch.zeros(D, K, POSCONV_CP, dtype=torch.float64)
    wg[:, :, :Cg] = wn.permute(0, 2, 1)
    hi, lo = _split_t(wg.reshape(D, K * POSCONV_CP).float(), three)
    return hi, lo, norm.float()


def posconv_unpack_dx(dxg, dx, B, T, G, pad_left, Tp):
    D = dx.shape[-1]
    Cg = D // G
    d4 = dx.view(B, T, G, Cg)
    for g in range(G):
        d4[:, :, g] += dxg[g, :, pad_left:pad_left + T, :Cg]


def posconv_weight_bwd(dwg, g, v, norm, dg_out=None, dv_out=None):
    D, Cg, K = v.shape
    dwn = dwg.double().reshape(D, K, POSCONV_CP)[:, :, :Cg].permute(0, 2, 1)   # [D, Cg, K]
    vd, nd, gd = v.double(), norm.double().view(1, 1, K), g.double().reshape(1, 1, K)
    s = (dwn * vd).sum((0, 1), keepdim=True)
    dg = (s / nd).reshape(g.shape).float()
    dv = (dwn * gd / nd - vd * gd * s / nd ** 3).float()
    if dg_out is not None:
        dg = dg_out.copy_(dg)
    if dv_out is not None:
        dv = dv_out.copy_(dv)
    return dg, dv


def act_bwd(dy, pre, mode, three=True, want_f32=False, want_split=True):
    v = (dy.double() * _dact(pre.double(), mode)).float()
    hi, lo = _split_t(v, three) if want_split else (None, None)
    return (v if want_f32 else None), hi, lo


# ---------------------------------------------------------------------------------------------
# conformer kernels
# ---------------------------------------------------------------------------------------------
def rope_table(inv_freq, T, head_dim):
    global LAUNCHES
    LAUNCHES += 1
    f = torch.arange(T, dtype=torch.float32)[:, None] * inv_freq.float()[None, :]
    emb = torch.cat((f, f), dim=-1)
    return emb.cos(), emb.sin()


def _rot_half(v, hd):
    r, n = v.shape
    u = v.view(r, n // hd, hd)
    half = hd // 2
    return torch.cat((-u[..., half:], u[..., :half]), dim=-1).reshape(r, n)


def layernorm_rope_fwd(x, gamma, beta, eps, cos, sin, T, head_dim, three=True):
    global LAUNCHES
    LAUNCHES += 1
    xd = x.double()
    mean = xd.mean(-1)
    var = ((xd - mean[:, None]) ** 2).mean(-1)
    rstd = 1.0 / torch.sqrt(var + eps)
    y = ((xd - mean[:, None]) * rstd[:, None] * gamma.double() + beta.double()).float()
    rows, n = x.shape
    t = torch.arange(rows) % T
    c = cos.double()[t].repeat(1, n // head_dim)
    s = sin.double()[t].repeat(1, n // head_dim)
    rot = (y.double() * c + _rot_half(y.double(), head_dim) * s).float()
    y_hi, y_lo = _split_t(y, three)
    r_hi, r_lo = _split_t(rot, three)
    return y_hi, y_lo, r_hi, r_lo, mean.float(), rstd.float()


def rope_bwd_add(d_rot, d_plain, cos, sin, T, head_dim):
    global LAUNCHES
    LAUNCHES += 1
    rows, n = d_rot.shape
    t = torch.arange(rows) % T
    c = cos.double()[t].repeat(1, n // head_dim)
    s = sin.double()[t].repeat(1, n // head_dim)
    d = d_rot.double()
    ys = d * s
    u = ys.view(rows, n // head_dim, head_dim)
    half = head_dim // 2
    rt = torch.cat((u[..., half:], -u[..., :half]), dim=-1).reshape(rows, n)
    out = d * c + rt
    if d_plain is not None:
        out = out + d_plain.double()
    return out.float()


def _glu(g):
    D = g.shape[1] // 2
    return g[:, :D].double() * torch.sigmoid(g[:, D:].double())


def glu_dwconv_fwd(g, w, B, T, want_stats=True):
    global LAUNCHES
    LAUNCHES += 1
    D = g.shape[1] // 2
    K = w.shape[-1]
    s = _glu(g).view(B, T, D).transpose(1, 2)
    out = F.conv1d(s, w.double().view(D, 1, K), None, padding=(K - 1) // 2, groups=D).transpose(1, 2).reshape(B * T, D)
    stats = torch.stack([out.sum(0), (out * out).sum(0)]) if want_stats else None
    return out.float(), stats


def bn_silu_fwd(x, stats, gamma, beta, eps, momentum, running_mean, running_var, num_batches_tracked, training, three=True):
    global LAUNCHES
    LAUNCHES += 1
    M, D = x.shape
    if training:
        mean = stats[0] / M
        var = (stats[1] / M - mean * mean).clamp_min(0.0)
        running_mean.mul_(1 - momentum).add_(momentum * mean.float())
        running_var.mul_(1 - momentum).add_(momentum * (var * M / max(M - 1, 1)).float())
        if num_batches_tracked is not None:
            num_batches_tracked += 1
    else:
        mean, var = running_mean.double(), running_var.double()
    rstd = 1.0 / torch.sqrt(var + eps)
    y = F.silu((x.double() - mean) * rstd * gamma.double() + beta.double()).float()
    hi, lo = _split_t(y, three)
    return hi, lo, mean.float(), rstd.float()


def bn_silu_bwd_stats(ds, x, mean, rstd, gamma, beta, dgamma_out=None, dbeta_out=None):
    global LAUNCHES
    LAUNCHES += 1
    xh = (x.double() - mean.double()) * rstd.double()
    dbn = ds.double() * _dact(xh * gamma.double() + beta.double(), "silu")
    dg, db = (dbn * xh).sum(0).float(), dbn.sum(0).float()
    if dgamma_out is not None:
        dg = dgamma_out.add_(dg)
    if dbeta_out is not None:
        db = dbeta_out.add_(db)
    return dg, db


def glu_dwconv_bwd(ds, x, g, w, mean, rstd, gamma, beta, dgamma, dbeta, B, T, training, three=True, dw_out=None):
    global LAUNCHES
    LAUNCHES += 1
    M, D = x.shape
    K = w.shape[-1]
    xh = (x.double() - mean.double()) * rstd.double()
    dbn = ds.double() * _dact(xh * gamma.double() + beta.double(), "silu")
    if training:
        dx = gamma.double() * rstd.double() * (dbn - dbeta.double() / M - xh * dgamma.double() / M)
    else:
        dx = gamma.double() * rstd.double() * dbn
    with torch.enable_grad():
        gl = g.detach().double().requires_grad_(True)
        wl = w.detach().double().view(D, 1, K).requires_grad_(True)
        s = (gl[:, :D] * torch.sigmoid(gl[:, D:])).view(B, T, D).transpose(1, 2)
        out = F.conv1d(s, wl, None, padding=(K - 1) // 2, groups=D).transpose(1, 2).reshape(M, D)
        ggl, gwl = torch.autograd.grad(out, (gl, wl), dx.detach())
    hi, lo = _split_t(ggl.float(), three)
    dw = gwl.float().view_as(w)
    if dw_out is not None:
        dw = dw_out.add_(dw)
    return hi, lo, dw
