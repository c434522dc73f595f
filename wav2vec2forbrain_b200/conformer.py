"""Forward / backward kernel schedule of the Wav2Vec2-Conformer (rotary) encoder
(transformers modeling_wav2vec2_conformer.py: encoder forward :651-716, block :568-631, attention +
RoPE :420-527, convolution module :360-417, feed-forward :336-357).

Block (all pre-LN, residual stream `h` in fp32):
    h += 1/2 FFN1(LN(h));  h += Dropout(MHSA(rope(LN(h)), LN(h)));  h += Conv(h);  h += 1/2 FFN2(LN(h));  h = LN(h)
RoPE is applied to the layer-normed state BEFORE the Q/K projections, V uses the un-rotated state
(SURVEY.md fact 7), so Q|K and V are two GEMMs with different A operands writing one qkv buffer.
`pos_conv_embed` exists in the state_dict but is never called by the reference (no gradient).

Called through engine.Engine (same `ops` indirection, so the CPU suite can run this schedule on
the ABI emulator).
"""
from __future__ import annotations

from typing import Dict

import torch

from . import engine as _e

Tensor = torch.Tensor
Q = "w2v_encoder.wav2vec2_conformer.encoder."
BLOCK_LN_EPS = 1e-5      # nn.LayerNorm default inside the blocks (:577,581,589,591); the encoder LN uses config eps
BN_EPS, BN_MOMENTUM = 1e-5, 0.1


def _seed(seed: int, layer: int, site: int) -> int:
    return _e._mix(seed, 2000 + 20 * layer + site)


def _w2d(eng, P, name: str):
    """bf16 operand copy of a pointwise Conv1d weight [out, in, 1] seen as [out, in]."""
    w = P[name]
    return eng.weights.get(name, _e._ver(w), lambda: _e.ops.split(w.view(w.shape[0], w.shape[1]), eng.three))


def _rope(eng, P, T: int, hd: int):
    inv = P[Q + "embed_positions.inv_freq"]
    return eng.weights.get(f"rope_{T}_{hd}", _e._ver(inv), lambda: _e.ops.rope_table(inv, T, hd))


def _ffn_forward(eng, P, pfx: str, ln: str, h: Tensor, M: int, training: bool, seed: int, layer: int, site: int, st: dict,
                 tag: str) -> Tensor:
    """h + 1/2 * Dropout(W2 . Dropout(SiLU(W1 . LN(h) + b1)) + b2)"""
    ops, cfg = _e.ops, eng.w2v
    D, F = cfg.hidden, cfg.ffn
    p_h = cfg.hidden_dropout if training else 0.0
    p_act = cfg.activation_dropout if training else 0.0
    _, u_hi, u_lo, m, r = ops.layernorm_fwd(h, P[ln + ".weight"], P[ln + ".bias"], BLOCK_LN_EPS, three=eng.three, want_f32=False)
    _, pre, f_hi, f_lo = eng._linear_fwd(u_hi, u_lo, eng._w(P, pfx + "intermediate_dense.weight"),
                                         P[pfx + "intermediate_dense.bias"], M, F, D, want_f32=False, want_split=True,
                                         want_pre=training, act=cfg.hidden_act, drop_p=p_act,
                                         drop_seed=_seed(seed, layer, site), pre_mode=1 if training else 0)
    # pre_mode 1: `pre` holds SiLU'(x) * dropout factor (the whole backward factor of this elementwise tail), so the
    # backward GEMM's epilogue is one multiply instead of a sigmoid and a mask hash per element
    out, _, _, _ = eng._linear_fwd(f_hi, f_lo, eng._w(P, pfx + "output_dense.weight"), P[pfx + "output_dense.bias"], M, D, F,
                                   residual=h, drop_p=p_h, drop_seed=_seed(seed, layer, site + 1), out_scale=0.5)
    st[tag] = dict(x=h, ln=(m, r), u_hi=u_hi, u_lo=u_lo, pre=pre, f_hi=f_hi, f_lo=f_lo)
    return out


def forward(eng, P, feats: Tensor, B: int, T: int, training: bool, seed: int, skip_layers=()):
    """feats [B*T, D] -> (h [B*T, D] after the encoder LayerNorm, saved state)."""
    ops, cfg, three = _e.ops, eng.w2v, eng.three
    D, NH = cfg.hidden, cfg.heads
    hd = D // NH
    M = B * T
    dev = feats.device
    p_h = cfg.hidden_dropout if training else 0.0
    p_att = cfg.attention_dropout if training else 0.0
    p_conv = cfg.conv_dropout if training else 0.0
    S: dict = {"B": B, "T": T, "layers": [], "seed": seed, "p": (p_h, p_att, p_conv), "training": training}
    h = ops.dropout(feats, p_h, _e._mix(seed, 200), three=three)[0] if p_h > 0 else feats     # :674
    cos, sin = _rope(eng, P, T, hd)
    skip = set(skip_layers)
    for i in range(cfg.layers):
        if i in skip:
            S["layers"].append(None)
            continue
        L = Q + f"layers.{i}."
        st: dict = {}
        # ---- 1/2 feed-forward 1 (:597-601) ----
        h = _ffn_forward(eng, P, L + "ffn1.", L + "ffn1_layer_norm", h, M, training, seed, i, 1, st, "ffn1")
        # ---- self-attention with rotary embedding (:604-612) ----
        sa = L + "self_attn."
        y_hi, y_lo, r_hi, r_lo, m, r = ops.layernorm_rope_fwd(h, P[L + "self_attn_layer_norm.weight"],
                                                              P[L + "self_attn_layer_norm.bias"], BLOCK_LN_EPS, cos, sin, T, hd,
                                                              three=three)
        wqk = eng._w_cat(P, sa + "qk", [sa + "linear_q.weight", sa + "linear_k.weight"])
        bqk = eng._cat_vec(P, [sa + "linear_q.bias", sa + "linear_k.bias"])
        qkv = torch.empty(M, 3 * D, dtype=torch.float32, device=dev)
        ops.gemm(ops.kmajor(r_hi, r_lo), ops.kmajor(wqk[0], wqk[1]), m_per_batch=M, n=2 * D, k=D, out_f32=qkv, ldc=3 * D,
                 bias=bqk)
        wv = eng._w(P, sa + "linear_v.weight")
        ops.gemm(ops.kmajor(y_hi, y_lo), ops.kmajor(wv[0], wv[1]), m_per_batch=M, n=D, k=D, out_f32=qkv[:, 2 * D:],
                 ldc=3 * D, bias=P[sa + "linear_v.bias"])
        _, c_hi, c_lo, probs = ops.attention_fwd(qkv, B, T, NH, hd, three=three, save_probs=training, drop_p=p_att,
                                                 seed=_seed(seed, i, 3))
        h_in = h
        h, _, _, _ = eng._linear_fwd(c_hi, c_lo, eng._w(P, sa + "linear_out.weight"), P[sa + "linear_out.bias"], M, D, D,
                                     residual=h_in, drop_p=p_att, drop_seed=_seed(seed, i, 4))      # self_attn_dropout (:583)
        st["attn"] = dict(x=h_in, ln=(m, r), y_hi=y_hi, y_lo=y_lo, r_hi=r_hi, r_lo=r_lo, qkv=qkv, probs=probs, c_hi=c_hi,
                          c_lo=c_lo)
        # ---- convolution module (:615-618, :398-417) ----
        cm = L + "conv_module."
        _, n_hi, n_lo, m, r = ops.layernorm_fwd(h, P[cm + "layer_norm.weight"], P[cm + "layer_norm.bias"], BLOCK_LN_EPS,
                                                three=three, want_f32=False)
        g, _, _, _ = eng._linear_fwd(n_hi, n_lo, _w2d(eng, P, cm + "pointwise_conv1.weight"), None, M, 2 * D, D)
        dwo, stats = ops.glu_dwconv_fwd(g, P[cm + "depthwise_conv.weight"], B, T, want_stats=training)
        s_hi, s_lo, bn_mean, bn_rstd = ops.bn_silu_fwd(dwo, stats, P[cm + "batch_norm.weight"], P[cm + "batch_norm.bias"], BN_EPS,
                                                       BN_MOMENTUM, P[cm + "batch_norm.running_mean"],
                                                       P[cm + "batch_norm.running_var"],
                                                       P.get(cm + "batch_norm.num_batches_tracked"), training, three=three)
        h_in = h
        h, _, _, _ = eng._linear_fwd(s_hi, s_lo, _w2d(eng, P, cm + "pointwise_conv2.weight"), None, M, D, D, residual=h_in,
                                     drop_p=p_conv, drop_seed=_seed(seed, i, 5))
        st["conv"] = dict(x=h_in, ln=(m, r), n_hi=n_hi, n_lo=n_lo, g=g, dwo=dwo, bn=(bn_mean, bn_rstd), s_hi=s_hi, s_lo=s_lo)
        # ---- 1/2 feed-forward 2 (:621-626) and the block's final LayerNorm (:627) ----
        h = _ffn_forward(eng, P, L + "ffn2.", L + "ffn2_layer_norm", h, M, training, seed, i, 6, st, "ffn2")
        h_in = h
        h, _, _, m, r = ops.layernorm_fwd(h_in, P[L + "final_layer_norm.weight"], P[L + "final_layer_norm.bias"], BLOCK_LN_EPS,
                                          three=three, want_split=False)
        st["final"] = dict(x=h_in, ln=(m, r))
        S["layers"].append(st)
    h_in = h
    h, _, _, m, r = ops.layernorm_fwd(h_in, P[Q + "layer_norm.weight"], P[Q + "layer_norm.bias"], cfg.ln_eps, three=three,
                                      want_split=False)                                              # :707
    S["enc_ln"] = dict(x=h_in, ln=(m, r))
    return h, S


def _v3(w):
    """[out, in] weight gradient seen as the Conv1d weight [out, in, 1] (None in frozen-w2v mode)."""
    return None if w is None else w.view(w.shape[0], w.shape[1], 1)


def _zeros(n: int, dev) -> Tensor:
    return _e.ops.zeros(n, device=dev)


def _gz(eng, name: str, n: int, dev) -> Tensor:
    """zero-initialised gradient accumulator of parameter `name` (its arena slot when one is attached)"""
    return eng._gout(name, (n,), dev, zero=True)


def _ln_bwd(eng, P, G, name: str, dy: Tensor, x: Tensor, ln, **kw):
    """LayerNorm backward; parameter gradients go to G[name.weight / name.bias]."""
    dev = dy.device
    n = x.shape[1]
    G[name + ".weight"], G[name + ".bias"] = _gz(eng, name + ".weight", n, dev), _gz(eng, name + ".bias", n, dev)
    return _e.ops.layernorm_bwd(dy, x, P[name + ".weight"], ln[0], ln[1], G[name + ".weight"], G[name + ".bias"],
                                three=eng.three, **kw)


def _ffn_backward(eng, P, G, pfx: str, ln: str, st: dict, g_hi, g_lo, db_out: Tensor, d_res: Tensor, M: int, training: bool,
                  seed: int, layer: int, site: int, **ln_kw):
    """Backward of out = x + 1/2 Dropout(FFN(LN(x))).  g = (mask * d out) as bf16 operand, db_out = its column
    sums, d_res = d out (fp32).  Returns layernorm_bwd's outputs for x (d x includes the residual path)."""
    ops, cfg = _e.ops, eng.w2v
    D, F = cfg.hidden, cfg.ffn
    p_act = cfg.activation_dropout if training else 0.0
    G[pfx + "output_dense.bias"] = ops.axpby(db_out, None, alpha=0.5, out=db_out)
    G[pfx + "output_dense.weight"] = eng._wgrad(g_hi, g_lo, st["f_hi"], st["f_lo"], D, F, M, alpha=0.5,
                                                name=pfx + "output_dense.weight")
    db1 = _gz(eng, pfx + "intermediate_dense.bias", F, g_hi.device)
    _, dp_hi, dp_lo = eng._dgrad(g_hi, g_lo, eng._w(P, pfx + "output_dense.weight"), M, D, F, want_f32=False, want_split=True,
                                 aux=st["pre"], aux_mode="mul", colsum=db1, alpha=0.5)
    G[pfx + "intermediate_dense.bias"] = db1
    G[pfx + "intermediate_dense.weight"] = eng._wgrad(dp_hi, dp_lo, st["u_hi"], st["u_lo"], F, D, M,
                                                      name=pfx + "intermediate_dense.weight")
    du, _, _ = eng._dgrad(dp_hi, dp_lo, eng._w(P, pfx + "intermediate_dense.weight"), M, F, D)
    return _ln_bwd(eng, P, G, ln, du, st["x"], st["ln"], dres=d_res, **ln_kw)


def backward(eng, P, S: dict, dh: Tensor, G: Dict[str, Tensor]) -> Tensor:
    """dh = gradient w.r.t. the encoder output [M, D]; fills G, returns the gradient w.r.t. feats."""
    ops, cfg, three = _e.ops, eng.w2v, eng.three
    D, NH = cfg.hidden, cfg.heads
    hd = D // NH
    B, T = S["B"], S["T"]
    M = B * T
    dev = dh.device
    seed, training = S["seed"], S["training"]
    p_h, p_att, p_conv = S["p"]
    cos, sin = _rope(eng, P, T, hd)
    dh, _, _ = _ln_bwd(eng, P, G, Q + "layer_norm", dh, S["enc_ln"]["x"], S["enc_ln"]["ln"], want_split=False)
    ds_ws = torch.empty(B, NH, T, T, dtype=torch.float32, device=dev)
    for i in reversed(range(cfg.layers)):
        eng.emit_grads(G)   # gradients of the layers above are complete: start their all-reduce (data parallel)
        st = S["layers"][i]
        if st is None:
            continue
        L = Q + f"layers.{i}."
        # ---- final LayerNorm; its masked copy is the operand of the ffn2 output GEMMs ----
        db = _gz(eng, L + "ffn2.output_dense.bias", D, dev)
        d4, g_hi, g_lo = _ln_bwd(eng, P, G, L + "final_layer_norm", dh, st["final"]["x"], st["final"]["ln"], dcol=db,
                                 br_drop_p=p_h, br_seed=_seed(seed, i, 7))
        # ---- 1/2 ffn2: x = conv output h3; the masked copy of d h3 feeds pointwise_conv2's backward ----
        d3, g_hi, g_lo = _ffn_backward(eng, P, G, L + "ffn2.", L + "ffn2_layer_norm", st["ffn2"], g_hi, g_lo, db, d4, M, training,
                                       seed, i, 6, br_drop_p=p_conv, br_seed=_seed(seed, i, 5))
        # ---- convolution module ----
        cm = L + "conv_module."
        cs = st["conv"]
        w2 = _w2d(eng, P, cm + "pointwise_conv2.weight")
        G[cm + "pointwise_conv2.weight"] = _v3(eng._wgrad(g_hi, g_lo, cs["s_hi"], cs["s_lo"], D, D, M,
                                                          name=cm + "pointwise_conv2.weight"))
        dsil, _, _ = eng._dgrad(g_hi, g_lo, w2, M, D, D)
        gam, bet = P[cm + "batch_norm.weight"], P[cm + "batch_norm.bias"]
        bn_mean, bn_rstd = cs["bn"]
        dgam, dbet = ops.bn_silu_bwd_stats(dsil, cs["dwo"], bn_mean, bn_rstd, gam, bet,
                                           dgamma_out=_gz(eng, cm + "batch_norm.weight", D, dev),
                                           dbeta_out=_gz(eng, cm + "batch_norm.bias", D, dev))
        G[cm + "batch_norm.weight"], G[cm + "batch_norm.bias"] = dgam, dbet
        dg_hi, dg_lo, ddw = ops.glu_dwconv_bwd(dsil, cs["dwo"], cs["g"], P[cm + "depthwise_conv.weight"], bn_mean, bn_rstd, gam,
                                               bet, dgam, dbet, B, T, training, three=three,
                                               dw_out=eng._gout(cm + "depthwise_conv.weight", P[cm + "depthwise_conv.weight"].shape,
                                                                dev, zero=True))
        G[cm + "depthwise_conv.weight"] = ddw
        G[cm + "pointwise_conv1.weight"] = _v3(eng._wgrad(dg_hi, dg_lo, cs["n_hi"], cs["n_lo"], 2 * D, D, M,
                                                          name=cm + "pointwise_conv1.weight"))
        dn, _, _ = eng._dgrad(dg_hi, dg_lo, _w2d(eng, P, cm + "pointwise_conv1.weight"), M, 2 * D, D)
        dbo = _gz(eng, L + "self_attn.linear_out.bias", D, dev)
        d2, g_hi, g_lo = _ln_bwd(eng, P, G, cm + "layer_norm", dn, cs["x"], cs["ln"], dres=d3, dcol=dbo, br_drop_p=p_att,
                                 br_seed=_seed(seed, i, 4))
        # ---- self-attention ----
        sa = L + "self_attn."
        at = st["attn"]
        G[sa + "linear_out.bias"] = dbo
        G[sa + "linear_out.weight"] = eng._wgrad(g_hi, g_lo, at["c_hi"], at["c_lo"], D, D, M, name=sa + "linear_out.weight")
        dctx, _, _ = eng._dgrad(g_hi, g_lo, eng._w(P, sa + "linear_out.weight"), M, D, D)
        dqkv, dq_hi, dq_lo = ops.attention_bwd(at["qkv"], at["probs"], dctx, ds_ws, B, T, NH, hd, three=three, drop_p=p_att,
                                               seed=_seed(seed, i, 3))
        dbq = eng._gout_cat([sa + nm + ".bias" for nm in ("linear_q", "linear_k", "linear_v")], (3 * D,), dev, zero=True)
        ops.colsum(dqkv, dbq)
        for j, nm in enumerate(("linear_q", "linear_k", "linear_v")):
            G[sa + nm + ".bias"] = dbq[j * D:(j + 1) * D]
        qk_hi, qk_lo = dq_hi[:, :2 * D], (None if dq_lo is None else dq_lo[:, :2 * D])
        v_hi, v_lo = dq_hi[:, 2 * D:], (None if dq_lo is None else dq_lo[:, 2 * D:])
        dwqk = eng._wgrad(qk_hi, qk_lo, at["r_hi"], at["r_lo"], 2 * D, D, M, name=[sa + "linear_q.weight", sa + "linear_k.weight"])
        if dwqk is not None:
            G[sa + "linear_q.weight"], G[sa + "linear_k.weight"] = dwqk[:D], dwqk[D:]
        G[sa + "linear_v.weight"] = eng._wgrad(v_hi, v_lo, at["y_hi"], at["y_lo"], D, D, M, name=sa + "linear_v.weight")
        wqk = eng._w_cat(P, sa + "qk", [sa + "linear_q.weight", sa + "linear_k.weight"])
        du_rot, _, _ = eng._dgrad(qk_hi, qk_lo, wqk, M, 2 * D, D)
        du_v, _, _ = eng._dgrad(v_hi, v_lo, eng._w(P, sa + "linear_v.weight"), M, D, D)
        du = ops.rope_bwd_add(du_rot, du_v, cos, sin, T, hd)
        db = _gz(eng, L + "ffn1.output_dense.bias", D, dev)
        d1, g_hi, g_lo = _ln_bwd(eng, P, G, L + "self_attn_layer_norm", du, at["x"], at["ln"], dres=d2, dcol=db, br_drop_p=p_h,
                                 br_seed=_seed(seed, i, 2))
        # ---- 1/2 ffn1 ----
        dh, _, _ = _ffn_backward(eng, P, G, L + "ffn1.", L + "ffn1_layer_norm", st["ffn1"], g_hi, g_lo, db, d1, M, training, seed,
                                 i, 1, want_split=False)
    if p_h > 0:
        dh = ops.dropout(dh, p_h, _e._mix(seed, 200), three=three)[0]
    return dh
