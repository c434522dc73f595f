"""CPU oracle for the b2p2t_gru+w2v / b2p2t_gru+w2v_conformer hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package (`wav2vec2forbrain_b200/`) may
import this module; only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py` do, and only as the checker / the timed CPU baseline.

It is a plain-PyTorch (CPU, fp32 or fp64) restatement of what the reference computes on the
path named by BASELINE.json.  Each function cites the reference file:line it follows
(paths are relative to the reference checkout; `TF:` = transformers, the third-party
dependency the reference pins at 4.35.2 (`environment.yaml:291`) whose encoder arithmetic is
not vendored in the reference tree, so its published algorithm is restated here).

Parity pinning: the reference ships no tests or golden vectors for this path,
so the oracle is pinned against outputs of the reference itself: `oracle/gen_golden.py`
imports the live reference classes from /root/reference in the build container, runs them on
seeded inputs and writes `tests/golden/*.pt`; `tests/test_oracle_golden.py` replays the
oracle against those files.  All weights come from `make_state_dict` (a seeded generator
with the reference's state_dict key names) so the same model can be rebuilt on the GPU box,
where /root/reference does not exist.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

N_DAYS = 24          # b2p2t_model.py:115
NEURAL_DIM = 256     # b2p2t_model.py:116
SMOOTH_TAPS = 20     # b2p2t_model.py:118


# ---------------------------------------------------------------------------------------
# architecture descriptions
# ---------------------------------------------------------------------------------------
@dataclass
class BrainEncoderCfg:
    """Flags of B2P2TBrainFeatureExtractorArgsModel (brain_feature_extractor.py:13-21,
    b2p2t_model.py:20-24)."""
    gru_hidden: int = 512
    gru_layers: int = 3
    bidirectional: bool = True
    fc_hidden: Sequence[int] = (256,)
    fc_act: str = "gelu"
    out_size: int = 768                 # PRETRAINED_LATENT_SIZES[checkpoint], base_args.py:5-11
    unfold_len: int = 32
    unfold_stride: int = 4
    smooth_sigma: float = 1.5
    dropout: float = 0.0
    learnable_h0: bool = False


@dataclass
class W2VCfg:
    """Subset of transformers' Wav2Vec2Config / Wav2Vec2ConformerConfig that the path reads."""
    kind: str = "base"                  # "base" (wav2vec2) | "conformer"
    hidden: int = 768
    layers: int = 12
    heads: int = 12
    ffn: int = 3072
    vocab: int = 32
    ln_eps: float = 1e-5
    pos_conv_k: int = 128
    pos_conv_groups: int = 16
    stable_ln: bool = False
    # conformer only
    dw_kernel: int = 31
    rope_base: int = 10000
    # dropouts (parity runs set all of them to 0)
    hidden_dropout: float = 0.0
    attention_dropout: float = 0.0
    activation_dropout: float = 0.0
    final_dropout: float = 0.0
    conv_dropout: float = 0.0
    layerdrop: float = 0.0


def base_cfg() -> W2VCfg:
    """facebook/wav2vec2-base-960h architecture."""
    return W2VCfg(kind="base", hidden=768, layers=12, heads=12, ffn=3072)


def conformer_large_cfg() -> W2VCfg:
    """facebook/wav2vec2-conformer-rope-large-960h-ft architecture."""
    return W2VCfg(kind="conformer", hidden=1024, layers=24, heads=16, ffn=4096)


# ---------------------------------------------------------------------------------------
# seeded weights with the reference's state_dict keys
# ---------------------------------------------------------------------------------------
def gaussian_taps(sigma: float, size: int = SMOOTH_TAPS, dtype=torch.float32) -> Tensor:
    """GaussianSmoothing.__init__ (b2p2t_model.py:41-68), dim=1: normalised Gaussian."""
    grid = torch.arange(size, dtype=torch.float32)
    mean = (size - 1) / 2
    k = 1 / (sigma * math.sqrt(2 * math.pi)) * torch.exp(-(((grid - mean) / sigma) ** 2) / 2)
    k = k / torch.sum(k)
    return k.to(dtype)


def make_state_dict(be: BrainEncoderCfg, w2v: W2VCfg, seed: int = 42,
                    include_unused: bool = True) -> Dict[str, Tensor]:
    """Deterministic non-trivial weights under the reference's parameter names.

    Unlike the library initialisers (zero biases, unit LayerNorm gains) every tensor gets
    random content so that a transposed bias or a dropped gain cannot hide.
    """
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}

    def rn(*shape, std=1.0):
        return torch.randn(*shape, generator=g) * std

    def un(*shape, bound=1.0):
        return (torch.rand(*shape, generator=g) * 2 - 1) * bound

    # --- brain encoder (b2p2t_model.py:114-136, brain_feature_extractor.py:24-54)
    p = "brain_encoder."
    sd[p + "day_weights"] = torch.eye(NEURAL_DIM).repeat(N_DAYS, 1, 1) + rn(N_DAYS, NEURAL_DIM, NEURAL_DIM, std=0.02)
    sd[p + "day_bias"] = rn(N_DAYS, 1, NEURAL_DIM, std=0.05)
    sd[p + "gaussian_smoother.weight"] = gaussian_taps(be.smooth_sigma).view(1, 1, -1).repeat(NEURAL_DIM, 1, 1)
    if include_unused:
        for d in range(N_DAYS):
            sd[p + f"inpLayer{d}.weight"] = torch.eye(NEURAL_DIM) + rn(NEURAL_DIM, NEURAL_DIM, std=0.02)
            sd[p + f"inpLayer{d}.bias"] = rn(NEURAL_DIM, std=0.02)
    e = p + "neural_decoder.encoder."
    ndir = 2 if be.bidirectional else 1
    H = be.gru_hidden
    sd[e + "hidden_start"] = rn(ndir * be.gru_layers, H)
    bound = 1.0 / math.sqrt(H)
    for layer in range(be.gru_layers):
        in_size = NEURAL_DIM * be.unfold_len if layer == 0 else H * ndir
        for sfx in ([""] + (["_reverse"] if be.bidirectional else [])):
            sd[e + f"gru.weight_ih_l{layer}{sfx}"] = un(3 * H, in_size, bound=bound)
            sd[e + f"gru.weight_hh_l{layer}{sfx}"] = un(3 * H, H, bound=bound)
            sd[e + f"gru.bias_ih_l{layer}{sfx}"] = un(3 * H, bound=bound)
            sd[e + f"gru.bias_hh_l{layer}{sfx}"] = un(3 * H, bound=bound)
    sizes = [H * ndir] + list(be.fc_hidden) + [be.out_size]
    for i in range(len(sizes) - 1):
        idx = 2 * i  # nn.Sequential(Linear, act, Linear, ...) — nn_helper.py:39-49
        b = 1.0 / math.sqrt(sizes[i])
        sd[e + f"fc.{idx}.weight"] = un(sizes[i + 1], sizes[i], bound=b)
        sd[e + f"fc.{idx}.bias"] = un(sizes[i + 1], bound=b)

    # --- w2v encoder
    Hd = w2v.hidden

    def lin(prefix, out_f, in_f, bias=True, std=0.02):
        sd[prefix + ".weight"] = rn(out_f, in_f, std=std)
        if bias:
            sd[prefix + ".bias"] = rn(out_f, std=0.02)

    def ln(prefix, n):
        sd[prefix + ".weight"] = 1.0 + rn(n, std=0.05)
        sd[prefix + ".bias"] = rn(n, std=0.05)

    if w2v.kind == "base":
        q = "w2v_encoder.wav2vec2.encoder."
        cin = Hd // w2v.pos_conv_groups
        sd[q + "pos_conv_embed.conv.bias"] = rn(Hd, std=0.02)
        v = rn(Hd, cin, w2v.pos_conv_k, std=math.sqrt(4.0 / (w2v.pos_conv_k * Hd)))
        sd[q + "pos_conv_embed.conv.parametrizations.weight.original1"] = v
        sd[q + "pos_conv_embed.conv.parametrizations.weight.original0"] = \
            v.norm(dim=(0, 1), keepdim=True) * (1.0 + rn(1, 1, w2v.pos_conv_k, std=0.05))
        ln(q + "layer_norm", Hd)
        for i in range(w2v.layers):
            L = q + f"layers.{i}."
            for nm in ("k_proj", "v_proj", "q_proj", "out_proj"):
                lin(L + "attention." + nm, Hd, Hd, std=0.03)
            ln(L + "layer_norm", Hd)
            lin(L + "feed_forward.intermediate_dense", w2v.ffn, Hd, std=0.03)
            lin(L + "feed_forward.output_dense", Hd, w2v.ffn, std=0.03)
            ln(L + "final_layer_norm", Hd)
        lin("w2v_encoder.lm_head", w2v.vocab, Hd, std=0.05)
    elif w2v.kind == "conformer":
        q = "w2v_encoder.wav2vec2_conformer.encoder."
        hd = Hd // w2v.heads
        sd[q + "embed_positions.inv_freq"] = 1.0 / (w2v.rope_base ** (torch.arange(0, hd, 2, dtype=torch.int64).float() / hd))
        if include_unused:  # constructed but never called (TF:wav2vec2_conformer…:645 vs :651-716)
            cin = Hd // w2v.pos_conv_groups
            sd[q + "pos_conv_embed.conv.bias"] = rn(Hd, std=0.02)
            v = rn(Hd, cin, w2v.pos_conv_k, std=0.01)
            sd[q + "pos_conv_embed.conv.parametrizations.weight.original1"] = v
            sd[q + "pos_conv_embed.conv.parametrizations.weight.original0"] = v.norm(dim=(0, 1), keepdim=True)
        ln(q + "layer_norm", Hd)
        for i in range(w2v.layers):
            L = q + f"layers.{i}."
            ln(L + "ffn1_layer_norm", Hd)
            lin(L + "ffn1.intermediate_dense", w2v.ffn, Hd, std=0.03)
            lin(L + "ffn1.output_dense", Hd, w2v.ffn, std=0.03)
            ln(L + "self_attn_layer_norm", Hd)
            for nm in ("linear_q", "linear_k", "linear_v", "linear_out"):
                lin(L + "self_attn." + nm, Hd, Hd, std=0.03)
            ln(L + "conv_module.layer_norm", Hd)
            sd[L + "conv_module.pointwise_conv1.weight"] = rn(2 * Hd, Hd, 1, std=0.03)
            sd[L + "conv_module.depthwise_conv.weight"] = rn(Hd, 1, w2v.dw_kernel, std=0.15)
            sd[L + "conv_module.batch_norm.weight"] = 1.0 + rn(Hd, std=0.05)
            sd[L + "conv_module.batch_norm.bias"] = rn(Hd, std=0.05)
            sd[L + "conv_module.batch_norm.running_mean"] = rn(Hd, std=0.05)
            sd[L + "conv_module.batch_norm.running_var"] = 1.0 + 0.1 * torch.rand(Hd, generator=g)
            sd[L + "conv_module.batch_norm.num_batches_tracked"] = torch.tensor(0, dtype=torch.int64)
            sd[L + "conv_module.pointwise_conv2.weight"] = rn(Hd, Hd, 1, std=0.03)
            ln(L + "ffn2_layer_norm", Hd)
            lin(L + "ffn2.intermediate_dense", w2v.ffn, Hd, std=0.03)
            lin(L + "ffn2.output_dense", Hd, w2v.ffn, std=0.03)
            ln(L + "final_layer_norm", Hd)
        lin("w2v_encoder.lm_head", w2v.vocab, Hd, std=0.05)
    else:
        raise ValueError(w2v.kind)
    return sd


def make_batch(B: int, T: int, S: int = 40, seed: int = 1234, ragged: bool = False):
    """Synthetic batch of SURVEY.md §8(d): input ~N(0,1) [B,T,256], day ids, CTC targets 4..31."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, T, NEURAL_DIM, generator=g)
    day = torch.randint(0, N_DAYS, (B,), generator=g)
    if ragged:
        in_lens = torch.linspace(T, int(T * 0.72), B).long()
        tgt_lens = torch.randint(max(1, S // 2), S + 1, (B,), generator=g)
        tgt_lens[0] = S
    else:
        in_lens = torch.full((B,), T, dtype=torch.long)
        tgt_lens = torch.full((B,), S, dtype=torch.long)
    tgt = torch.randint(4, 32, (B, S), generator=g)
    for b in range(B):
        x[b, in_lens[b]:] = 0
        tgt[b, tgt_lens[b]:] = 0
    return {"input": x, "target": tgt, "day_idxs": day, "input_lens": in_lens, "target_lens": tgt_lens}


# ---------------------------------------------------------------------------------------
# stage 1: smooth -> day affine -> softsign -> unfold      (b2p2t_model.py:138-177)
# ---------------------------------------------------------------------------------------
def smooth(x: Tensor, taps: Tensor) -> Tensor:
    """GaussianSmoothing.forward (b2p2t_model.py:82-90): depthwise cross-correlation over time,
    `padding="same"` for an even kernel = 9 zeros left, 10 right."""
    B, T, C = x.shape
    K = taps.numel()
    left = (K - 1) // 2
    xp = F.pad(x, (0, 0, left, K - 1 - left))
    y = torch.zeros_like(x)
    for k in range(K):
        y = y + taps[k] * xp[:, k:k + T, :]
    return y


def day_affine_softsign(y: Tensor, day_idxs: Tensor, day_w: Tensor, day_b: Tensor) -> Tensor:
    """b2p2t_model.py:154-159: einsum('btd,bdk->btk', y, W[day]) + b[day], then Softsign."""
    z = torch.bmm(y, day_w[day_idxs]) + day_b[day_idxs].unsqueeze(1)
    return z / (1 + z.abs())


def unfold(s: Tensor, klen: int, stride: int) -> Tensor:
    """b2p2t_model.py:108-113,161-167: nn.Unfold((klen,1), stride) -> feature index c*klen+k."""
    B, T, C = s.shape
    Tp = (T - klen) // stride + 1
    idx = (torch.arange(Tp)[:, None] * stride + torch.arange(klen)[None, :])   # [Tp, klen]
    u = s[:, idx, :]                       # [B, Tp, klen, C]
    return u.permute(0, 1, 3, 2).reshape(B, Tp, C * klen)


def processed_lens(input_lens: Tensor, klen: int, stride: int) -> Tensor:
    """b2p2t_model.py:169-174: ((len - klen) / stride).to(int32) — truncation, = T' - 1."""
    return ((input_lens - klen) / stride).to(torch.int32)


# ---------------------------------------------------------------------------------------
# stage 2: GRU + FC                       (brain_feature_extractor.py:39-67, nn_helper.py:31-49)
# ---------------------------------------------------------------------------------------
def gru_direction(x: Tensor, w_ih, w_hh, b_ih, b_hh, h0: Optional[Tensor], reverse: bool) -> Tensor:
    """torch.nn.GRU equations, gates ordered r,z,n."""
    B, T, _ = x.shape
    H = w_hh.shape[1]
    h = x.new_zeros(B, H) if h0 is None else h0
    gi_all = x @ w_ih.t() + b_ih
    outs: List[Optional[Tensor]] = [None] * T
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        gi = gi_all[:, t]
        gh = h @ w_hh.t() + b_hh
        r = torch.sigmoid(gi[:, :H] + gh[:, :H])
        z = torch.sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
        n = torch.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
        h = (1 - z) * n + z * h
        outs[t] = h
    return torch.stack(outs, dim=1)


def gru_stack(x: Tensor, sd: Dict[str, Tensor], prefix: str, cfg: BrainEncoderCfg,
              training: bool = False, impl: str = "loop") -> Tensor:
    """brain_feature_extractor.py:61-65.  `impl="aten"` calls torch's own CPU GRU (what the
    reference executes) — used for speed by the CPU baseline and cross-checked against the
    explicit loop in tests."""
    dirs = [""] + (["_reverse"] if cfg.bidirectional else [])
    B = x.shape[0]
    if impl == "aten":
        flat = []
        for layer in range(cfg.gru_layers):
            for sfx in dirs:
                for nm in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"):
                    flat.append(sd[f"{prefix}gru.{nm}_l{layer}{sfx}"])
        ndir = len(dirs)
        if cfg.learnable_h0:
            h0 = sd[prefix + "hidden_start"].unsqueeze(1).repeat(1, B, 1)
        else:
            h0 = x.new_zeros(cfg.gru_layers * ndir, B, cfg.gru_hidden)
        out, _ = torch._VF.gru(x, h0, flat, True, cfg.gru_layers, cfg.dropout if training else 0.0,
                               training, cfg.bidirectional, True)
        return out
    h = x
    for layer in range(cfg.gru_layers):
        outs = []
        for di, sfx in enumerate(dirs):
            h0 = None
            if cfg.learnable_h0:
                h0 = sd[prefix + "hidden_start"][layer * len(dirs) + di].unsqueeze(0).expand(B, -1)
            outs.append(gru_direction(
                h, sd[f"{prefix}gru.weight_ih_l{layer}{sfx}"], sd[f"{prefix}gru.weight_hh_l{layer}{sfx}"],
                sd[f"{prefix}gru.bias_ih_l{layer}{sfx}"], sd[f"{prefix}gru.bias_hh_l{layer}{sfx}"],
                h0, reverse=(di == 1)))
        h = torch.cat(outs, dim=-1)
        if training and cfg.dropout > 0 and layer < cfg.gru_layers - 1:
            h = F.dropout(h, cfg.dropout, True)
    return h


def gelu(x: Tensor) -> Tensor:
    """transformers GELUActivation = exact erf GELU (nn_helper.py:48 ACT2FN['gelu'])."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def fc_stack(h: Tensor, sd: Dict[str, Tensor], prefix: str, cfg: BrainEncoderCfg) -> Tensor:
    """nn_helper.py:31-49: Linear, act, ..., Linear (no activation after the last)."""
    n = len(cfg.fc_hidden) + 1
    for i in range(n):
        h = h @ sd[f"{prefix}fc.{2 * i}.weight"].t() + sd[f"{prefix}fc.{2 * i}.bias"]
        if i < n - 1:
            assert cfg.fc_act == "gelu"
            h = gelu(h)
    return h


def brain_encoder(batch: dict, sd: Dict[str, Tensor], cfg: BrainEncoderCfg, training=False,
                  gru_impl="loop", stages: Optional[dict] = None):
    """B2P2TModel.forward (b2p2t_model.py:138-177) -> (features [B,T',out], logit_lens)."""
    p = "brain_encoder."
    taps = sd[p + "gaussian_smoother.weight"][0, 0]
    y = smooth(batch["input"], taps)
    s = day_affine_softsign(y, batch["day_idxs"], sd[p + "day_weights"], sd[p + "day_bias"][:, 0])
    u = unfold(s, cfg.unfold_len, cfg.unfold_stride)
    e = p + "neural_decoder.encoder."
    g = gru_stack(u, sd, e, cfg, training, gru_impl)
    f = fc_stack(g, sd, e, cfg)
    lens = processed_lens(batch["input_lens"], cfg.unfold_len, cfg.unfold_stride) \
        if batch.get("input_lens") is not None else None
    if stages is not None:
        stages.update(smooth=y, softsign=s, gru=g, features=f)
    return f, lens


# ---------------------------------------------------------------------------------------
# stage 3a: wav2vec2 encoder (TF:wav2vec2/modeling_wav2vec2.py)
# ---------------------------------------------------------------------------------------
def layer_norm(x: Tensor, w: Tensor, b: Tensor, eps: float) -> Tensor:
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * w + b


def weight_norm_dim2(g: Tensor, v: Tensor) -> Tensor:
    """torch weight_norm(dim=2): w = g * v / ||v|| with the norm over dims (0,1) per tap
    (TF:wav2vec2/…:326-368)."""
    return v * (g / v.norm(dim=(0, 1), keepdim=True))


def pos_conv_embed(x: Tensor, sd, q: str, cfg: W2VCfg) -> Tensor:
    """Wav2Vec2PositionalConvEmbedding.forward: grouped conv1d (pad k//2), drop the last
    frame for even k (Wav2Vec2SamePadLayer), GELU."""
    pre = q + "pos_conv_embed.conv."
    if pre + "parametrizations.weight.original0" in sd:
        w = weight_norm_dim2(sd[pre + "parametrizations.weight.original0"],
                             sd[pre + "parametrizations.weight.original1"])
    else:  # torch<2.1 spelling
        w = weight_norm_dim2(sd[pre + "weight_g"], sd[pre + "weight_v"])
    k = cfg.pos_conv_k
    y = F.conv1d(x.transpose(1, 2), w, sd[pre + "bias"], padding=k // 2, groups=cfg.pos_conv_groups)
    if k % 2 == 0:
        y = y[:, :, :-1]
    return gelu(y).transpose(1, 2)


def mha(h: Tensor, sd, a: str, heads: int, names=("q_proj", "k_proj", "v_proj", "out_proj"),
        q_in: Optional[Tensor] = None, p_drop: float = 0.0, training=False) -> Tensor:
    """Wav2Vec2Attention.forward, eager, no mask (TF:wav2vec2/…:466-549).  `q_in` lets the
    conformer feed the rotated state to Q and K only."""
    B, T, Hd = h.shape
    hd = Hd // heads
    qk_src = h if q_in is None else q_in
    q = qk_src @ sd[a + names[0] + ".weight"].t() + sd[a + names[0] + ".bias"]
    k = qk_src @ sd[a + names[1] + ".weight"].t() + sd[a + names[1] + ".bias"]
    v = h @ sd[a + names[2] + ".weight"].t() + sd[a + names[2] + ".bias"]
    q = q.view(B, T, heads, hd).transpose(1, 2)
    k = k.view(B, T, heads, hd).transpose(1, 2)
    v = v.view(B, T, heads, hd).transpose(1, 2)
    s = (q @ k.transpose(2, 3)) * (hd ** -0.5)
    p = torch.softmax(s, dim=-1)
    if training and p_drop > 0:
        p = F.dropout(p, p_drop, True)
    o = (p @ v).transpose(1, 2).reshape(B, T, Hd)
    return o @ sd[a + names[3] + ".weight"].t() + sd[a + names[3] + ".bias"]


def _drop(x, p, training):
    return F.dropout(x, p, True) if (training and p > 0) else x


def w2v_base_encoder(x: Tensor, sd, cfg: W2VCfg, training=False,
                     skip_layers: Optional[Sequence[bool]] = None, stages: Optional[dict] = None) -> Tensor:
    """Wav2Vec2Encoder.forward (post-LN; TF:wav2vec2/…:668-727) or the StableLayerNorm variant
    (:730-803).  LayerDrop decisions come in through `skip_layers` (the reference draws
    `torch.rand([])` per layer on the CPU generator)."""
    q = "w2v_encoder.wav2vec2.encoder."
    h = x + pos_conv_embed(x, sd, q, cfg)
    if not cfg.stable_ln:
        h = layer_norm(h, sd[q + "layer_norm.weight"], sd[q + "layer_norm.bias"], cfg.ln_eps)
    h = _drop(h, cfg.hidden_dropout, training)
    if stages is not None:
        stages["enc_in"] = h
    for i in range(cfg.layers):
        if skip_layers is not None and skip_layers[i]:
            continue
        L = q + f"layers.{i}."
        if not cfg.stable_ln:
            a = mha(h, sd, L + "attention.", cfg.heads, p_drop=cfg.attention_dropout, training=training)
            h = layer_norm(h + _drop(a, cfg.hidden_dropout, training),
                           sd[L + "layer_norm.weight"], sd[L + "layer_norm.bias"], cfg.ln_eps)
            f = gelu(h @ sd[L + "feed_forward.intermediate_dense.weight"].t()
                     + sd[L + "feed_forward.intermediate_dense.bias"])
            f = _drop(f, cfg.activation_dropout, training)
            f = f @ sd[L + "feed_forward.output_dense.weight"].t() + sd[L + "feed_forward.output_dense.bias"]
            h = layer_norm(h + _drop(f, cfg.hidden_dropout, training),
                           sd[L + "final_layer_norm.weight"], sd[L + "final_layer_norm.bias"], cfg.ln_eps)
        else:
            a = mha(layer_norm(h, sd[L + "layer_norm.weight"], sd[L + "layer_norm.bias"], cfg.ln_eps),
                    sd, L + "attention.", cfg.heads, p_drop=cfg.attention_dropout, training=training)
            h = h + _drop(a, cfg.hidden_dropout, training)
            u = layer_norm(h, sd[L + "final_layer_norm.weight"], sd[L + "final_layer_norm.bias"], cfg.ln_eps)
            f = gelu(u @ sd[L + "feed_forward.intermediate_dense.weight"].t()
                     + sd[L + "feed_forward.intermediate_dense.bias"])
            f = _drop(f, cfg.activation_dropout, training)
            f = f @ sd[L + "feed_forward.output_dense.weight"].t() + sd[L + "feed_forward.output_dense.bias"]
            h = h + _drop(f, cfg.hidden_dropout, training)
        if stages is not None:
            stages[f"layer{i}"] = h
    if cfg.stable_ln:
        h = layer_norm(h, sd[q + "layer_norm.weight"], sd[q + "layer_norm.bias"], cfg.ln_eps)
    return h


# ---------------------------------------------------------------------------------------
# stage 3b: wav2vec2-conformer encoder (TF:wav2vec2_conformer/…)
# ---------------------------------------------------------------------------------------
def rope_tables(T: int, inv_freq: Tensor, dtype) -> tuple:
    """Wav2Vec2ConformerRotaryPositionalEmbedding.forward (TF:…conformer.py:125-156)."""
    t = torch.arange(T).type_as(inv_freq)
    freqs = torch.einsum("i,j->ij", t, inv_freq)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def apply_rope(u: Tensor, cos: Tensor, sin: Tensor, heads: int) -> Tensor:
    """_apply_rotary_embedding (TF:…conformer.py:509-527): per 64-wide head slice,
    u*cos + rotate_half(u)*sin with rotate_half = cat(-u[half:], u[:half])."""
    B, T, Hd = u.shape
    hd = Hd // heads
    v = u.view(B, T, heads, hd)
    half = hd // 2
    rot = torch.cat((-v[..., half:], v[..., :half]), dim=-1)
    out = v * cos[None, :, None, :] + rot * sin[None, :, None, :]
    return out.reshape(B, T, Hd)


def silu(x):
    return x * torch.sigmoid(x)


def conformer_ffn(h, sd, p, cfg, training):
    """Wav2Vec2ConformerFeedForward (TF:…conformer.py:336-357), hidden_act='swish'."""
    f = silu(h @ sd[p + "intermediate_dense.weight"].t() + sd[p + "intermediate_dense.bias"])
    f = _drop(f, cfg.activation_dropout, training)
    f = f @ sd[p + "output_dense.weight"].t() + sd[p + "output_dense.bias"]
    return _drop(f, cfg.hidden_dropout, training)


def conformer_conv_module(h, sd, p, cfg: W2VCfg, training, bn_stats_out: Optional[dict] = None):
    """Wav2Vec2ConformerConvolutionModule.forward (TF:…conformer.py:398-417): LN -> pw1 ->
    GLU -> depthwise k=31 pad 15 -> BatchNorm1d -> SiLU -> pw2 -> dropout.  BatchNorm in
    training mode normalises with biased batch statistics over all B*T' positions."""
    c = layer_norm(h, sd[p + "layer_norm.weight"], sd[p + "layer_norm.bias"], 1e-5)
    c = c @ sd[p + "pointwise_conv1.weight"][:, :, 0].t()
    Hd = cfg.hidden
    c = c[..., :Hd] * torch.sigmoid(c[..., Hd:])
    k = cfg.dw_kernel
    c = F.conv1d(c.transpose(1, 2), sd[p + "depthwise_conv.weight"], None, padding=(k - 1) // 2, groups=Hd)
    c = c.transpose(1, 2)
    if training:
        mean = c.mean(dim=(0, 1))
        var = ((c - mean) ** 2).mean(dim=(0, 1))
        if bn_stats_out is not None:
            n = c.shape[0] * c.shape[1]
            bn_stats_out[p] = (mean.detach(), (var * n / max(n - 1, 1)).detach())
    else:
        mean, var = sd[p + "batch_norm.running_mean"], sd[p + "batch_norm.running_var"]
    c = (c - mean) / torch.sqrt(var + 1e-5) * sd[p + "batch_norm.weight"] + sd[p + "batch_norm.bias"]
    c = silu(c)
    c = c @ sd[p + "pointwise_conv2.weight"][:, :, 0].t()
    return _drop(c, cfg.conv_dropout, training)


def conformer_encoder(x: Tensor, sd, cfg: W2VCfg, training=False,
                      skip_layers: Optional[Sequence[bool]] = None, stages: Optional[dict] = None,
                      bn_stats_out: Optional[dict] = None) -> Tensor:
    """Wav2Vec2ConformerEncoder.forward (TF:…conformer.py:651-716), rotary positions,
    block = Wav2Vec2ConformerEncoderLayer.forward (:593-631)."""
    q = "w2v_encoder.wav2vec2_conformer.encoder."
    h = _drop(x, cfg.hidden_dropout, training)
    cos, sin = rope_tables(x.shape[1], sd[q + "embed_positions.inv_freq"], x.dtype)
    for i in range(cfg.layers):
        if skip_layers is not None and skip_layers[i]:
            continue
        L = q + f"layers.{i}."
        u = layer_norm(h, sd[L + "ffn1_layer_norm.weight"], sd[L + "ffn1_layer_norm.bias"], 1e-5)
        h = h + 0.5 * conformer_ffn(u, sd, L + "ffn1.", cfg, training)
        u = layer_norm(h, sd[L + "self_attn_layer_norm.weight"], sd[L + "self_attn_layer_norm.bias"], 1e-5)
        a = mha(u, sd, L + "self_attn.", cfg.heads,
                names=("linear_q", "linear_k", "linear_v", "linear_out"),
                q_in=apply_rope(u, cos, sin, cfg.heads), p_drop=cfg.attention_dropout, training=training)
        h = h + _drop(a, cfg.attention_dropout, training)   # self_attn_dropout uses attention_dropout (:583)
        h = h + conformer_conv_module(h, sd, L + "conv_module.", cfg, training, bn_stats_out)
        u = layer_norm(h, sd[L + "ffn2_layer_norm.weight"], sd[L + "ffn2_layer_norm.bias"], 1e-5)
        h = h + 0.5 * conformer_ffn(u, sd, L + "ffn2.", cfg, training)
        h = layer_norm(h, sd[L + "final_layer_norm.weight"], sd[L + "final_layer_norm.bias"], 1e-5)
        if stages is not None:
            stages[f"layer{i}"] = h
    return layer_norm(h, sd[q + "layer_norm.weight"], sd[q + "layer_norm.bias"], cfg.ln_eps)


# ---------------------------------------------------------------------------------------
# stage 4: lm_head, CTC, greedy decode
# ---------------------------------------------------------------------------------------
def lm_head(h, sd, cfg: W2VCfg, training=False):
    """w2v_custom_feat_extractor.py:149-152 / w2v_conformer_custom_feat_extractor.py:72-75."""
    h = _drop(h, cfg.final_dropout, training)
    return h @ sd["w2v_encoder.lm_head.weight"].t() + sd["w2v_encoder.lm_head.bias"]


def ctc_nll(logits_b: Tensor, labels: Tensor, T_len: int) -> Tensor:
    """Log-space CTC forward (alpha) recursion for one sample; blank=0.
    Differentiable restatement of torch.nn.CTCLoss's per-sample negative log-likelihood."""
    lp = torch.log_softmax(logits_b, dim=-1)
    S = labels.numel()
    ext = torch.zeros(2 * S + 1, dtype=torch.long)
    ext[1::2] = labels
    L = ext.numel()
    neg = torch.full((), -float("inf"), dtype=lp.dtype)
    alpha = [neg] * L
    if T_len > 0:
        alpha[0] = lp[0, ext[0]]
        if L > 1:
            alpha[1] = lp[0, ext[1]]
    for t in range(1, T_len):
        new = []
        for s in range(L):
            terms = [alpha[s]]
            if s >= 1:
                terms.append(alpha[s - 1])
            if s >= 2 and ext[s] != 0 and ext[s] != ext[s - 2]:
                terms.append(alpha[s - 2])
            st = torch.stack(terms)
            m = st.max()
            if torch.isinf(m):
                new.append(neg)
            else:
                new.append(m + torch.log(torch.exp(st - m).sum()) + lp[t, ext[s]])
        alpha = new
    if T_len == 0:
        return -neg if L > 1 else torch.zeros((), dtype=lp.dtype)
    ends = [alpha[L - 1]] + ([alpha[L - 2]] if L > 1 else [])
    st = torch.stack(ends)
    m = st.max()
    if torch.isinf(m):
        return -neg
    return -(m + torch.log(torch.exp(st - m).sum()))


def ctc_loss(logits: Tensor, targets: Tensor, logit_lens: Tensor, target_lens: Tensor, impl="aten") -> Tensor:
    """w2v_custom_feat_extractor.py:59,70,81-90: targets<1 -> -100 (ignored beyond target_lens),
    CTCLoss(blank=0, reduction='mean', zero_infinity=True) on log_softmax(logits).transpose(0,1):
    mean_b( nll_b / max(target_len_b, 1) ), infinite nll -> 0."""
    if impl == "aten":
        tg = torch.where(targets < 1, torch.tensor(-100), targets)
        return F.ctc_loss(torch.log_softmax(logits, -1).transpose(0, 1), tg, logit_lens.long(),
                          target_lens.long(), blank=0, reduction="mean", zero_infinity=True)
    losses = []
    for b in range(logits.shape[0]):
        S = int(target_lens[b])
        nll = ctc_nll(logits[b], targets[b, :S], int(logit_lens[b]))
        if torch.isinf(nll):
            nll = torch.zeros_like(nll)
        losses.append(nll / max(S, 1))
    return torch.stack(losses).mean()


def greedy_decode(logits: Tensor) -> List[List[int]]:
    """evaluator.py:108-120 + TF:wav2vec2/tokenization_wav2vec2.py:296-345: argmax over ALL
    frames (no length mask), collapse repeats, drop <pad>=0."""
    ids = logits.argmax(-1)
    out = []
    for row in ids.tolist():
        seq, prev = [], None
        for t in row:
            if t != prev and t != 0:
                seq.append(t)
            prev = t
        out.append(seq)
    return out


# ---------------------------------------------------------------------------------------
# whole model
# ---------------------------------------------------------------------------------------
@dataclass
class OracleOutput:
    logits: Tensor
    loss: Optional[Tensor]
    logit_lens: Optional[Tensor]
    hidden_states: Tensor
    stages: dict = field(default_factory=dict)


def forward(batch: dict, sd: Dict[str, Tensor], be: BrainEncoderCfg, w2v: W2VCfg, training=False,
            skip_layers=None, gru_impl="loop", ctc_impl="aten", keep_stages=False) -> OracleOutput:
    """W2VBrainEncoderModel.forward (w2v_custom_feat_extractor.py:65-123) /
    W2VConformerBrainEncoderModel.forward (w2v_conformer_custom_feat_extractor.py:36-59)."""
    stages = {} if keep_stages else None
    feats, lens = brain_encoder(batch, sd, be, training, gru_impl, stages)
    if w2v.kind == "base":
        h = w2v_base_encoder(feats, sd, w2v, training, skip_layers, stages)
    else:
        h = conformer_encoder(feats, sd, w2v, training, skip_layers, stages)
    hd = _drop(h, w2v.final_dropout, training)
    logits = hd @ sd["w2v_encoder.lm_head.weight"].t() + sd["w2v_encoder.lm_head.bias"]
    loss = None
    if batch.get("target") is not None and lens is not None and batch.get("target_lens") is not None:
        loss = ctc_loss(logits, batch["target"], lens, batch["target_lens"], ctc_impl)
    return OracleOutput(logits, loss, lens, hd, stages or {})


BUFFER_SUFFIXES = ("gaussian_smoother.weight", "inv_freq", "running_mean", "running_var", "num_batches_tracked")


def is_buffer(key: str) -> bool:
    """state_dict entries that are registered buffers, not nn.Parameters."""
    return key.endswith(BUFFER_SUFFIXES)


def train_step_grads(batch, sd, be, w2v, gru_impl="aten", dtype=torch.float32):
    """loss.backward() of train_loop.py:67-70 with every dropout at 0: returns
    (OracleOutput, {param name: grad}).  Parameters the reference never touches in forward
     get no entry."""
    p = {k: (v.detach().clone().to(dtype).requires_grad_(not is_buffer(k)) if v.is_floating_point() else v)
         for k, v in sd.items()}
    b = {k: (v.to(dtype) if (isinstance(v, Tensor) and v.is_floating_point()) else v) for k, v in batch.items()}
    out = forward(b, p, be, w2v, training=True, gru_impl=gru_impl)
    out.loss.backward()
    grads = {k: v.grad for k, v in p.items() if isinstance(v, Tensor) and v.is_floating_point() and v.grad is not None}
    return out, grads
