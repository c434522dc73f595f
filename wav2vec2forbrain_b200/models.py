"""Model classes with the reference's API, state_dict keys and call contract, running on the
B200 kernels.

Mirrors (same names, constructor arguments, forward contract and error behaviour):
  ModelOutput, B2TModel                        src/model/b2tmodel.py:9-21
  B2tSampleBatch                               src/datasets/batch_types.py:5-36
  B2P2TModel / GaussianSmoothing               src/model/b2p2t_model.py:27-183
  BrainFeatureExtractor, B2TBrainFeatureExtractor, bfe_w_preprocessing_from_config
                                               src/model/brain_feature_extractor.py:24-123
  W2VBrainEncoderModel, Wav2Vec2WithoutFeatExtrForCTC
                                               src/model/w2v_custom_feat_extractor.py:22-191
  W2VConformerBrainEncoderModel, Wav2Vec2ConformerWithoutFeatExtrForCTC
                                               src/model/w2v_conformer_custom_feat_extractor.py:16-112
Parameters are ordinary nn.Parameters under the reference's names, so
optimizers, clip_grad_norm_, state_dict()/load_state_dict(strict=True) and checkpoints keep
working; the arithmetic is three autograd nodes (brain encoder, w2v encoder + head, CTC) whose
forward/backward are explicit kernel schedules (engine.py).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, NamedTuple, Optional

import torch
from torch import nn

from . import engine as _engine
from . import ops as _ops
from .config import (N_DAYS, NEURAL_DIM, PRETRAINED_LATENT_SIZES, SMOOTH_TAPS, BrainEncoderConfig, W2VConfig, base_config,
                     conformer_large_config)

Tensor = torch.Tensor


@dataclass
class ModelOutput:
    logits: Tensor
    metrics: dict
    loss: Optional[Tensor] = None
    logit_lens: Optional[Tensor] = None
    hidden_states: Optional[Tensor] = None


class SampleBatch(NamedTuple):
    input: Tensor
    target: Optional[Tensor]

    def cuda(self):
        copy = self._replace(input=self.input.cuda(), target=self.target.cuda() if self.target is not None else None)
        if hasattr(self, "__dict__"):
            for key, value in self.__dict__.items():
                copy.__setattr__(key, value.cuda() if isinstance(value, Tensor) else value)
        return copy

    def copy_and_change(self, **diff):
        copy = self._replace(**diff)
        for key, value in self.__dict__.items():
            copy.__setattr__(key, value)
        return copy


class B2tSampleBatch(SampleBatch):
    day_idxs: Tensor
    input_lens: Tensor
    target_lens: Optional[Tensor]


class B2TModel(nn.Module):
    def forward(self, batch) -> ModelOutput:   # pragma: no cover - interface
        raise NotImplementedError


class _LazyMetrics(dict):
    """`metrics` dict of ModelOutput whose float values are materialised (device->host sync) only
    when somebody reads them.  The reference calls ctc_loss.item() inside forward
    (w2v_custom_feat_extractor.py:94), a stream sync per step that readers who never look at the
    metrics should not pay for; readers (evaluator.track_batch, history.MetricEntry) see plain floats.

    Every read path materialises: __iter__ is overridden on purpose, which also takes CPython off its
    dict-subclass fast paths (dict(m), {**m}, m.copy(), json.dump), so none of them can see a tensor."""

    def set_lazy(self, key, tensor):
        dict.__setitem__(self, key, tensor)

    def _mat_all(self):
        for k in list(dict.keys(self)):
            v = dict.__getitem__(self, k)
            if isinstance(v, Tensor):
                dict.__setitem__(self, k, v.item())

    def __getitem__(self, key):
        self._mat_all()
        return dict.__getitem__(self, key)

    def get(self, key, default=None):
        self._mat_all()
        return dict.get(self, key, default)

    def __iter__(self):
        self._mat_all()
        return dict.__iter__(self)

    def keys(self):
        self._mat_all()
        return dict.keys(self)

    def items(self):
        self._mat_all()
        return dict.items(self)

    def values(self):
        self._mat_all()
        return dict.values(self)

    def copy(self):
        self._mat_all()
        return dict(dict.items(self))

    def __reduce__(self):
        self._mat_all()
        return (dict, (dict(dict.items(self)),))

    def __repr__(self):
        self._mat_all()
        return dict.__repr__(self)


# ---------------------------------------------------------------------------------------------
# shared runtime state: one Engine per top-level model, dropout / LayerDrop seeds
# ---------------------------------------------------------------------------------------------
class _Runtime:
    def __init__(self, be: BrainEncoderConfig, w2v: Optional[W2VConfig], precision: str, seed: int = 42):
        self.engine = _engine.Engine(be, w2v if w2v is not None else base_config(), precision)
        self.step = 0
        self.seed = seed
        self.rank = 0
        # LayerDrop decisions come from a host generator seeded identically on every rank
        self.layerdrop_gen = torch.Generator().manual_seed(seed)
        self.augment = (0.0, 0.0)   # (whiteNoiseSD, constantOffsetSD); (0, 0) = the reference's observable behaviour
        # frozen-w2v training (unfreeze_strategy == "brain_encoder"): skip the w2v weight gradients nobody reads
        # (set by compat's _create_model swap and by bench.py --workload frozen)
        self.skip_w2v_wgrad = False
        # flat gradient arena (flat.GradArena): built lazily from the owning top-level model's parameters
        self.use_arena = True
        self._owner = None          # weakref to the W2V*BrainEncoderModel
        self._arena = None

    def grad_arena(self):
        """The model's flat gradient buffer (None for a stand-alone brain encoder or when switched off)."""
        if not self.use_arena or self._owner is None:
            return None
        if self._arena is not None:       # invalidated by _TopModel._apply (.cuda() / .to() / dtype changes)
            return self._arena
        owner = self._owner()
        if owner is None:
            return None
        named = [(n, p) for n, p in owner.named_parameters() if p.requires_grad]
        if not named or (not named[0][1].is_cuda and not _EMULATED):
            return None
        from .flat import GradArena
        self._arena = GradArena(named, _engine.Engine.grad_groups(self.engine.be, self.engine.w2v), named[0][1].device)
        self.engine.grad_arena = self._arena
        return self._arena

    def begin_node(self, prefix: str) -> bool:
        """Start of an autograd node's backward: switch the engine to the arena slots of `prefix` (zeroed with one
        memset) when none of those parameters holds a .grad yet."""
        eng = self.engine
        arena = self.grad_arena()
        on = arena is not None and arena.free_for(prefix)
        eng.grad_arena, eng._arena_on = arena, on
        if on:
            arena.zero_prefix(prefix)
        return on

    def end_node(self, G, prefix: str) -> None:
        eng = self.engine
        eng.finalize_grads(G)
        eng.sync_grads(G, prefix)
        eng._arena_on = False

    def next_seed(self) -> int:
        self.step += 1
        return _engine._mix(self.seed * 1000003 + self.rank, self.step)

    def layerdrop_skips(self, n_layers: int, p: float, training: bool):
        if not training or p <= 0.0:
            return ()
        # transformers modeling_wav2vec2.py:702-704: one torch.rand([]) per layer per forward
        r = torch.rand(n_layers, generator=self.layerdrop_gen)
        return tuple(i for i in range(n_layers) if r[i].item() < p)


def _named_tensors(module: nn.Module, prefix: str) -> Dict[str, Tensor]:
    """name -> parameter / buffer of `module`, in named_parameters() + named_buffers() order.  Walking the 700-tensor
    encoder cost ~2.5 ms of Python per forward, so the walk is cached and only re-validated: every cached object must
    still be the one registered under its name (optimizer steps and load_state_dict write in place; .to() / .cuda() /
    re-assignment replace the objects and trigger a new walk)."""
    cache = module.__dict__.get("_b2t_named")
    if cache is not None:
        names, objs, where = cache
        for (reg, attr), o in zip(where, objs):
            if reg.get(attr) is not o:
                break
        else:
            return dict(zip(names, objs))
    names, objs, where, memo = [], [], [], set()
    for members in ("_parameters", "_buffers"):
        for mod_prefix, mod in module.named_modules():
            reg = getattr(mod, members)
            for attr, v in reg.items():
                if v is None or id(v) in memo:      # a shared tensor is listed once, like named_parameters()
                    continue
                memo.add(id(v))
                names.append(prefix + (mod_prefix + "." if mod_prefix else "") + attr)
                objs.append(v)
                where.append((reg, attr))
    module.__dict__["_b2t_named"] = (names, objs, where)
    return dict(zip(names, objs))


class _BrainFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rt: _Runtime, names, training, seed, x, day, *params):
        P = dict(zip(names, params))
        feats, saved = rt.engine.brain_forward(P, x, day, training, seed, rt.augment if training else (0.0, 0.0))
        ctx.rt, ctx.names, ctx.saved, ctx.P = rt, names, saved, P
        return feats.view(saved["B"], saved["Tp"], -1)

    @staticmethod
    def backward(ctx, dfeats):
        G: Dict[str, Tensor] = {}
        d = dfeats.reshape(-1, dfeats.shape[-1]).contiguous()
        with _ops.zero_arena(("brain", id(ctx.rt)), d.device):
            ctx.rt.begin_node("brain_encoder.")
            try:
                ctx.rt.engine.brain_backward(ctx.P, ctx.saved, d, G)
                ctx.rt.end_node(G, "brain_encoder.")
            finally:
                ctx.rt.engine._arena_on = False
        return (None, None, None, None, None, None) + tuple(G.get(n) for n in ctx.names)


class _EncoderFn(torch.autograd.Function):
    """w2v encoder + final dropout + lm_head."""

    @staticmethod
    def forward(ctx, rt: _Runtime, names, training, seed, skip, feats, *params):
        P = dict(zip(names, params))
        eng = rt.engine
        B, T, D = feats.shape
        f2 = feats.reshape(B * T, D).contiguous()
        if eng.w2v.kind == "base":
            h, Se = eng.w2v_base_forward(P, f2, B, T, training, seed, skip)
        else:
            from . import conformer
            h, Se = conformer.forward(eng, P, f2, B, T, training, seed, skip)
        logits, hd, Sh = eng.head_forward(P, h, B, T, training, seed)
        ctx.rt, ctx.names, ctx.P, ctx.Se, ctx.Sh = rt, names, P, Se, Sh
        return logits, hd.view(B, T, D)

    @staticmethod
    def backward(ctx, dlogits, dhidden):
        eng = ctx.rt.engine
        G: Dict[str, Tensor] = {}
        skip = ctx.rt.skip_w2v_wgrad
        with _ops.zero_arena(("w2v", id(ctx.rt)), dlogits.device):
            eng._skip_wgrad = skip
            ctx.rt.begin_node("w2v_encoder.")
            try:
                dh = eng.head_backward(ctx.P, ctx.Sh, dlogits.contiguous(), G)
                if dhidden is not None:
                    dh = dh + eng_dropout_bwd(eng, ctx.Sh, dhidden)
                if eng.w2v.kind == "base":
                    dfeats = eng.w2v_base_backward(ctx.P, ctx.Se, dh, G)
                else:
                    from . import conformer
                    dfeats = conformer.backward(eng, ctx.P, ctx.Se, dh, G)
                if skip:
                    G = {}      # data gradient only: no w2v parameter receives a .grad
                ctx.rt.end_node(G, "w2v_encoder.")
            finally:
                eng._skip_wgrad = False
                eng._arena_on = False
        B, T = ctx.Sh["B"], ctx.Sh["T"]
        return (None, None, None, None, None, dfeats.view(B, T, -1)) + tuple(G.get(n) for n in ctx.names)


def eng_dropout_bwd(eng, Sh, dhidden):
    d = dhidden.reshape(-1, dhidden.shape[-1]).contiguous()
    if Sh["p_f"] > 0:
        d = _engine.ops.dropout(d, Sh["p_f"], Sh["seed"], three=eng.three)[0]
    return d


class _CTCFn(torch.autograd.Function):
    """log_softmax + CTCLoss(blank=0, reduction='mean', zero_infinity=True) in one kernel that also emits
    d loss / d logits (w2v_custom_feat_extractor.py:59,81-90)."""

    @staticmethod
    def forward(ctx, logits, targets, logit_lens, target_lens):
        loss, nll, grad = _engine.ops.ctc_fwd_bwd(logits.contiguous(), targets, logit_lens, target_lens,
                                                  want_grad=ctx.needs_input_grad[0])
        ctx.grad = grad
        return loss

    @staticmethod
    def backward(ctx, dloss):
        return ctx.grad * dloss, None, None, None


# ---------------------------------------------------------------------------------------------
# brain encoder modules
# ---------------------------------------------------------------------------------------------
def gaussian_taps(sigma: float, size: int = SMOOTH_TAPS) -> Tensor:
    """GaussianSmoothing.__init__ (b2p2t_model.py:41-68) for dim=1."""
    grid = torch.arange(size, dtype=torch.float32)
    mean = (size - 1) / 2
    k = 1 / (sigma * math.sqrt(2 * math.pi)) * torch.exp(-(((grid - mean) / sigma) ** 2) / 2)
    return k / torch.sum(k)


class GaussianSmoothing(nn.Module):
    def __init__(self, channels: int, kernel_size: int, sigma: float, dim: int = 1):
        super().__init__()
        if dim != 1:
            raise RuntimeError("Only 1 dimension is supported on this path.")
        self.register_buffer("weight", gaussian_taps(sigma, kernel_size).view(1, 1, -1).repeat(channels, 1, 1))
        self.groups = channels


class _GRUParams(nn.Module):
    """Parameter container with nn.GRU's parameter names and initialisation (no cuDNN behind it)."""

    def __init__(self, in_size: int, hidden: int, layers: int, bias: bool, bidirectional: bool):
        super().__init__()
        ndir = 2 if bidirectional else 1
        bound = 1.0 / math.sqrt(hidden)
        for l in range(layers):
            k = in_size if l == 0 else hidden * ndir
            for sfx in [""] + (["_reverse"] if bidirectional else []):
                self.register_parameter(f"weight_ih_l{l}{sfx}", nn.Parameter(torch.empty(3 * hidden, k).uniform_(-bound, bound)))
                self.register_parameter(f"weight_hh_l{l}{sfx}", nn.Parameter(torch.empty(3 * hidden, hidden).uniform_(-bound, bound)))
                if bias:
                    self.register_parameter(f"bias_ih_l{l}{sfx}", nn.Parameter(torch.empty(3 * hidden).uniform_(-bound, bound)))
                    self.register_parameter(f"bias_hh_l{l}{sfx}", nn.Parameter(torch.empty(3 * hidden).uniform_(-bound, bound)))


class _Act(nn.Module):
    """parameter-free placeholder that keeps nn.Sequential indices 1, 3, ... occupied (nn_helper.py:39-49)."""


def create_fully_connected(input_size: int, output_size: int, hidden_sizes=(), activation: str = "gelu") -> nn.Sequential:
    layers = []
    sizes = [input_size] + list(hidden_sizes) + [output_size]
    for i in range(len(sizes) - 1):
        layers.append(nn.Linear(sizes[i], sizes[i + 1]))
        if i < len(sizes) - 2:
            layers.append(_Act())
    return nn.Sequential(*layers)


class BrainFeatureExtractor(nn.Module):
    def __init__(self, config, in_size: int, wav2vec_checkpoint: str):
        super().__init__()
        self.config = config
        self.num_directions = 2 if config.encoder_bidirectional else 1
        self.hidden_start = nn.Parameter(torch.randn(self.num_directions * config.encoder_num_gru_layers,
                                                     config.encoder_gru_hidden_size))
        self.gru = _GRUParams(in_size, config.encoder_gru_hidden_size, config.encoder_num_gru_layers, config.encoder_bias,
                              config.encoder_bidirectional)
        self.fc = create_fully_connected(config.encoder_gru_hidden_size * self.num_directions,
                                         PRETRAINED_LATENT_SIZES[wav2vec_checkpoint], config.encoder_fc_hidden_sizes,
                                         config.encoder_fc_activation_function)


class B2TBrainFeatureExtractor(B2TModel):
    def __init__(self, config, wav2vec_checkpoint: str, in_size: int):
        super().__init__()
        self.encoder = BrainFeatureExtractor(config, in_size, wav2vec_checkpoint)


def _be_config_from_args(config, wav2vec_checkpoint: str) -> BrainEncoderConfig:
    return BrainEncoderConfig(
        gru_hidden=config.encoder_gru_hidden_size, gru_layers=config.encoder_num_gru_layers,
        bidirectional=config.encoder_bidirectional, bias=getattr(config, "encoder_bias", True),
        fc_hidden=tuple(config.encoder_fc_hidden_sizes), fc_act=config.encoder_fc_activation_function,
        out_size=PRETRAINED_LATENT_SIZES[wav2vec_checkpoint], unfold_len=config.unfolder_kernel_len,
        unfold_stride=config.unfolder_stride_len, smooth_sigma=config.gaussian_smooth_width,
        dropout=config.encoder_dropout, learnable_h0=config.encoder_learnable_inital_state,
        nonlinearity=config.input_layer_nonlinearity)


class B2P2TModel(B2TModel):
    """Day-dependent preprocessing + GRU brain encoder (b2p2t_model.py:93-183) on the B200 kernels."""

    def __init__(self, config, neural_decoder: B2TBrainFeatureExtractor, pad_token_id=0, wav2vec_checkpoint: str = None,
                 precision: str = "bf16x3"):
        super().__init__()
        self.config = config
        self.pad_token_id = pad_token_id
        if config.input_layer_nonlinearity != "softsign":
            raise NotImplementedError("Only softsign is currently supported as input layer nonlinearity")
        self.gaussian_smoother = GaussianSmoothing(NEURAL_DIM, SMOOTH_TAPS, config.gaussian_smooth_width, dim=1)
        self.day_weights = nn.Parameter(torch.eye(NEURAL_DIM).repeat(N_DAYS, 1, 1))
        self.day_bias = nn.Parameter(torch.zeros(N_DAYS, 1, NEURAL_DIM))
        self.neural_decoder = neural_decoder
        for d in range(N_DAYS):   # created by the reference, never used in forward
            lin = nn.Linear(NEURAL_DIM, NEURAL_DIM)
            lin.weight = nn.Parameter(lin.weight.detach() + torch.eye(NEURAL_DIM))
            setattr(self, "inpLayer" + str(d), lin)
        self._prefix = "brain_encoder."
        self._rt: Optional[_Runtime] = None
        self._ckpt = wav2vec_checkpoint
        self._precision = precision

    def _runtime(self) -> _Runtime:
        if self._rt is None:
            self._rt = _Runtime(_be_config_from_args(self.config, self._ckpt), None, self._precision)
        return self._rt

    @classmethod
    def get_in_size_after_preprocessing(cls, unfolder_kernel_len: int):
        return NEURAL_DIM * unfolder_kernel_len

    def _features(self, batch, seed: Optional[int] = None) -> Tensor:
        rt = self._runtime()
        x = batch.input
        if not x.is_cuda and not _EMULATED:
            raise RuntimeError("wav2vec2forbrain_b200 runs on CUDA (sm_100a) only: call batch.cuda() first")
        day = batch.day_idxs.to(device=x.device, dtype=torch.int32)
        tensors = _named_tensors(self, self._prefix)
        names = tuple(tensors)
        training = self.training
        seed = rt.next_seed() if seed is None else seed
        if torch.is_grad_enabled():
            return _BrainFn.apply(rt, names, training, seed, x.float().contiguous(), day, *tensors.values())
        feats, saved = rt.engine.brain_forward(tensors, x.float().contiguous(), day, training, seed,
                                               rt.augment if training else (0.0, 0.0))
        return feats.view(saved["B"], saved["Tp"], -1)

    def forward(self, batch) -> ModelOutput:
        feats = self._features(batch)
        out = ModelOutput(logits=feats, metrics={})
        if hasattr(batch, "input_lens"):
            out.logit_lens = self._runtime().engine.processed_lens(batch.input_lens)
        return out


def bfe_w_preprocessing_from_config(config, brain_encoder_path: Optional[str], wav2vec_checkpoint: str,
                                    precision: str = "bf16x3") -> B2P2TModel:
    model = B2P2TModel(config,
                       B2TBrainFeatureExtractor(config, wav2vec_checkpoint,
                                                B2P2TModel.get_in_size_after_preprocessing(config.unfolder_kernel_len)),
                       wav2vec_checkpoint=wav2vec_checkpoint, precision=precision).cuda()
    if brain_encoder_path is not None:
        state = torch.load(brain_encoder_path, map_location="cuda")
        for key in [k for k in state if k.startswith("neural_decoder.discriminator") or k.startswith("neural_decoder.suc_for_ctc")]:
            del state[key]
        model.load_state_dict(state, strict=True)
    return model


# ---------------------------------------------------------------------------------------------
# wav2vec2 / conformer encoder parameter trees (names of transformers' modules)
# ---------------------------------------------------------------------------------------------
class _Tree(nn.Module):
    """Generic nested container: registers parameters / buffers under dotted names."""

    def add(self, dotted: str, tensor: Tensor, buffer: bool = False):
        head, _, rest = dotted.partition(".")
        if not rest:
            if buffer:
                self.register_buffer(head, tensor)
            else:
                self.register_parameter(head, nn.Parameter(tensor))
            return
        if head not in self._modules:
            self.add_module(head, _Tree())
        self._modules[head].add(rest, tensor, buffer)


def _normal(*shape, std=0.02):
    return torch.randn(*shape) * std


def _build_w2v_tree(cfg: W2VConfig) -> _Tree:
    """Random init following transformers' _init_weights (Linear N(0, 0.02), LayerNorm 1/0, pos-conv
    N(0, 2*sqrt(1/(k*in))) with weight_norm g = ||v||)."""
    t = _Tree()
    D = cfg.hidden

    def lin(p, out_f, in_f, bias=True):
        t.add(p + ".weight", _normal(out_f, in_f))
        if bias:
            t.add(p + ".bias", torch.zeros(out_f))

    def ln(p, n):
        t.add(p + ".weight", torch.ones(n))
        t.add(p + ".bias", torch.zeros(n))

    if cfg.kind == "base":
        q = "wav2vec2.encoder."
    else:
        q = "wav2vec2_conformer.encoder."
        hd = D // cfg.heads
        t.add(q + "embed_positions.inv_freq",
              1.0 / (cfg.rope_base ** (torch.arange(0, hd, 2, dtype=torch.int64).float() / hd)), buffer=True)
    cin = D // cfg.pos_conv_groups
    t.add(q + "pos_conv_embed.conv.bias", torch.zeros(D))
    v = _normal(D, cin, cfg.pos_conv_k, std=2 * math.sqrt(1 / (cfg.pos_conv_k * cin)))
    t.add(q + "pos_conv_embed.conv.parametrizations.weight.original0", v.norm(dim=(0, 1), keepdim=True))
    t.add(q + "pos_conv_embed.conv.parametrizations.weight.original1", v)
    ln(q + "layer_norm", D)
    for i in range(cfg.layers):
        L = q + f"layers.{i}."
        if cfg.kind == "base":
            for nm in ("k_proj", "v_proj", "q_proj", "out_proj"):
                lin(L + "attention." + nm, D, D)
            ln(L + "layer_norm", D)
            lin(L + "feed_forward.intermediate_dense", cfg.ffn, D)
            lin(L + "feed_forward.output_dense", D, cfg.ffn)
            ln(L + "final_layer_norm", D)
        else:
            ln(L + "ffn1_layer_norm", D)
            lin(L + "ffn1.intermediate_dense", cfg.ffn, D)
            lin(L + "ffn1.output_dense", D, cfg.ffn)
            ln(L + "self_attn_layer_norm", D)
            for nm in ("linear_q", "linear_k", "linear_v", "linear_out"):
                lin(L + "self_attn." + nm, D, D)
            ln(L + "conv_module.layer_norm", D)
            t.add(L + "conv_module.pointwise_conv1.weight", _normal(2 * D, D, 1, std=math.sqrt(1 / D)))
            t.add(L + "conv_module.depthwise_conv.weight", _normal(D, 1, cfg.dw_kernel, std=math.sqrt(1 / cfg.dw_kernel)))
            t.add(L + "conv_module.batch_norm.weight", torch.ones(D))
            t.add(L + "conv_module.batch_norm.bias", torch.zeros(D))
            t.add(L + "conv_module.batch_norm.running_mean", torch.zeros(D), buffer=True)
            t.add(L + "conv_module.batch_norm.running_var", torch.ones(D), buffer=True)
            t.add(L + "conv_module.batch_norm.num_batches_tracked", torch.tensor(0, dtype=torch.int64), buffer=True)
            t.add(L + "conv_module.pointwise_conv2.weight", _normal(D, D, 1, std=math.sqrt(1 / D)))
            ln(L + "ffn2_layer_norm", D)
            lin(L + "ffn2.intermediate_dense", cfg.ffn, D)
            lin(L + "ffn2.output_dense", D, cfg.ffn)
            ln(L + "final_layer_norm", D)
    lin("lm_head", cfg.vocab, D)
    return t


_ARCH = {
    "facebook/wav2vec2-base-960h": base_config,
    "facebook/wav2vec2-base-100h": base_config,
    "facebook/wav2vec2-conformer-rope-large-960h-ft": conformer_large_config,
}


def register_architecture(name: str, factory, latent_size: int) -> None:
    """Tests and offline runs: make `wav2vec_checkpoint=name` resolve to a local architecture."""
    _ARCH[name] = factory
    PRETRAINED_LATENT_SIZES[name] = latent_size


class _W2VEncoderBase(nn.Module):
    """Common part of Wav2Vec2WithoutFeatExtrForCTC / Wav2Vec2ConformerWithoutFeatExtrForCTC."""

    def __init__(self, cfg: W2VConfig):
        super().__init__()
        self.cfg = cfg
        tree = _build_w2v_tree(cfg)
        for name, mod in tree._modules.items():
            self.add_module(name, mod)
        self._prefix = "w2v_encoder."
        self._rt: Optional[_Runtime] = None

    def _run(self, feats: Tensor):
        rt = self._rt
        if rt is None:
            raise RuntimeError("w2v_encoder must be owned by a W2V*BrainEncoderModel before it is called")
        tensors = _named_tensors(self, self._prefix)
        names = tuple(tensors)
        training = self.training
        seed = rt.next_seed()
        skip = rt.layerdrop_skips(self.cfg.layers, self.cfg.layerdrop, training)
        if torch.is_grad_enabled():
            return _EncoderFn.apply(rt, names, training, seed, skip, feats, *tensors.values())
        eng = rt.engine
        B, T, D = feats.shape
        f2 = feats.reshape(B * T, D).contiguous()
        if self.cfg.kind == "base":
            h, _ = eng.w2v_base_forward(tensors, f2, B, T, training, seed, skip)
        else:
            from . import conformer
            h, _ = conformer.forward(eng, tensors, f2, B, T, training, seed, skip)
        logits, hd, _ = eng.head_forward(tensors, h, B, T, training, seed)
        return logits, hd.view(B, T, D)


class Wav2Vec2WithoutFeatExtrForCTC(_W2VEncoderBase):
    def forward(self, x: Tensor):
        return self._run(x)          # (logits, hidden_states)  w2v_custom_feat_extractor.py:144-153


class Wav2Vec2ConformerWithoutFeatExtrForCTC(_W2VEncoderBase):
    def forward(self, x: Tensor):
        return self._run(x)[0]       # logits only  w2v_conformer_custom_feat_extractor.py:67-76


def _resolve_arch(wav2vec_checkpoint: str, **overrides) -> W2VConfig:
    if wav2vec_checkpoint not in _ARCH:
        raise KeyError(f"unknown wav2vec_checkpoint {wav2vec_checkpoint!r}: offline box, only architectures in "
                       f"{sorted(_ARCH)} (or register_architecture) can be constructed")
    return _ARCH[wav2vec_checkpoint](**overrides)


def _load_pretrained_if_present(module: nn.Module, wav2vec_checkpoint: str) -> bool:
    """The reference calls from_pretrained(checkpoint) (network / HF cache).  Offline: load a local
    state_dict if `wav2vec_checkpoint` is a directory holding pytorch_model.bin, else keep random init."""
    import os
    path = os.path.join(wav2vec_checkpoint, "pytorch_model.bin")
    if os.path.isdir(wav2vec_checkpoint) and os.path.exists(path):
        sd = torch.load(path, map_location="cpu")
        own = module.state_dict()
        module.load_state_dict({k: v for k, v in sd.items() if k in own and own[k].shape == v.shape}, strict=False)
        return True
    return False


# True only inside the CPU test-suite, which substitutes tests/abi_emulator.py for `engine.ops` to check the host-side
# schedule, the arena and the data-parallel logic without a GPU; the product never sets it (and without the emulator a
# CPU tensor fails in the first kernel wrapper anyway)
_EMULATED = False


class _TopModel(B2TModel):
    def _apply(self, fn, *a, **kw):
        out = super()._apply(fn, *a, **kw)
        rt = getattr(self, "_rt", None)
        if rt is not None:
            rt._arena = None            # parameters may have moved: the arena is rebuilt on the next backward
            rt.engine.grad_arena = None
        return out

    def _attach(self, brain_encoder: B2P2TModel, w2v_cfg: W2VConfig, precision: str):
        be_cfg = _be_config_from_args(brain_encoder.config, brain_encoder._ckpt)
        rt = _Runtime(be_cfg, w2v_cfg, precision)
        import weakref
        rt._owner = weakref.ref(self)
        brain_encoder._rt = rt
        self.w2v_encoder._rt = rt
        self._rt = rt

    @property
    def runtime(self) -> _Runtime:
        return self._rt

    def _forward_impl(self, batch, with_hidden: bool) -> ModelOutput:
        targets = batch.target
        assert targets is not None
        feats = self.brain_encoder._features(batch)
        logit_lens = self._rt.engine.processed_lens(batch.input_lens) if hasattr(batch, "input_lens") else None
        res = self.w2v_encoder._run(feats)
        logits, hidden = res
        ctc_loss = None
        target_lens = getattr(batch, "target_lens", None)
        if target_lens is not None and logit_lens is not None:
            ctc_loss = _CTCFn.apply(logits, targets.to(logits.device), logit_lens.to(logits.device),
                                    target_lens.to(logits.device))
        metrics = _LazyMetrics()
        if ctc_loss is not None:
            metrics.set_lazy("ctc_loss", ctc_loss.detach())
        return ModelOutput(logits, metrics, loss=ctc_loss, logit_lens=logit_lens if with_hidden else None,
                           hidden_states=hidden if with_hidden else None)


class W2VBrainEncoderModel(_TopModel):
    def __init__(self, config, brain_encoder: B2P2TModel, wav2vec_checkpoint: str, head=None, skip_loading_weights=False,
                 pre_w2v_head_for_additional_loss=None, additonal_loss_weight=None, additional_loss_squared=False,
                 precision: str = "bf16x3", **arch_overrides):
        super().__init__()
        if head is not None or pre_w2v_head_for_additional_loss is not None:
            # dead on the kept experiments (b2t_gru_w2v_experiment.py:99-105); SURVEY.md §8b
            raise NotImplementedError("head / pre_w2v_head_for_additional_loss are not part of the accelerated path")
        self.brain_encoder = brain_encoder
        cfg = _resolve_arch(wav2vec_checkpoint, stable_ln=getattr(config, "w2v_do_stable_layer_norm", False), **arch_overrides)
        if cfg.kind != "base":
            raise ValueError("W2VBrainEncoderModel expects a wav2vec2 (non-conformer) checkpoint")
        self.w2v_encoder = Wav2Vec2WithoutFeatExtrForCTC(cfg)
        if skip_loading_weights:
            print("Skipping loading weights for Wav2Vec2WithoutFeatExtrForCTC\n")
        else:
            _load_pretrained_if_present(self.w2v_encoder, wav2vec_checkpoint)
        self.head = head
        self.pre_w2v_head_for_additional_loss = pre_w2v_head_for_additional_loss
        self._attach(brain_encoder, cfg, precision)

    def forward(self, batch) -> ModelOutput:
        return self._forward_impl(batch, with_hidden=True)


class W2VConformerBrainEncoderModel(_TopModel):
    def __init__(self, brain_encoder: B2P2TModel, wav2vec_checkpoint: str, precision: str = "bf16x3", **arch_overrides):
        super().__init__()
        self.brain_encoder = brain_encoder
        cfg = _resolve_arch(wav2vec_checkpoint, **arch_overrides)
        if cfg.kind != "conformer":
            raise ValueError("W2VConformerBrainEncoderModel expects a wav2vec2-conformer checkpoint")
        self.w2v_encoder = Wav2Vec2ConformerWithoutFeatExtrForCTC(cfg)
        _load_pretrained_if_present(self.w2v_encoder, wav2vec_checkpoint)
        self._attach(brain_encoder, cfg, precision)

    def forward(self, batch) -> ModelOutput:
        return self._forward_impl(batch, with_hidden=False)


@dataclass
class BrainEncoderArgs:
    """Plain-attribute stand-in for B2P2TBrainFeatureExtractorArgsModel + W2VBrainEncoderModelArgs
    (brain_feature_extractor.py:13-21,90-93; b2p2t_model.py:20-24; w2v_custom_feat_extractor.py:18-19)."""
    encoder_gru_hidden_size: int = 256
    encoder_bidirectional: bool = True
    encoder_num_gru_layers: int = 2
    encoder_bias: bool = True
    encoder_dropout: float = 0.0
    encoder_learnable_inital_state: bool = False
    encoder_fc_hidden_sizes: tuple = ()
    encoder_fc_activation_function: str = "gelu"
    input_layer_nonlinearity: str = "softsign"
    unfolder_kernel_len: int = 32
    unfolder_stride_len: int = 4
    gaussian_smooth_width: float = 0.3
    w2v_do_stable_layer_norm: bool = False
