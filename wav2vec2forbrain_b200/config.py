"""Architecture descriptions consumed by the engine (host-side mirror of the reference's flag models:
src/model/brain_feature_extractor.py:13-21, src/model/b2p2t_model.py:20-24 and the subset of
transformers' Wav2Vec2Config / Wav2Vec2ConformerConfig that the path reads)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Sequence

N_DAYS = 24          # b2p2t_model.py:115
NEURAL_DIM = 256     # b2p2t_model.py:116
SMOOTH_TAPS = 20     # b2p2t_model.py:118

# src/args/base_args.py:5-11
PRETRAINED_LATENT_SIZES = {
    "facebook/wav2vec2-base-100h": 768,
    "facebook/wav2vec2-base-960h": 768,
    "facebook/wav2vec2-large-960h": 1024,
    "facebook/wav2vec2-large-960h-lv60-self": 1024,
    "facebook/wav2vec2-conformer-rope-large-960h-ft": 1024,
}


@dataclass
class BrainEncoderConfig:
    gru_hidden: int = 512
    gru_layers: int = 3
    bidirectional: bool = True
    bias: bool = True
    fc_hidden: Sequence[int] = (256,)
    fc_act: str = "gelu"
    out_size: int = 768
    unfold_len: int = 32
    unfold_stride: int = 4
    smooth_sigma: float = 1.5
    dropout: float = 0.0
    learnable_h0: bool = False
    nonlinearity: str = "softsign"

    @property
    def ndir(self) -> int:
        return 2 if self.bidirectional else 1


@dataclass
class W2VConfig:
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
    hidden_act: str = "gelu"
    # conformer only
    dw_kernel: int = 31
    rope_base: int = 10000
    # dropouts
    hidden_dropout: float = 0.1
    attention_dropout: float = 0.1
    activation_dropout: float = 0.1
    final_dropout: float = 0.1
    conv_dropout: float = 0.1
    layerdrop: float = 0.1

    def no_dropout(self) -> "W2VConfig":
        from dataclasses import replace
        return replace(self, hidden_dropout=0.0, attention_dropout=0.0, activation_dropout=0.0, final_dropout=0.0,
                       conv_dropout=0.0, layerdrop=0.0)


def base_config(**kw) -> W2VConfig:
    """facebook/wav2vec2-base-960h architecture."""
    return W2VConfig(kind="base", **kw)


def conformer_large_config(**kw) -> W2VConfig:
    """facebook/wav2vec2-conformer-rope-large-960h-ft architecture."""
    d = dict(kind="conformer", hidden=1024, layers=24, heads=16, ffn=4096, hidden_act="swish", layerdrop=0.1)
    d.update(kw)
    return W2VConfig(**d)
