"""Generate tests/golden/*.pt by running the LIVE reference (imported from /root/reference).

Run in the build container only (the GPU box has no /root/reference):

    python oracle/gen_golden.py            # writes tests/golden/*.pt

The reference's own classes are used unmodified; the harness-side patches are the ones listed
in SURVEY.md §8(c): local HF config construction (no network), `.cuda()` -> identity on this
CPU-only box, eager attention, all dropout probabilities 0 so that gradients are deterministic.
Weights are injected with `load_state_dict(strict=True)` from `oracle.b2t_oracle.make_state_dict`
so that every consumer (oracle, CUDA path) can rebuild the identical model from a seed.
"""
from __future__ import annotations

import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
REF = os.environ.get("B2T_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

from oracle import b2t_oracle as O  # noqa: E402

# ---- harness-side patches (reference files untouched) --------------------------------------
torch.Tensor.cuda = lambda self, *a, **k: self
torch.nn.Module.cuda = lambda self, *a, **k: self

import transformers  # noqa: E402
from transformers import Wav2Vec2Config, Wav2Vec2ConformerConfig  # noqa: E402

_CFG_REGISTRY: dict = {}


def _hf_cfg(w: O.W2VCfg, **kw):
    common = dict(hidden_size=w.hidden, num_hidden_layers=w.layers, num_attention_heads=w.heads,
                  intermediate_size=w.ffn, vocab_size=w.vocab, layer_norm_eps=w.ln_eps,
                  num_conv_pos_embeddings=w.pos_conv_k, num_conv_pos_embedding_groups=w.pos_conv_groups,
                  hidden_dropout=0.0, attention_dropout=0.0, activation_dropout=0.0, final_dropout=0.0,
                  feat_proj_dropout=0.0, layerdrop=0.0, pad_token_id=0)
    if w.kind == "base":
        return Wav2Vec2Config(hidden_act="gelu", **common, **kw)
    return Wav2Vec2ConformerConfig(hidden_act="swish", position_embeddings_type="rotary",
                                   rotary_embedding_base=w.rope_base, conv_depthwise_kernel_size=w.dw_kernel,
                                   conformer_conv_dropout=0.0, **common, **kw)


Wav2Vec2Config.from_pretrained = classmethod(lambda cls, name, **kw: _hf_cfg(_CFG_REGISTRY[name], **kw))
Wav2Vec2ConformerConfig.from_pretrained = classmethod(lambda cls, name, **kw: _hf_cfg(_CFG_REGISTRY[name], **kw))

from src.args import base_args  # noqa: E402
from src.datasets.batch_types import B2tSampleBatch  # noqa: E402
from src.model.brain_feature_extractor import (  # noqa: E402
    B2P2TBrainFeatureExtractorArgsModel, bfe_w_preprocessing_from_config)
from src.model.w2v_custom_feat_extractor import W2VBrainEncoderModel, W2VBrainEncoderModelArgs  # noqa: E402
from src.model import w2v_conformer_custom_feat_extractor as conf_mod  # noqa: E402

conf_mod.Wav2Vec2ConformerWithoutFeatExtrForCTC.from_pretrained = classmethod(
    lambda cls, name, config=None, **kw: cls(config))


class _Cfg(B2P2TBrainFeatureExtractorArgsModel, W2VBrainEncoderModelArgs):
    pass


def build_reference(name: str, be: O.BrainEncoderCfg, w2v: O.W2VCfg, sd):
    _CFG_REGISTRY[name] = w2v
    base_args.PRETRAINED_LATENT_SIZES[name] = be.out_size
    import src.model.brain_feature_extractor as bfe
    bfe.PRETRAINED_LATENT_SIZES[name] = be.out_size
    cfg = _Cfg(encoder_gru_hidden_size=be.gru_hidden, encoder_num_gru_layers=be.gru_layers,
               encoder_bidirectional=be.bidirectional, encoder_fc_hidden_sizes=list(be.fc_hidden),
               encoder_dropout=0.0, encoder_learnable_inital_state=be.learnable_h0,
               gaussian_smooth_width=be.smooth_sigma, unfolder_kernel_len=be.unfold_len,
               unfolder_stride_len=be.unfold_stride, w2v_do_stable_layer_norm=w2v.stable_ln)
    brain = bfe_w_preprocessing_from_config(cfg, None, name)
    if w2v.kind == "base":
        m = W2VBrainEncoderModel(cfg, brain, name, None, True)
    else:
        m = conf_mod.W2VConformerBrainEncoderModel(brain, name)
    m.w2v_encoder.config._attn_implementation = "eager"
    missing = m.load_state_dict(sd, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return m


def ref_batch(b: dict):
    rb = B2tSampleBatch(b["input"].clone(), b["target"].clone())
    rb.day_idxs = b["day_idxs"].clone()
    rb.input_lens = b["input_lens"].clone()
    rb.target_lens = b["target_lens"].clone()
    return rb


def run_case(name, be, w2v, B, T, S, ragged, full_tensors, seed_w=42, seed_x=1234):
    sd = O.make_state_dict(be, w2v, seed=seed_w)
    batch = O.make_batch(B, T, S, seed=seed_x, ragged=ragged)
    m = build_reference(name, be, w2v, sd)
    # eval first: the train pass below updates the conformer's BatchNorm running statistics
    m.eval()
    with torch.no_grad():
        out_eval = m.forward(ref_batch(batch))
        feats = m.brain_encoder.forward(ref_batch(batch)).logits
    m.train()
    out = m.forward(ref_batch(batch))
    out.loss.backward()
    grads = {k: p.grad.detach().clone() for k, p in m.named_parameters() if p.grad is not None}
    no_grad = sorted(k for k, p in m.named_parameters() if p.grad is None)
    ids = out_eval.logits.argmax(-1)
    rec = {
        "case": name, "be": be.__dict__, "w2v": w2v.__dict__,
        "B": B, "T": T, "S": S, "ragged": ragged, "seed_w": seed_w, "seed_x": seed_x,
        "logits_train": out.logits.detach().clone(), "loss_train": out.loss.detach().clone(),
        "logits_eval": out_eval.logits.clone(), "loss_eval": out_eval.loss.clone(),
        "argmax_eval": ids.to(torch.int16),
        "logit_lens": None if out.logit_lens is None else out.logit_lens.clone(),
        "params_without_grad": no_grad,
        "grad_norm": {k: v.double().norm().item() for k, v in grads.items()},
        "torch": torch.__version__, "transformers": transformers.__version__,
    }
    if full_tensors:
        rec["features"] = feats.clone()
        # tensors above 64K elements are stored as a stride-11 sample (key + "::11") to keep fixtures small
        rec["grads"] = {(k if v.numel() <= 65536 else k + "::11"): (v if v.numel() <= 65536 else v.flatten()[::11].clone())
                        for k, v in grads.items()}
        rec["state_dict_keys"] = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    else:
        rec["features_head"] = feats[:, :4, :64].clone()
        # a deterministic sample of every gradient (first 32 + strided 32 entries)
        rec["grad_sample"] = {k: torch.cat([v.flatten()[:32], v.flatten()[::max(1, v.numel() // 32)][:32]]).clone()
                              for k, v in grads.items()}
        # a wider strided sample: the 64-entry sample is too noisy an estimate of a tensor's rel-L2 error
        # once that error sits within 2x of the tolerance (conformer-large, 24 layers of bf16x3 GEMMs)
        rec["grad_sample_512"] = {k: v.flatten()[::max(1, v.numel() // 512)][:512].clone() for k, v in grads.items()}
    if w2v.kind == "conformer":
        rec["bn_running_mean_l0"] = m.state_dict()[
            "w2v_encoder.wav2vec2_conformer.encoder.layers.0.conv_module.batch_norm.running_mean"].clone()
        rec["bn_running_var_l0"] = m.state_dict()[
            "w2v_encoder.wav2vec2_conformer.encoder.layers.0.conv_module.batch_norm.running_var"].clone()
    return rec


TINY_BE = dict(gru_hidden=32, gru_layers=2, fc_hidden=(24,), out_size=48, smooth_sigma=1.5)
TINY_W2V = dict(hidden=48, layers=2, heads=4, ffn=96, pos_conv_k=16, pos_conv_groups=4)


def cases():
    yield ("tiny_base", O.BrainEncoderCfg(**TINY_BE), O.W2VCfg(kind="base", **TINY_W2V), 3, 100, 6, True, True)
    yield ("tiny_base_h0_uni", O.BrainEncoderCfg(**{**TINY_BE, "learnable_h0": True, "bidirectional": False,
                                                     "fc_hidden": (), "smooth_sigma": 0.3}),
           O.W2VCfg(kind="base", **TINY_W2V), 2, 64, 4, False, True)
    yield ("tiny_base_stable_ln", O.BrainEncoderCfg(**TINY_BE),
           O.W2VCfg(kind="base", stable_ln=True, **TINY_W2V), 3, 100, 6, True, True)
    yield ("tiny_conformer", O.BrainEncoderCfg(**{**TINY_BE, "out_size": 64}),
           O.W2VCfg(kind="conformer", hidden=64, layers=2, heads=4, ffn=96, dw_kernel=7), 3, 100, 6, True, True)
    if os.environ.get("B2T_GOLDEN_FULL", "1") == "1":
        yield ("base_cfg1", O.BrainEncoderCfg(), O.base_cfg(), 8, 500, 40, False, False)
        yield ("conformer_large_b2", O.BrainEncoderCfg(out_size=1024), O.conformer_large_cfg(), 2, 500, 40, False, False)


def main():
    outdir = os.path.join(REPO, "tests", "golden")
    os.makedirs(outdir, exist_ok=True)
    only = set(sys.argv[1:])
    for name, be, w2v, B, T, S, ragged, full in cases():
        if only and name not in only:
            continue
        torch.manual_seed(0)
        rec = run_case(name, be, w2v, B, T, S, ragged, full)
        path = os.path.join(outdir, name + ".pt")
        torch.save(rec, path)
        print(f"{name}: loss={rec['loss_train'].item():.6f} logits={tuple(rec['logits_train'].shape)} "
              f"no-grad params={len(rec['params_without_grad'])} -> {path} ({os.path.getsize(path) / 1e6:.2f} MB)")


if __name__ == "__main__":
    main()
