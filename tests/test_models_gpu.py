"""Drop-in boundary: model classes with the reference's constructor/forward/state_dict contract, on the GPU."""
import pytest
import torch

from tests.golden_util import golden_grad, load_case, rel_l2

pytestmark = pytest.mark.gpu


def build_from_golden(name):
    from wav2vec2forbrain_b200 import models
    from wav2vec2forbrain_b200.config import W2VConfig
    rec, _, _, sd, batch = load_case(name)
    be, w = rec["be"], rec["w2v"]
    ck = "golden/" + name

    def factory(**kw):
        d = dict(kind=w["kind"], hidden=w["hidden"], layers=w["layers"], heads=w["heads"], ffn=w["ffn"], vocab=w["vocab"],
                 pos_conv_k=w["pos_conv_k"], pos_conv_groups=w["pos_conv_groups"], dw_kernel=w["dw_kernel"],
                 hidden_act="gelu" if w["kind"] == "base" else "swish", hidden_dropout=0.0, attention_dropout=0.0,
                 activation_dropout=0.0, final_dropout=0.0, conv_dropout=0.0, layerdrop=0.0)
        d.update(kw)
        return W2VConfig(**d)
    models.register_architecture(ck, factory, be["out_size"])
    args = models.BrainEncoderArgs(encoder_gru_hidden_size=be["gru_hidden"], encoder_num_gru_layers=be["gru_layers"],
                                   encoder_bidirectional=be["bidirectional"], encoder_fc_hidden_sizes=tuple(be["fc_hidden"]),
                                   encoder_learnable_inital_state=be["learnable_h0"], gaussian_smooth_width=be["smooth_sigma"],
                                   encoder_dropout=0.0, w2v_do_stable_layer_norm=w["stable_ln"])
    brain = models.bfe_w_preprocessing_from_config(args, None, ck)
    if w["kind"] == "base":
        m = models.W2VBrainEncoderModel(args, brain, ck, None, True).cuda()
    else:
        m = models.W2VConformerBrainEncoderModel(brain, ck).cuda()
    return m, rec, sd, batch, models


def make_batch(models, batch):
    b = models.B2tSampleBatch(batch["input"], batch["target"])
    b.day_idxs, b.input_lens, b.target_lens = batch["day_idxs"], batch["input_lens"], batch["target_lens"]
    return b.cuda()


@pytest.mark.parametrize("name", ["tiny_base", "tiny_base_stable_ln", "tiny_base_h0_uni", "tiny_conformer"])
def test_state_dict_and_training_step(name):
    m, rec, sd, batch, models = build_from_golden(name)
    assert {k: tuple(v.shape) for k, v in m.state_dict().items()} == rec["state_dict_keys"]
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    conformer = rec["w2v"]["kind"] == "conformer"
    # eval mode / no_grad path first (the train pass below moves the conformer's BatchNorm running statistics, and the
    # golden's eval outputs were recorded before its train pass): greedy decode ids identical to the reference's argmax
    m.eval()
    with torch.no_grad():
        ev = m.forward(make_batch(models, batch))
    assert rel_l2(ev.logits.cpu(), rec["logits_eval"]) < 2e-4
    assert torch.equal(ev.logits.argmax(-1).cpu().to(torch.int16), rec["argmax_eval"])
    # sub-module calls used by the analysis helpers (latent_analysis_util.py:196-201)
    with torch.no_grad():
        enc = m.brain_encoder.forward(make_batch(models, batch))
        res = m.w2v_encoder.forward(enc.logits)
        lg = res if conformer else res[0]      # conformer: logits only (w2v_conformer_custom_feat_extractor.py:67-76)
        assert isinstance(res, torch.Tensor) == conformer
    assert rel_l2(enc.logits.cpu(), rec["features"]) < 2e-4
    assert rel_l2(lg.cpu(), rec["logits_eval"]) < 2e-4
    m.train()
    out = m.forward(make_batch(models, batch))
    assert isinstance(out.metrics["ctc_loss"], float)
    out.loss.backward()
    assert rel_l2(out.logits.detach().cpu(), rec["logits_train"]) < 2e-4
    assert abs(out.loss.item() - rec["loss_train"].item()) < 2e-4 * abs(rec["loss_train"].item())
    if conformer:
        # W2VConformerBrainEncoderModel.forward returns neither logit_lens nor hidden_states
        # (w2v_conformer_custom_feat_extractor.py:55-59)
        assert out.logit_lens is None and out.hidden_states is None
    else:
        assert torch.equal(out.logit_lens.cpu(), rec["logit_lens"])
        assert out.hidden_states.shape[:2] == out.logits.shape[:2]
    got = {k: p.grad for k, p in m.named_parameters() if p.grad is not None}
    assert sorted(k for k, p in m.named_parameters() if p.grad is None) == rec["params_without_grad"]
    gmax = max(rec["grad_norm"].values())
    for k in got:
        ref, sl = golden_grad(rec, k)
        err = (sl(got[k].cpu()).double() - ref.double()).norm().item()
        assert err <= 1e-3 * ref.double().norm().item() + 1e-6 * gmax, k


def test_adam_step_moves_only_trained_parameters():
    m, rec, sd, batch, models = build_from_golden("tiny_base")
    m.load_state_dict(sd, strict=True)
    before = {k: v.detach().clone() for k, v in m.named_parameters()}
    opt = torch.optim.Adam([{"params": m.brain_encoder.parameters(), "lr": 1e-3},
                            {"params": m.w2v_encoder.parameters(), "lr": 1e-4}], weight_decay=1e-4)
    m.train()
    for _ in range(2):
        opt.zero_grad()
        out = m.forward(make_batch(models, batch))
        out.loss.backward()
        torch.nn.utils.clip_grad_norm_(m.parameters(), 10.0)
        opt.step()
    moved = {k for k, v in m.named_parameters() if not torch.equal(v, before[k])}
    assert moved == {k for k in before} - set(rec["params_without_grad"])


DROPOUT_FD_KEYS = {
    "tiny_base": ["brain_encoder.day_weights", "brain_encoder.neural_decoder.encoder.gru.weight_hh_l0",
                  "w2v_encoder.wav2vec2.encoder.layers.0.feed_forward.intermediate_dense.weight",
                  "w2v_encoder.wav2vec2.encoder.layers.1.attention.q_proj.weight", "w2v_encoder.lm_head.weight"],
    # conformer dropout sites: half-FFN epilogue dropout (ffn1/ffn2 output_dense), attention dropout inside the
    # attention kernel and on linear_out, conv-module dropout behind pointwise_conv2, activation dropout
    "tiny_conformer": ["brain_encoder.neural_decoder.encoder.gru.weight_ih_l1",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.0.ffn1.intermediate_dense.weight",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.0.ffn2.output_dense.weight",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.1.self_attn.linear_q.weight",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.1.self_attn.linear_v.weight",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.0.conv_module.pointwise_conv1.weight",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.1.conv_module.depthwise_conv.weight",
                       "w2v_encoder.wav2vec2_conformer.encoder.layers.0.conv_module.pointwise_conv2.weight",
                       "w2v_encoder.lm_head.weight"],
}


@pytest.mark.parametrize("case", ["tiny_base", "tiny_conformer"])
def test_dropout_training_is_finite_and_seeded(case):
    """train mode with the reference's dropout probabilities: masks differ per step, the same seed regenerates the
    same masks (mask round trip: forward twice bit-identical per trial, backward regenerates them from the seed),
    loss stays finite, and the whole dropout-aware backward is checked by directional finite differences."""
    from wav2vec2forbrain_b200 import engine, ops
    from tests.engine_util import product_configs
    from dataclasses import replace
    rec, _, _, sd, batch = load_case(case)
    be, w2v = product_configs(rec)
    be = replace(be, dropout=0.3)
    w2v = replace(w2v, hidden_dropout=0.1, attention_dropout=0.1, activation_dropout=0.1, final_dropout=0.1,
                  conv_dropout=0.1)
    eng = engine.Engine(be, w2v, "bf16x3")
    P = {k: v.cuda() for k, v in sd.items()}
    x, day = batch["input"].cuda(), batch["day_idxs"].to("cuda", torch.int32)
    lens = eng.processed_lens(batch["input_lens"]).cuda()
    tg, tl = batch["target"].cuda(), batch["target_lens"].cuda()

    def loss_at(Pd, seed):
        logits, _, _, saved = eng.forward(Pd, x, day, True, seed)
        loss, nll, dl = ops.ctc_fwd_bwd(logits, tg, lens, tl)
        loss_at.nll = nll.clone()
        return loss, saved, dl
    l1, saved, dl = loss_at(P, 5)
    n1 = loss_at.nll
    l1b, _, _ = loss_at(P, 5)
    n1b = loss_at.nll
    l2, _, _ = loss_at(P, 6)
    # the forward is deterministic per trial (same seed -> bit-identical negative log-likelihoods); the batch
    # mean is an atomicAdd over trials, so the scalar may differ by an ulp with the arrival order
    assert torch.isfinite(l1) and torch.equal(n1, n1b) and abs(l1.item() - l1b.item()) <= 2e-6 * abs(l1.item())
    assert abs(l1.item() - l2.item()) > 1e-5 * abs(l1.item())
    G = eng.backward(P, saved, dl)
    # directional derivative along the gradient of a few parameter tensors
    for k in DROPOUT_FD_KEYS[case]:
        g = G[k]
        eps = 1e-2 / g.norm().item()
        Pp, Pm = dict(P), dict(P)
        Pp[k] = P[k] + eps * g
        Pm[k] = P[k] - eps * g
        fd = (loss_at(Pp, 5)[0].item() - loss_at(Pm, 5)[0].item()) / (2 * eps)
        an = (g * g).sum().item()
        assert abs(fd - an) < 0.05 * abs(an) + 1e-7, (k, fd, an)


def test_layerdrop_skipped_layer_backward():
    """LayerDrop through the model classes (transformers modeling_wav2vec2.py:702-704): the layers the shared host
    generator skips receive no gradient, every other parameter's gradient equals the oracle's with the same
    layers skipped."""
    from oracle import b2t_oracle as O
    from wav2vec2forbrain_b200 import models
    from wav2vec2forbrain_b200.config import W2VConfig
    rec, be_o, w_o, sd, batch = load_case("tiny_base")
    w = rec["w2v"]
    ck = "golden/tiny_base_layerdrop"
    models.register_architecture(ck, lambda **kw: W2VConfig(**dict(dict(
        kind="base", hidden=w["hidden"], layers=w["layers"], heads=w["heads"], ffn=w["ffn"], vocab=w["vocab"],
        pos_conv_k=w["pos_conv_k"], pos_conv_groups=w["pos_conv_groups"], hidden_dropout=0.0, attention_dropout=0.0,
        activation_dropout=0.0, final_dropout=0.0, layerdrop=0.5), **kw)), rec["be"]["out_size"])
    be = rec["be"]
    args = models.BrainEncoderArgs(encoder_gru_hidden_size=be["gru_hidden"], encoder_num_gru_layers=be["gru_layers"],
                                   encoder_fc_hidden_sizes=tuple(be["fc_hidden"]), gaussian_smooth_width=be["smooth_sigma"])
    brain = models.bfe_w_preprocessing_from_config(args, None, ck)
    m = models.W2VBrainEncoderModel(args, brain, ck, None, True).cuda()
    m.load_state_dict(sd, strict=True)
    m.train()
    rt = m.runtime
    seen = []
    real = rt.layerdrop_skips
    rt.layerdrop_skips = lambda *a, **k: seen.append(real(*a, **k)) or seen[-1]
    tried = 0
    while True:     # draw until the generator skips some but not all layers (p = 0.5, 2 layers)
        m.zero_grad(set_to_none=True)
        out = m.forward(make_batch(models, batch))
        tried += 1
        if 0 < len(seen[-1]) < w["layers"] or tried > 20:
            break
    skips = tuple(seen[-1])
    assert 0 < len(skips) < w["layers"]
    out.loss.backward()
    p = {k: (v.detach().clone().requires_grad_(not O.is_buffer(k)) if v.is_floating_point() else v) for k, v in sd.items()}
    ref = O.forward(batch, p, be_o, w_o, training=True, skip_layers=[i in skips for i in range(w["layers"])], gru_impl="loop")
    ref.loss.backward()
    ref_grads = {k: v.grad for k, v in p.items() if torch.is_tensor(v) and v.is_floating_point() and v.grad is not None}
    got = {k: q.grad for k, q in m.named_parameters() if q.grad is not None}
    for i in skips:
        assert not any(f".layers.{i}." in k for k in got)
    assert set(got) == set(ref_grads)
    assert rel_l2(out.logits.detach().cpu(), ref.logits) < 2e-4
    gmax = max(v.double().norm().item() for v in ref_grads.values())
    for k, g in ref_grads.items():
        err = (got[k].double().cpu() - g.double()).norm().item()
        assert err <= 1e-3 * g.double().norm().item() + 1e-6 * gmax, k
