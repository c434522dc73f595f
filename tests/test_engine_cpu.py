"""Host-side schedule of the engine (operand views, epilogue wiring, hand-derived backward)
checked against the live-reference goldens through the CPU emulator of the C ABI."""
import pytest
import torch

from tests import abi_emulator
from tests.engine_util import check_against_golden, run_engine_train_step
from tests.golden_util import load_case, rel_l2


@pytest.fixture()
def engine_mod(monkeypatch):
    from wav2vec2forbrain_b200 import engine
    monkeypatch.setattr(engine, "ops", abi_emulator)
    return engine


@pytest.mark.parametrize("name", ["tiny_base", "tiny_base_h0_uni", "tiny_base_stable_ln", "tiny_conformer"])
def test_engine_schedule_matches_reference(engine_mod, name):
    rec, _, _, sd, batch = load_case(name)
    logits, loss, lens, G, feats, _, _ = run_engine_train_step(engine_mod, rec, sd, batch, "cpu")
    assert rel_l2(feats.view(rec["features"].shape), rec["features"]) < 5e-5
    check_against_golden(rec, logits, loss, lens, G, tol_logits=5e-5, tol_grad=3e-4)


@pytest.mark.parametrize("name", ["tiny_base", "tiny_conformer"])
def test_frozen_w2v_backward_skips_weight_gradients_only(engine_mod, name):
    """unfreeze_strategy == "brain_encoder" (b2t_gru_w2v_experiment.py:129-133): with Engine._skip_wgrad the
    encoder backward issues no weight-gradient GEMM, and the data gradient it hands to the brain encoder — hence
    every brain-encoder gradient — is unchanged."""
    from tests.engine_util import product_configs
    rec, _, _, sd, batch = load_case(name)
    be, w2v = product_configs(rec)
    eng = engine_mod.Engine(be, w2v, "bf16x3")
    x, day = batch["input"], batch["day_idxs"].to(torch.int32)
    logits, _, _, saved = eng.forward(sd, x, day, training=True, seed=1)
    lens = eng.processed_lens(batch["input_lens"])
    _, _, dl = engine_mod.ops.ctc_fwd_bwd(logits, batch["target"], lens, batch["target_lens"])
    G_full = eng.backward(sd, saved, dl)
    calls = []
    real = engine_mod.ops.gemm
    engine_mod.ops.gemm = lambda a, b, **kw: (calls.append((a.mn_major, b.mn_major)), real(a, b, **kw))[1]
    try:
        G: dict = {}
        eng._skip_wgrad = True
        dh = eng.head_backward(sd, saved["head"], dl, G)
        if w2v.kind == "base":
            dfeats = eng.w2v_base_backward(sd, saved["enc"], dh, G)
        else:
            from wav2vec2forbrain_b200 import conformer
            dfeats = conformer.backward(eng, sd, saved["enc"], dh, G)
        eng._skip_wgrad = False
        n_enc = len(calls)
        Gb: dict = {}
        eng.brain_backward(sd, saved["brain"], dfeats, Gb)
    finally:
        engine_mod.ops.gemm = real
        eng._skip_wgrad = False
    # weight gradients are the GEMMs with both operands contracted over rows (MN-major x MN-major)
    assert not any(a and b for a, b in calls[:n_enc])
    assert any(a and b for a, b in calls[n_enc:])
    # (the depthwise-conv weight gradient falls out of the fused conv-module backward kernel and is simply dropped)
    assert not any(k.endswith(".weight") and v is not None and v.dim() >= 2 and "depthwise" not in k for k, v in G.items())
    for k, v in Gb.items():
        assert torch.equal(v, G_full[k]), k


@pytest.mark.parametrize("name", ["tiny_base", "tiny_conformer"])
def test_flat_gradient_arena_through_the_model_classes(engine_mod, monkeypatch, name):
    """The model classes on the emulator: with the flat gradient arena every .grad is a view of one buffer at the
    parameter's slot (no copy at the end of the node except for a handful of tiny GRU bias tensors), and the values are
    bit-identical to the arena-less path."""
    from wav2vec2forbrain_b200 import models
    from wav2vec2forbrain_b200.config import W2VConfig
    monkeypatch.setattr(models, "_EMULATED", True)
    rec, _, _, sd, batch = load_case(name)
    be, w = rec["be"], rec["w2v"]
    ck = "golden/arena_" + name
    models.register_architecture(ck, lambda **kw: W2VConfig(**dict(dict(
        kind=w["kind"], hidden=w["hidden"], layers=w["layers"], heads=w["heads"], ffn=w["ffn"], vocab=w["vocab"],
        pos_conv_k=w["pos_conv_k"], pos_conv_groups=w["pos_conv_groups"], dw_kernel=w["dw_kernel"],
        hidden_act="gelu" if w["kind"] == "base" else "swish", hidden_dropout=0.0, attention_dropout=0.0,
        activation_dropout=0.0, final_dropout=0.0, conv_dropout=0.0, layerdrop=0.0), **kw)), be["out_size"])
    args = models.BrainEncoderArgs(encoder_gru_hidden_size=be["gru_hidden"], encoder_num_gru_layers=be["gru_layers"],
                                   encoder_fc_hidden_sizes=tuple(be["fc_hidden"]), gaussian_smooth_width=be["smooth_sigma"])
    brain = models.B2P2TModel(args, models.B2TBrainFeatureExtractor(args, ck, 256 * 32), wav2vec_checkpoint=ck)
    m = models.W2VBrainEncoderModel(args, brain, ck, None, True) if w["kind"] == "base" else \
        models.W2VConformerBrainEncoderModel(brain, ck)
    m.load_state_dict(sd, strict=True)
    m.train()
    mb = models.B2tSampleBatch(batch["input"], batch["target"])
    mb.day_idxs, mb.input_lens, mb.target_lens = batch["day_idxs"], batch["input_lens"], batch["target_lens"]

    def grads(use_arena):
        m.runtime.use_arena = use_arena
        m.zero_grad(set_to_none=True)
        m.runtime.step = 0
        m.forward(mb).loss.backward()
        return {k: p.grad for k, p in m.named_parameters() if p.grad is not None}
    plain = {k: v.clone() for k, v in grads(False).items()}
    assert m.runtime.engine.arena_copies == 0
    flat = grads(True)
    arena = m.runtime.grad_arena()
    assert set(flat) == set(plain)
    for k, g in flat.items():
        assert arena.owns(k, g), k                      # .grad IS the arena slot
        assert torch.equal(g, plain[k]), k
    n_gru_bias_hh = sum(1 for k in flat if ".gru.bias_hh_" in k)
    assert m.runtime.engine.arena_copies <= n_gru_bias_hh
    # a second step reuses the same slots (stable gradient addresses for the optimizer tables)
    ptrs = {k: g.data_ptr() for k, g in flat.items()}
    flat2 = grads(True)
    assert {k: g.data_ptr() for k, g in flat2.items()} == ptrs
    # parameters that still hold a .grad (no zero_grad): the engine must not overwrite their slots -> accumulation works
    before = {k: g.clone() for k, g in flat2.items()}
    m.runtime.step = 0
    m.forward(mb).loss.backward()
    for k, p in m.named_parameters():
        if p.grad is not None:
            assert torch.allclose(p.grad, 2 * before[k], rtol=1e-6, atol=1e-12), k


@pytest.mark.parametrize("name", ["tiny_base", "tiny_conformer"])
def test_training_step_leaves_no_reference_cycles(engine_mod, monkeypatch, name):
    """dist.tune_gc() makes garbage collections rare on purpose, so a step must not create cyclic garbage: a cycle that
    holds activations would keep device memory alive until the next collection and grow the allocator pool mid-run."""
    import gc
    from wav2vec2forbrain_b200 import models
    from wav2vec2forbrain_b200.config import W2VConfig
    monkeypatch.setattr(models, "_EMULATED", True)
    rec, _, _, sd, batch = load_case(name)
    be, w = rec["be"], rec["w2v"]
    ck = "golden/cycles_" + name
    models.register_architecture(ck, lambda **kw: W2VConfig(**dict(dict(
        kind=w["kind"], hidden=w["hidden"], layers=w["layers"], heads=w["heads"], ffn=w["ffn"], vocab=w["vocab"],
        pos_conv_k=w["pos_conv_k"], pos_conv_groups=w["pos_conv_groups"], dw_kernel=w["dw_kernel"],
        hidden_act="gelu" if w["kind"] == "base" else "swish", layerdrop=0.3), **kw)), be["out_size"])
    args = models.BrainEncoderArgs(encoder_gru_hidden_size=be["gru_hidden"], encoder_num_gru_layers=be["gru_layers"],
                                   encoder_fc_hidden_sizes=tuple(be["fc_hidden"]), gaussian_smooth_width=be["smooth_sigma"])
    brain = models.B2P2TModel(args, models.B2TBrainFeatureExtractor(args, ck, 256 * 32), wav2vec_checkpoint=ck)
    m = models.W2VBrainEncoderModel(args, brain, ck, None, True) if w["kind"] == "base" else \
        models.W2VConformerBrainEncoderModel(brain, ck)
    m.train()
    mb = models.B2tSampleBatch(batch["input"], batch["target"])
    mb.day_idxs, mb.input_lens, mb.target_lens = batch["day_idxs"], batch["input_lens"], batch["target_lens"]
    opt = torch.optim.SGD([p for p in m.parameters() if p.requires_grad], lr=1e-4)

    def step():
        opt.zero_grad(set_to_none=True)
        out = m.forward(mb)
        out.loss.backward()
        _ = out.metrics["ctc_loss"]          # the lazy metrics dict is read as the Trainer reads it
        opt.step()
    for _ in range(2):
        step()
    gc.collect()
    was_enabled = gc.isenabled()
    gc.disable()
    try:
        for _ in range(3):
            step()
        found = gc.collect()
    finally:
        if was_enabled:
            gc.enable()
    assert found == 0, f"{found} unreachable objects after 3 training steps"
