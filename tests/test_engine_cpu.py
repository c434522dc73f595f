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
