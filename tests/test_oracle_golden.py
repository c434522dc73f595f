"""Pins the CPU oracle (oracle/b2t_oracle.py) to outputs of the live reference recorded in
tests/golden/*.pt by oracle/gen_golden.py (the reference has no tests of its own)."""
import pytest
import torch

from oracle import b2t_oracle as O
from tests.golden_util import golden_grad, load_case, rel_l2

TINY = ["tiny_base", "tiny_base_h0_uni", "tiny_base_stable_ln", "tiny_conformer"]


@pytest.mark.parametrize("name", TINY)
def test_tiny_forward_and_grads(name):
    rec, be, w2v, sd, batch = load_case(name)
    out, grads = O.train_step_grads(batch, sd, be, w2v, gru_impl="loop")
    assert rel_l2(out.logits, rec["logits_train"]) < 2e-5
    assert abs(out.loss.item() - rec["loss_train"].item()) < 2e-5 * abs(rec["loss_train"].item())
    if rec["logit_lens"] is not None:
        assert torch.equal(out.logit_lens, rec["logit_lens"])
    # same set of parameters receives a gradient
    ref_keys = {k.split("::")[0] for k in rec["grads"]}
    assert set(grads) == ref_keys
    gmax = max(rec["grad_norm"].values())
    for k in sorted(ref_keys):
        ref, sl = golden_grad(rec, k)
        got = sl(grads[k])
        err = (got.double() - ref.double()).norm().item()
        assert err <= 2e-4 * ref.double().norm().item() + 1e-6 * gmax, k


@pytest.mark.parametrize("name", TINY)
def test_tiny_eval_and_greedy(name):
    rec, be, w2v, sd, batch = load_case(name)
    with torch.no_grad():
        out = O.forward(batch, sd, be, w2v, training=False, gru_impl="aten", ctc_impl="loop")
    assert rel_l2(out.logits, rec["logits_eval"]) < 2e-5
    assert abs(out.loss.item() - rec["loss_eval"].item()) < 1e-4 * abs(rec["loss_eval"].item())
    assert torch.equal(out.logits.argmax(-1).to(torch.int16), rec["argmax_eval"])
    feats, _ = O.brain_encoder(batch, sd, be)
    assert rel_l2(feats, rec["features"]) < 2e-5


def test_state_dict_keys_match_reference():
    rec, be, w2v, sd, _ = load_case("tiny_base")
    assert {k: tuple(v.shape) for k, v in sd.items()} == rec["state_dict_keys"]
    rec, be, w2v, sd, _ = load_case("tiny_conformer")
    assert {k: tuple(v.shape) for k, v in sd.items()} == rec["state_dict_keys"]


def test_conformer_bn_running_stats():
    rec, be, w2v, sd, batch = load_case("tiny_conformer")
    stats = {}
    feats, _ = O.brain_encoder(batch, sd, be)
    O.conformer_encoder(feats, sd, w2v, training=True, bn_stats_out=stats)
    p = "w2v_encoder.wav2vec2_conformer.encoder.layers.0.conv_module."
    mean, var_unbiased = stats[p]
    rm = 0.9 * sd[p + "batch_norm.running_mean"] + 0.1 * mean
    rv = 0.9 * sd[p + "batch_norm.running_var"] + 0.1 * var_unbiased
    assert rel_l2(rm, rec["bn_running_mean_l0"]) < 1e-5
    assert rel_l2(rv, rec["bn_running_var_l0"]) < 1e-5


def test_base_cfg1_full_size():
    """BASELINE.json configs[0]: B=8, T=500, GRU 3x512 bi + FC[256] + wav2vec2-base arch."""
    rec, be, w2v, sd, batch = load_case("base_cfg1")
    out, grads = O.train_step_grads(batch, sd, be, w2v, gru_impl="aten")
    assert rel_l2(out.logits, rec["logits_train"]) < 5e-5
    assert abs(out.loss.item() - rec["loss_train"].item()) < 5e-5 * abs(rec["loss_train"].item())
    assert out.logits.shape == (8, 118, 32) and int(out.logit_lens[0]) == 117
    assert set(grads) == set(rec["grad_norm"])
    gmax = max(rec["grad_norm"].values())
    for k, ref in rec["grad_sample"].items():
        v = grads[k]
        got = torch.cat([v.flatten()[:32], v.flatten()[::max(1, v.numel() // 32)][:32]])
        assert (got - ref).norm().item() <= 1e-3 * ref.norm().item() + 1e-6 * gmax, k
        assert abs(v.double().norm().item() - rec["grad_norm"][k]) <= 1e-3 * rec["grad_norm"][k] + 1e-6 * gmax, k
    with torch.no_grad():
        ev = O.forward(batch, sd, be, w2v, training=False, gru_impl="aten")
    assert torch.equal(ev.logits.argmax(-1).to(torch.int16), rec["argmax_eval"])


@pytest.mark.slow
def test_conformer_large_full_size():
    rec, be, w2v, sd, batch = load_case("conformer_large_b2")
    with torch.no_grad():
        out = O.forward(batch, sd, be, w2v, training=True, gru_impl="aten")
    assert rel_l2(out.logits, rec["logits_train"]) < 5e-5
    assert abs(out.loss.item() - rec["loss_train"].item()) < 5e-5 * abs(rec["loss_train"].item())


def test_ctc_loop_matches_aten():
    g = torch.Generator().manual_seed(3)
    logits = torch.randn(4, 30, 32, generator=g)
    tgt = torch.randint(4, 32, (4, 9), generator=g)
    tgt[1, 3] = tgt[1, 4]  # repeated label
    tl = torch.tensor([9, 6, 1, 0])
    ll = torch.tensor([30, 25, 29, 7])
    for b in range(4):
        tgt[b, tl[b]:] = 0
    a = O.ctc_loss(logits, tgt, ll, tl, "aten")
    b_ = O.ctc_loss(logits, tgt, ll, tl, "loop")
    assert abs(a.item() - b_.item()) < 1e-5
    # infeasible alignment: zero_infinity
    ll2 = torch.tensor([5, 25, 29, 7])
    a2 = O.ctc_loss(logits, tgt, ll2, tl, "aten")
    b2 = O.ctc_loss(logits, tgt, ll2, tl, "loop")
    assert abs(a2.item() - b2.item()) < 1e-5


def test_greedy_decode_collapse():
    logits = torch.full((1, 7, 32), -5.0)
    for t, k in enumerate([0, 5, 5, 0, 5, 7, 7]):
        logits[0, t, k] = 1.0
    assert O.greedy_decode(logits) == [[5, 5, 7]]
