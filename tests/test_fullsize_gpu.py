"""Parity at the BENCHMARKED configurations (BASELINE.json configs[1..3]) through the product model classes.

Every number bench.py prints sits on one of these shapes: wav2vec2-base B=64 (full fine-tuning), the same model
B=128 with the w2v weight gradients skipped (frozen), conformer-large B=32.  Batch size changes the code path
(GEMM tile shape from the wave count, split-K of the weight gradients, GRU cluster waves, attention grid), so the
small goldens do not cover them.  The pinned CPU oracle (oracle/b2t_oracle.py, tests/test_oracle_golden.py) runs
the same step on the host cores inside the test; tolerance = north_star's 1e-3 relative on logits / loss / every
parameter gradient (absolute floor 1e-6 * largest gradient norm for gradients whose true value is ~0), greedy
decode ids identical in eval mode.  The worst per-tensor errors are written to a JSON record
(gpurun_out/parity_r02.json on the GPU box; committed as profiles/parity_r02.json).
"""
import json
import os
import time

import pytest
import torch

from oracle import b2t_oracle as O
from tests.golden_util import rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-3   # BASELINE.json north_star tolerance
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CASES = {
    # name: (kind, per-GPU batch, frozen, ragged batch)
    "base_b64": ("base", 64, False, True),
    "frozen_b128": ("base", 128, True, False),
    "conformer_large_b32": ("conformer", 32, False, False),
}


def _record(name, entry):
    path = os.environ.get("B2T_PARITY_OUT")
    if path is None:
        d = os.path.join(REPO, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "parity_r02.json")
    rec = {}
    if os.path.exists(path):
        try:
            with open(path) as f:
                rec = json.load(f)
        except Exception:
            rec = {}
    rec[name] = entry
    with open(path, "w") as f:
        json.dump(rec, f, indent=1, sort_keys=True)


def _build(kind, frozen):
    from wav2vec2forbrain_b200 import models
    nodrop = dict(hidden_dropout=0.0, attention_dropout=0.0, activation_dropout=0.0, final_dropout=0.0, conv_dropout=0.0,
                  layerdrop=0.0)
    args = models.BrainEncoderArgs(encoder_gru_hidden_size=512, encoder_num_gru_layers=3, encoder_fc_hidden_sizes=(256,),
                                   gaussian_smooth_width=1.5, encoder_dropout=0.0)
    if kind == "conformer":
        ck = "facebook/wav2vec2-conformer-rope-large-960h-ft"
        brain = models.bfe_w_preprocessing_from_config(args, None, ck)
        m = models.W2VConformerBrainEncoderModel(brain, ck, **nodrop)
        be_o, w_o = O.BrainEncoderCfg(out_size=1024), O.conformer_large_cfg()
    else:
        ck = "facebook/wav2vec2-base-960h"
        brain = models.bfe_w_preprocessing_from_config(args, None, ck)
        m = models.W2VBrainEncoderModel(args, brain, ck, None, True, **nodrop)
        be_o, w_o = O.BrainEncoderCfg(out_size=768), O.base_cfg()
    m = m.cuda()
    m.runtime.skip_w2v_wgrad = frozen
    return m, models, be_o, w_o


@pytest.mark.parametrize("name", list(CASES))
def test_benchmarked_config_matches_oracle(name):
    kind, B, frozen, ragged = CASES[name]
    torch.set_num_threads(os.cpu_count() or 1)
    m, models, be_o, w_o = _build(kind, frozen)
    sd = O.make_state_dict(be_o, w_o, seed=42)
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    hb = O.make_batch(B, 500, 40, seed=1234, ragged=ragged)
    batch = models.B2tSampleBatch(hb["input"], hb["target"])
    batch.day_idxs, batch.input_lens, batch.target_lens = hb["day_idxs"], hb["input_lens"], hb["target_lens"]
    batch = batch.cuda()

    # ---- eval mode first (BatchNorm running statistics still at their initial values on both sides) ----
    t0 = time.time()
    with torch.no_grad():
        ref_eval = O.forward(hb, sd, be_o, w_o, training=False, gru_impl="aten")
    m.eval()
    with torch.no_grad():
        ev = m.forward(batch)
    from wav2vec2forbrain_b200 import ops
    am, ids, lens = ops.greedy_decode(ev.logits)
    e_eval = rel_l2(ev.logits.cpu(), ref_eval.logits)
    assert e_eval < TOL, e_eval
    ref_am = ref_eval.logits.argmax(-1)
    n_diff = int((am.cpu().long() != ref_am).sum())
    assert n_diff == 0, f"{n_diff} of {ref_am.numel()} greedy frames differ"
    ref_ids = O.greedy_decode(ref_eval.logits)
    ids_c, lens_c = ids.cpu(), lens.cpu()
    for b in range(B):
        assert ids_c[b, :lens_c[b]].tolist() == ref_ids[b]

    # ---- one training step: logits, loss, every parameter gradient ----
    ref, ref_grads = O.train_step_grads(hb, sd, be_o, w_o, gru_impl="aten")
    t_oracle = time.time() - t0
    m.train()
    out = m.forward(batch)
    out.loss.backward()
    torch.cuda.synchronize()
    e_logits = rel_l2(out.logits.detach().cpu(), ref.logits)
    e_loss = abs(out.loss.item() - ref.loss.item()) / abs(ref.loss.item())
    got = {k: p.grad for k, p in m.named_parameters() if p.grad is not None}
    if frozen:
        assert all(k.startswith("brain_encoder.") for k in got), "frozen mode: no w2v parameter may receive a gradient"
        want = {k for k in ref_grads if k.startswith("brain_encoder.")}
    else:
        want = set(ref_grads)
    assert set(got) == want, sorted(set(got) ^ want)[:8]
    gmax = max(v.double().norm().item() for v in ref_grads.values())
    worst, worst_key, errs = 0.0, None, {}
    for k in sorted(want):
        g = ref_grads[k].double()
        err = (got[k].double().cpu() - g).norm().item()
        nrm = g.norm().item()
        r = err / (nrm + 1e-6 * gmax / TOL)        # relative error with the absolute floor folded in: pass <=> r <= TOL
        errs[k] = r
        if r > worst:
            worst, worst_key = r, k
    top = sorted(errs.items(), key=lambda kv: -kv[1])[:5]
    _record(name, {"config": {"kind": kind, "per_gpu_batch": B, "frozen_w2v": frozen, "ragged": ragged, "T": 500,
                              "precision": "bf16x3"},
                   "tolerance": TOL, "logits_rel_l2_train": e_logits, "logits_rel_l2_eval": e_eval, "loss_rel": e_loss,
                   "loss": out.loss.item(), "oracle_loss": ref.loss.item(), "greedy_frames": int(ref_am.numel()),
                   "greedy_frames_differing": n_diff, "parameter_gradients_checked": len(want),
                   "worst_gradient_rel_err": worst, "worst_gradient": worst_key,
                   "top5_gradient_rel_err": [[k, v] for k, v in top], "oracle_cpu_seconds": t_oracle})
    assert e_logits < TOL, e_logits
    assert e_loss < TOL, e_loss
    bad = [(k, v) for k, v in errs.items() if v > TOL]
    assert not bad, bad[:8]
    print(f"{name}: logits {e_logits:.2e} / eval {e_eval:.2e}, loss {e_loss:.2e}, worst grad {worst:.2e} ({worst_key}), "
          f"{len(want)} gradients, oracle {t_oracle:.1f} s")
