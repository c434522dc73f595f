"""Shared by the CPU (emulated ABI) and GPU (real kernels) engine parity tests."""
import torch

from tests.golden_util import golden_grad, load_case, rel_l2
from wav2vec2forbrain_b200.config import BrainEncoderConfig, W2VConfig


def product_configs(rec):
    be = BrainEncoderConfig(**{k: rec["be"][k] for k in ("gru_hidden", "gru_layers", "bidirectional", "fc_hidden", "fc_act",
                                                         "out_size", "unfold_len", "unfold_stride", "smooth_sigma", "dropout",
                                                         "learnable_h0")})
    w = rec["w2v"]
    w2v = W2VConfig(kind=w["kind"], hidden=w["hidden"], layers=w["layers"], heads=w["heads"], ffn=w["ffn"], vocab=w["vocab"],
                    ln_eps=w["ln_eps"], pos_conv_k=w["pos_conv_k"], pos_conv_groups=w["pos_conv_groups"],
                    stable_ln=w["stable_ln"], dw_kernel=w["dw_kernel"], rope_base=w["rope_base"],
                    hidden_act="gelu" if w["kind"] == "base" else "swish",
                    hidden_dropout=0.0, attention_dropout=0.0, activation_dropout=0.0, final_dropout=0.0, conv_dropout=0.0,
                    layerdrop=0.0)
    return be, w2v


def run_engine_train_step(engine_mod, rec, sd, batch, device, precision="bf16x3"):
    """forward(training) + CTC + backward through the engine.  -> (logits, loss, logit_lens, grads, feats)"""
    be, w2v = product_configs(rec)
    eng = engine_mod.Engine(be, w2v, precision=precision)
    P = {k: v.to(device) for k, v in sd.items()}
    x = batch["input"].to(device)
    day = batch["day_idxs"].to(device=device, dtype=torch.int32)
    logits, hidden, feats, saved = eng.forward(P, x, day, training=True, seed=1)
    lens = eng.processed_lens(batch["input_lens"]).to(device)
    loss, nll, dlogits = engine_mod.ops.ctc_fwd_bwd(logits, batch["target"].to(device), lens,
                                                    batch["target_lens"].to(device))
    G = eng.backward(P, saved, dlogits)
    return logits, loss, lens, G, feats, eng, P


def check_against_golden(rec, logits, loss, lens, G, tol_logits, tol_grad, feats=None):
    assert rel_l2(logits.cpu(), rec["logits_train"]) < tol_logits
    assert abs(loss.item() - rec["loss_train"].item()) < tol_logits * abs(rec["loss_train"].item())
    if rec["logit_lens"] is not None:
        assert torch.equal(lens.cpu(), rec["logit_lens"])
    if "grads" in rec:
        ref_keys = {k.split("::")[0] for k in rec["grads"]}
        assert set(G) == ref_keys, (sorted(set(G) ^ ref_keys))
        gmax = max(rec["grad_norm"].values())
        worst = 0.0
        for k in sorted(ref_keys):
            ref, sl = golden_grad(rec, k)
            got = sl(G[k].cpu())
            err = (got.double() - ref.double()).norm().item()
            bound = tol_grad * ref.double().norm().item() + 1e-3 * tol_grad * gmax
            assert err <= bound, (k, err, ref.double().norm().item())
            worst = max(worst, err / max(ref.double().norm().item(), 1e-30))
        return worst
    # sampled gradients (full-size cases)
    assert set(G) == set(rec["grad_norm"]), sorted(set(G) ^ set(rec["grad_norm"]))
    gmax = max(rec["grad_norm"].values())
    wide = "grad_sample_512" in rec
    for k, ref in rec["grad_sample_512" if wide else "grad_sample"].items():
        v = G[k].cpu()
        if wide:
            got = v.flatten()[::max(1, v.numel() // 512)][:512]
        else:
            got = torch.cat([v.flatten()[:32], v.flatten()[::max(1, v.numel() // 32)][:32]])
        assert (got - ref).norm().item() <= tol_grad * ref.norm().item() + 1e-3 * tol_grad * gmax, k
        assert abs(v.double().norm().item() - rec["grad_norm"][k]) <= tol_grad * rec["grad_norm"][k] + 1e-3 * tol_grad * gmax, k
    return None
