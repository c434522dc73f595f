"""Conformer kernels (csrc/conformer.cu) against fp64 torch restatements, and the conformer engine
schedule end to end against goldens recorded from the live reference (tiny + conformer-large B=2)."""
import pytest
import torch
import torch.nn.functional as F

from tests import abi_emulator as E
from tests.engine_util import check_against_golden, run_engine_train_step
from tests.golden_util import load_case, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-3   # BASELINE.json north_star tolerance


def _ops():
    from wav2vec2forbrain_b200 import ops
    return ops


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, generator=g, device="cuda") * scale


def full(hi, lo):
    return hi.double().cpu() + (lo.double().cpu() if lo is not None else 0.0)


@pytest.mark.parametrize("n,hd,T", [(64, 16, 25), (1024, 64, 118), (96, 32, 7)])
def test_layernorm_rope(n, hd, T):
    ops = _ops()
    B = 3
    x = rnd(B * T, n, seed=1, scale=2.0) + 0.3
    gamma, beta = rnd(n, seed=2) * 0.2 + 1.0, rnd(n, seed=3) * 0.1
    inv = 1.0 / (10000 ** (torch.arange(0, hd, 2).float() / hd))
    cos, sin = ops.rope_table(inv.cuda(), T, hd)
    cos_r, sin_r = E.rope_table(inv, T, hd)
    assert rel_l2(cos.cpu(), cos_r) < 1e-6 and rel_l2(sin.cpu(), sin_r) < 1e-6
    y_hi, y_lo, r_hi, r_lo, mean, rstd = ops.layernorm_rope_fwd(x, gamma, beta, 1e-5, cos, sin, T, hd)
    ref = F.layer_norm(x.double().cpu(), (n,), gamma.double().cpu(), beta.double().cpu(), 1e-5)
    t = torch.arange(B * T) % T
    c = cos_r.double()[t].repeat(1, n // hd)
    s = sin_r.double()[t].repeat(1, n // hd)
    rot = ref * c + E._rot_half(ref, hd) * s
    assert rel_l2(full(y_hi, y_lo), ref) < 1e-5
    assert rel_l2(full(r_hi, r_lo), rot) < 1e-5
    # backward of the rotation: <R u, d> == <u, R^T d>
    d = rnd(B * T, 2 * n, seed=4)
    dplain = rnd(B * T, n, seed=5)
    out = ops.rope_bwd_add(d[:, :n], dplain, cos, sin, T, hd)
    ref_b = E.rope_bwd_add(d[:, :n].cpu(), dplain.cpu(), cos_r, sin_r, T, hd)
    assert rel_l2(out.cpu(), ref_b) < 1e-5
    u = rnd(B * T, n, seed=6).double().cpu()
    lhs = ((u * c + E._rot_half(u, hd) * s) * d[:, :n].double().cpu()).sum()
    rhs = (u * (ref_b.double() - dplain.double().cpu())).sum()
    assert abs(lhs - rhs) < 1e-4 * abs(lhs)


@pytest.mark.parametrize("B,T,D,K", [(2, 118, 1024, 31), (3, 25, 64, 7), (1, 300, 96, 31)])
@pytest.mark.parametrize("training", [True, False])
def test_conv_module_kernels(B, T, D, K, training):
    ops = _ops()
    M = B * T
    g = rnd(M, 2 * D, seed=1)
    w = rnd(D, 1, K, seed=2, scale=0.3)
    gamma, beta = rnd(D, seed=3) * 0.2 + 1.0, rnd(D, seed=4) * 0.1
    rm, rv = rnd(D, seed=5) * 0.1, torch.rand(D, device="cuda") + 0.5
    nbt = torch.zeros((), dtype=torch.int64, device="cuda")
    rm_c, rv_c, nbt_c = rm.cpu().clone(), rv.cpu().clone(), nbt.cpu().clone()
    # forward
    dwo, stats = ops.glu_dwconv_fwd(g, w, B, T, want_stats=training)
    dwo_r, stats_r = E.glu_dwconv_fwd(g.cpu(), w.cpu(), B, T, want_stats=training)
    assert rel_l2(dwo.cpu(), dwo_r) < 1e-5
    if training:
        assert rel_l2(stats.cpu(), stats_r) < 1e-5
    s_hi, s_lo, mean, rstd = ops.bn_silu_fwd(dwo, stats, gamma, beta, 1e-5, 0.1, rm, rv, nbt, training)
    h_r, l_r, mean_r, rstd_r = E.bn_silu_fwd(dwo_r, stats_r, gamma.cpu(), beta.cpu(), 1e-5, 0.1, rm_c, rv_c, nbt_c, training)
    assert rel_l2(full(s_hi, s_lo), full(h_r, l_r)) < 2e-5
    assert rel_l2(mean.cpu(), mean_r) < 1e-5 and rel_l2(rstd.cpu(), rstd_r) < 1e-5
    assert rel_l2(rm.cpu(), rm_c) < 1e-5 and rel_l2(rv.cpu(), rv_c) < 1e-5
    assert int(nbt) == int(nbt_c) == (1 if training else 0)
    # against torch's own BatchNorm1d + SiLU
    bn = torch.nn.BatchNorm1d(D).double()
    bn.weight.data, bn.bias.data = gamma.double().cpu(), beta.double().cpu()
    if not training:
        bn.running_mean.data, bn.running_var.data = rm_c.double(), rv_c.double()
        bn.eval()
    y_t = F.silu(bn(dwo_r.double().view(B, T, D).transpose(1, 2)).transpose(1, 2).reshape(M, D))
    assert rel_l2(full(s_hi, s_lo), y_t) < 2e-5
    # backward
    ds = rnd(M, D, seed=7)
    dgam, dbet = ops.bn_silu_bwd_stats(ds, dwo, mean, rstd, gamma, beta)
    dgam_r, dbet_r = E.bn_silu_bwd_stats(ds.cpu(), dwo_r, mean_r, rstd_r, gamma.cpu(), beta.cpu())
    assert rel_l2(dgam.cpu(), dgam_r) < 1e-4 and rel_l2(dbet.cpu(), dbet_r) < 1e-4
    dg_hi, dg_lo, dw = ops.glu_dwconv_bwd(ds, dwo, g, w, mean, rstd, gamma, beta, dgam, dbet, B, T, training)
    h_r, l_r, dw_r = E.glu_dwconv_bwd(ds.cpu(), dwo_r, g.cpu(), w.cpu(), mean_r, rstd_r, gamma.cpu(), beta.cpu(), dgam_r, dbet_r,
                                      B, T, training)
    assert rel_l2(full(dg_hi, dg_lo), full(h_r, l_r)) < 5e-5
    assert rel_l2(dw.cpu(), dw_r) < 1e-4


def test_gemm_out_scale_and_ln_dres():
    ops = _ops()
    m, n, k = 200, 96, 128
    a, b, res = rnd(m, k, seed=1), rnd(n, k, seed=2), rnd(m, n, seed=3)
    bias = rnd(n, seed=4)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    out = torch.empty(m, n, device="cuda")
    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, bias=bias, residual=res, out_scale=0.5)
    ref = 0.5 * (a.double() @ b.double().t() + bias.double()) + res.double()
    assert rel_l2(out.cpu(), ref.cpu()) < 2e-5
    x, dy, dres = rnd(m, n, seed=5), rnd(m, n, seed=6), rnd(m, n, seed=7)
    gamma, beta = rnd(n, seed=8) * 0.2 + 1.0, rnd(n, seed=9)
    _, _, _, mean, rstd = ops.layernorm_fwd(x, gamma, beta, 1e-5)
    dgm, dbt, dcol = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    dx, hi, lo = ops.layernorm_bwd(dy, x, gamma, mean, rstd, dgm, dbt, dcol, dres=dres)
    xr = x.double().cpu().requires_grad_(True)
    F.layer_norm(xr, (n,), gamma.double().cpu(), beta.double().cpu(), 1e-5).backward(dy.double().cpu())
    ref = xr.grad + dres.double().cpu()
    assert rel_l2(dx.cpu(), ref) < 1e-5
    assert rel_l2(full(hi, lo), ref) < 2e-5
    assert rel_l2(dcol.cpu(), ref.sum(0)) < 1e-4


def test_tiny_conformer_bf16x3():
    from wav2vec2forbrain_b200 import engine
    rec, _, _, sd, batch = load_case("tiny_conformer")
    logits, loss, lens, G, feats, _, _ = run_engine_train_step(engine, rec, sd, batch, "cuda")
    worst = check_against_golden(rec, logits, loss, lens, G, tol_logits=2e-4, tol_grad=TOL)
    print(f"tiny_conformer: worst per-parameter grad rel-L2 {worst:.2e}")


def test_conformer_large_b2_bf16x3():
    """wav2vec2-conformer-rope-large architecture (24 layers, 1024 wide), B=2, T=500, against the live reference."""
    from wav2vec2forbrain_b200 import engine, ops
    rec, _, _, sd, batch = load_case("conformer_large_b2")
    logits, loss, lens, G, feats, eng, P = run_engine_train_step(engine, rec, sd, batch, "cuda")
    err = rel_l2(logits.cpu(), rec["logits_train"])
    print(f"conformer_large_b2 logits rel-L2 {err:.2e}, loss {loss.item():.6f} vs {rec['loss_train'].item():.6f}")
    check_against_golden(rec, logits, loss, lens, G, tol_logits=TOL, tol_grad=TOL)
    p = "w2v_encoder.wav2vec2_conformer.encoder.layers.0.conv_module.batch_norm."
    assert rel_l2(P[p + "running_mean"].cpu(), rec["bn_running_mean_l0"]) < 1e-4
    assert rel_l2(P[p + "running_var"].cpu(), rec["bn_running_var_l0"]) < 1e-4
    # eval mode uses the BatchNorm running statistics; the golden's eval pass ran BEFORE its train pass
    # (oracle/gen_golden.py), so evaluate on a fresh copy of the buffers
    P = {k: v.cuda() for k, v in sd.items()}
    x = batch["input"].cuda()
    day = batch["day_idxs"].to("cuda", torch.int32)
    ev_logits, _, _, _ = eng.forward(P, x, day, training=False)
    am, ids, n = ops.greedy_decode(ev_logits)
    assert rel_l2(ev_logits.cpu(), rec["logits_eval"]) < TOL
    assert torch.equal(am.cpu().to(torch.int16), rec["argmax_eval"])
