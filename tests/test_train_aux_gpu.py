"""GPU parity of the kernels either side of the model step (csrc/train_aux.cu), through the
C ABI, against the pinned CPU oracle (oracle/train_aux_oracle.py) and torch.optim.Adam / clip_grad_norm_ themselves.
Integer results (edit-distance counts, collated ids) are bit-exact; floating point within the stated tolerance."""
import numpy as np
import pytest
import torch

from oracle import train_aux_oracle as A
from tests.test_train_aux_cpu import _random_ids, _rel_l2

pytestmark = pytest.mark.gpu


# ---------------------------------------------------------------------------------------------------------
# n1: fused Adam + clip
# ---------------------------------------------------------------------------------------------------------
def _param_sets(seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = [(37, 5), (301,), (64, 64), (1,), (70001,), (300, 1024), (3,)]
    return [torch.randn(*s, generator=g) for s in shapes], g


def test_fused_adam_matches_torch_adam():
    """3 steps, two LR groups, L2 weight decay, eps (b2t_gru_w2v_experiment.py:138-145); parameters whose gradient is
    None in some step are skipped in that step (their step counts drift, as with LayerDrop); tolerance 1e-6 rel-L2."""
    from wav2vec2forbrain_b200.optim import FusedAdam
    ps, g = _param_sets()
    ref = [torch.nn.Parameter(p.clone().cuda()) for p in ps]
    mine = [torch.nn.Parameter(p.clone().cuda()) for p in ps]
    mk = lambda P, cls, **kw: cls([{"params": P[:4]}, {"params": P[4:], "lr": 1e-5}], lr=1e-4, weight_decay=1e-4, eps=1e-8, **kw)
    o_ref, o_mine = mk(ref, torch.optim.Adam), mk(mine, FusedAdam)
    sched = [torch.optim.lr_scheduler.LambdaLR(o, [lambda e: 1.0, lambda e: 0.5 * (e + 1)]) for o in (o_ref, o_mine)]
    for step in range(4):
        grads = [torch.randn(p.shape, generator=g) * (10.0 ** (i % 3 - 1)) for i, p in enumerate(ps)]
        for i, (a, b) in enumerate(zip(ref, mine)):
            skip = (step == 1 and i in (2, 5)) or (step == 2 and i == 5)
            a.grad = None if skip else grads[i].cuda()
            b.grad = None if skip else grads[i].cuda()
        o_ref.step()
        o_mine.step()
        for s in sched:
            s.step()
        for i, (a, b) in enumerate(zip(ref, mine)):
            assert _rel_l2(b.detach().cpu().numpy(), a.detach().cpu().numpy()) < 1e-6, (step, i)
            if a in o_ref.state and o_ref.state[a]:
                assert _rel_l2(o_mine.state[b]["exp_avg"].cpu().numpy(), o_ref.state[a]["exp_avg"].cpu().numpy()) < 1e-6
                assert _rel_l2(o_mine.state[b]["exp_avg_sq"].cpu().numpy(), o_ref.state[a]["exp_avg_sq"].cpu().numpy()) < 1e-6
    assert o_mine.launches == 4                      # one kernel launch per step for all tensors and groups
    # the update itself against the oracle restatement as well (one tensor, one step from fresh state)
    p0 = ps[4].numpy()
    q = torch.nn.Parameter(ps[4].clone().cuda())
    q.grad = torch.ones_like(q) * 0.25
    FusedAdam([q], lr=1e-3, weight_decay=0.1).step()
    want, _, _ = A.adam_step(p0, np.full_like(p0, 0.25), np.zeros_like(p0), np.zeros_like(p0), 1, 1e-3, weight_decay=0.1)
    assert _rel_l2(q.detach().cpu().numpy(), want) < 1e-6


def test_fused_adam_many_step_classes():
    """more distinct (group, step) classes than one launch carries -> several launches, same result as torch"""
    from wav2vec2forbrain_b200.optim import MAX_GROUPS, FusedAdam
    n = MAX_GROUPS + 5
    g = torch.Generator().manual_seed(3)
    base = [torch.randn(50, generator=g) for _ in range(n)]
    ref = [torch.nn.Parameter(p.clone().cuda()) for p in base]
    mine = [torch.nn.Parameter(p.clone().cuda()) for p in base]
    o_ref, o_mine = torch.optim.Adam(ref, lr=1e-3), FusedAdam(mine, lr=1e-3)
    for step in range(n):
        for i in range(n):
            gr = torch.randn(50, generator=g).cuda() if i <= step else None     # parameter i joins at step i
            ref[i].grad = gr
            mine[i].grad = None if gr is None else gr.clone()
        o_ref.step()
        o_mine.step()
    for a, b in zip(ref, mine):
        assert _rel_l2(b.detach().cpu().numpy(), a.detach().cpu().numpy()) < 1e-6


@pytest.mark.parametrize("max_norm", [0.5, 1e9])
def test_fused_clip_grad_norm(max_norm):
    from wav2vec2forbrain_b200.optim import clip_grad_norm_
    ps, g = _param_sets(1)
    ref = [torch.nn.Parameter(p.clone().cuda()) for p in ps]
    mine = [torch.nn.Parameter(p.clone().cuda()) for p in ps]
    for i, (a, b) in enumerate(zip(ref, mine)):
        gr = None if i == 3 else torch.randn(a.shape, generator=g).cuda()
        a.grad, b.grad = gr, (None if gr is None else gr.clone())
    t_ref = torch.nn.utils.clip_grad_norm_(ref, max_norm)
    t_mine = clip_grad_norm_(mine, max_norm)
    assert t_mine.is_cuda and abs(t_mine.item() - t_ref.item()) < 1e-6 * t_ref.item()
    for a, b in zip(ref, mine):
        if a.grad is not None:
            assert _rel_l2(b.grad.cpu().numpy(), a.grad.cpu().numpy()) < 1e-6
    want, tot = A.clip_grad_norm([p.grad.cpu().numpy() for p in ref if p.grad is not None], 1e30)
    assert abs(tot - torch.nn.utils.clip_grad_norm_(ref, 1e30).item()) < 1e-5 * tot


# ---------------------------------------------------------------------------------------------------------
# n3: WER / CER counts
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("B,T,S,special_p", [(32, 118, 40, 0.05), (5, 7, 3, 0.3), (64, 118, 80, 0.0), (3, 300, 120, 0.1)])
def test_ctc_error_counts_bit_exact(B, T, S, special_p):
    from wav2vec2forbrain_b200.evaluator import DEFAULT_VOCAB, ctc_error_counts, vocab_table
    rng = np.random.default_rng(B * 1000 + T)
    pred = _random_ids(rng, B, T, special_p)
    tgt = np.zeros((B, S), dtype=np.int64)
    for b in range(B):
        n = int(rng.integers(0, S + 1)) if b else S          # ragged labels, one empty-capable, one full
        body = rng.choice(np.asarray([4, 4, 27] + list(range(5, 32))), n)
        tgt[b, :n] = body
    tgt[B - 1] = 0                                           # an all-padding label (0 target words / characters)
    want, ps, ls = A.error_counts(pred, tgt, cut_eos=True)
    table = vocab_table(DEFAULT_VOCAB).cuda()
    counts = torch.zeros(4, dtype=torch.int64, device="cuda")
    per = torch.zeros(B, 4, dtype=torch.int32, device="cuda")
    ctc_error_counts(torch.from_numpy(pred).to("cuda", torch.int32), torch.from_numpy(tgt).cuda(), table, counts, per_trial=per)
    assert per.cpu().long().tolist() == want.tolist()
    assert counts.cpu().tolist() == want.sum(0).tolist()
    # without the eos cut (DefaultEvaluator) and without clean-up
    want2, _, _ = A.error_counts(pred, tgt, cut_eos=False, cleanup=False)
    per.zero_()
    ctc_error_counts(torch.from_numpy(pred).to("cuda", torch.int32), torch.from_numpy(tgt).cuda(), table, counts, eos_id=-1,
                     cleanup=False, per_trial=per)
    assert per.cpu().long().tolist() == want2.tolist()


def test_gpu_evaluator_tracks_batches_without_host_sync_until_evaluate():
    from wav2vec2forbrain_b200.evaluator import GpuEvaluator
    from wav2vec2forbrain_b200.models import B2tSampleBatch, ModelOutput
    rng = np.random.default_rng(9)
    ev = GpuEvaluator("train")
    want_batches = []
    for i in range(3):
        B, T, S = 8, 50, 20
        logits = torch.randn(B, T, 32, generator=torch.Generator().manual_seed(i)) * 3
        tgt = torch.from_numpy(rng.integers(4, 32, (B, S)))
        tgt[:, 15:] = 0
        out = ModelOutput(logits.cuda(), {"ctc_loss": torch.tensor(1.5 + i).cuda()}, loss=torch.tensor(1.5 + i).cuda())
        ev.track_batch(out, B2tSampleBatch(None, tgt.cuda()))
        c, _, _ = A.error_counts(logits.argmax(-1).numpy(), tgt.numpy())
        want_batches.append(c.sum(0))
    res = ev.evaluate()
    for b, w in zip(res["batches"], want_batches):
        assert abs(b["metrics"]["word_error_rate"] - w[2] / w[3]) < 1e-12
        assert abs(b["metrics"]["char_error_rate"] - w[0] / w[1]) < 1e-12
    assert [round(b["loss"], 6) for b in res["batches"]] == [1.5, 2.5, 3.5]
    assert abs(res["average"]["ctc_loss"] - 2.5) < 1e-6
    logits = torch.randn(2, 30, 32, generator=torch.Generator().manual_seed(5)) * 3
    preds, _ = ev.decode_predictions(ModelOutput(logits.cuda(), {}))
    assert preds == [A.cut_after_eos(A.decode_string(r.tolist(), True)) for r in logits.argmax(-1)]


# ---------------------------------------------------------------------------------------------------------
# n4: z-scoring, collation, loader
# ---------------------------------------------------------------------------------------------------------
def _trials(n=23, C=256, seed=2):
    rng = np.random.default_rng(seed)
    feats = [(rng.normal(1.0, 2.0, (int(rng.integers(40, 200)), C)) * rng.uniform(0.5, 3.0, C)).astype(np.float32) for _ in range(n)]
    texts = ["".join(rng.choice(list("ABCDEFGHI ' ,"), int(rng.integers(0, 30)))) for _ in range(n)]
    days = rng.integers(0, 24, n).tolist()
    blocks = [np.arange(0, 7), np.arange(7, 15), np.arange(15, n)]
    return feats, texts, days, blocks


def test_trial_store_zscore_and_collate_match_oracle():
    from wav2vec2forbrain_b200.pipeline import TrialStore
    feats, texts, days, blocks = _trials()
    store = TrialStore.from_trials(feats, texts, days, "cuda", blocks=blocks)
    z = A.block_zscore([f.astype(np.float64) for f in feats], blocks)
    got = store.frames.cpu().numpy()
    off = store.offsets.cpu().numpy()
    for i, zi in enumerate(z):
        # tolerance: the reference computes in float64 and stores float32 -> 1e-6 relative to the z-scored scale
        assert np.abs(got[off[i]:off[i + 1]] - zi.astype(np.float32)).max() < 2e-6 * max(1.0, np.abs(zi).max()), i
    labels = [A.tokenize_label(t) for t in texts]
    idx = [3, 20, 11, 0]
    want = A.collate([f.astype(np.float32) for f in z], labels, days, idx)
    b = store.collate(torch.tensor(idx, dtype=torch.int32, device="cuda"), want["input"].shape[1], max(1, want["target"].shape[1]))
    assert np.abs(b.input.cpu().numpy() - want["input"]).max() < 2e-6 * np.abs(want["input"]).max()
    assert b.target.cpu().numpy()[:, :want["target"].shape[1]].tolist() == want["target"].tolist()
    assert b.input_lens.cpu().tolist() == want["input_lens"].tolist()
    assert b.target_lens.cpu().tolist() == want["target_lens"].tolist()
    assert b.day_idxs.cpu().tolist() == want["day_idxs"].tolist()
    # padding frames are exact zeros
    for r, i in enumerate(idx):
        assert torch.count_nonzero(b.input[r, feats[i].shape[0]:]) == 0


def test_device_loader_epoch_and_sharding():
    from wav2vec2forbrain_b200.pipeline import DeviceLoader, TrialStore
    feats, texts, days, _ = _trials(n=50, C=8)
    store = TrialStore.from_trials(feats, texts, days, "cuda")
    loader = DeviceLoader(store, batch_size=8, seed=1, bucket_width=32)
    seen = 0
    for batch in loader:
        assert batch.input.is_cuda and batch.input.shape[0] == batch.day_idxs.shape[0] <= 8
        assert int(batch.input_lens.max()) == batch.input.shape[1]
        seen += batch.input.shape[0]
    assert seen == 50
    plain = DeviceLoader(store, batch_size=8, seed=1, bucket_width=None)
    assert loader.padded_fraction() < plain.padded_fraction()
    shards = [DeviceLoader(store, 8, seed=1, rank=r, world=2) for r in range(2)]
    b0, b1 = shards[0].epoch_batches(), shards[1].epoch_batches()
    assert len(b0) == len(b1) and not set(np.concatenate(b0).tolist()) & set(np.concatenate(b1).tolist())


def test_model_trains_from_the_device_loader():
    """end to end: TrialStore -> DeviceLoader -> product model -> FusedAdam + fused clip -> GpuEvaluator"""
    from tests.test_models_gpu import build_from_golden
    from wav2vec2forbrain_b200.evaluator import GpuEvaluator
    from wav2vec2forbrain_b200.optim import FusedAdam, clip_grad_norm_
    from wav2vec2forbrain_b200.pipeline import DeviceLoader, TrialStore
    m, rec, sd, batch, models = build_from_golden("tiny_base")
    m.load_state_dict(sd, strict=True)
    rng = np.random.default_rng(4)
    feats = [rng.normal(0.5, 2.0, (int(rng.integers(160, 200)), 256)).astype(np.float32) for _ in range(12)]
    texts = ["".join(rng.choice(list("ABCDE "), int(rng.integers(3, 9)))).strip() or "AB" for _ in range(12)]
    days = rng.integers(0, 24, 12).tolist()
    store = TrialStore.from_trials(feats, texts, days, "cuda", blocks=[np.arange(0, 12)])
    loader = DeviceLoader(store, batch_size=4, seed=3, bucket_width=64)
    opt = FusedAdam([{"params": m.brain_encoder.parameters()}, {"params": m.w2v_encoder.parameters(), "lr": 1e-4}], lr=1e-3,
                    weight_decay=1e-4)
    ev = GpuEvaluator("train")
    m.train()
    losses = []
    for epoch in range(3):
        for b in loader:
            opt.zero_grad()
            out = m.forward(b)
            out.loss.backward()
            clip_grad_norm_(m.parameters(), 10.0)
            opt.step()
            ev.track_batch(out, b)
    res = ev.evaluate()
    losses = [b["loss"] for b in res["batches"]]
    assert len(losses) == 9 and all(np.isfinite(losses))
    assert np.mean(losses[-3:]) < np.mean(losses[:3])          # it learns
    assert 0.0 <= res["average"]["char_error_rate"]
