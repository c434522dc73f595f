"""CPU suite for the rows either side of the model step: the oracle
(oracle/train_aux_oracle.py) pinned against the third-party code the reference calls (torch.optim.Adam,
clip_grad_norm_, the installed transformers Wav2Vec2CTCTokenizer), the host-side logic of the product (bucketing,
label tokenisation, table layout), and the struct layouts of the new C-ABI entries.  No GPU compute."""
import ctypes
import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest
import torch

from oracle import train_aux_oracle as A

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(REPO, "include", "b2t_sm100.h")


def _rel_l2(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


# ---------------------------------------------------------------------------------------------------------
# n1
# ---------------------------------------------------------------------------------------------------------
def test_oracle_adam_matches_torch_adam_two_groups():
    """3 steps, 2 parameter groups with different learning rates, weight_decay and eps as create_optimizer passes them
    (b2t_gru_w2v_experiment.py:138-145); a parameter without gradient is left alone."""
    g = torch.Generator().manual_seed(0)
    ps = [torch.randn(37, 5, generator=g), torch.randn(301, generator=g), torch.randn(64, 64, generator=g)]
    unused = torch.nn.Parameter(torch.randn(9, generator=g))
    tp = [torch.nn.Parameter(p.clone()) for p in ps]
    opt = torch.optim.Adam([{"params": tp[:2] + [unused]}, {"params": tp[2:], "lr": 1e-5}], lr=1e-4, weight_decay=1e-4, eps=1e-8)
    mine = [(p.numpy().copy(), np.zeros(p.shape, np.float32), np.zeros(p.shape, np.float32)) for p in ps]
    lrs = [1e-4, 1e-4, 1e-5]
    before = unused.detach().clone()
    for step in range(1, 4):
        grads = [torch.randn(p.shape, generator=g) * (10.0 ** (i - 1)) for i, p in enumerate(ps)]
        for p, gr in zip(tp, grads):
            p.grad = gr.clone()
        opt.step()
        mine = [A.adam_step(p, gr.numpy(), m, v, step, lr, eps=1e-8, weight_decay=1e-4) for (p, m, v), gr, lr in zip(mine, grads, lrs)]
        for (p, m, v), t in zip(mine, tp):
            for mine_x, ref_x in ((p, t.detach()), (m, opt.state[t]["exp_avg"]), (v, opt.state[t]["exp_avg_sq"])):
                assert _rel_l2(mine_x, ref_x.numpy()) < 1e-6       # fp32 rounding of single elements aside
    assert torch.equal(unused.detach(), before)


@pytest.mark.parametrize("max_norm", [0.5, 1e6])
def test_oracle_clip_matches_torch(max_norm):
    g = torch.Generator().manual_seed(1)
    ps = [torch.nn.Parameter(torch.randn(n, generator=g)) for n in (10, 3000, 77)]
    for p in ps:
        p.grad = torch.randn(p.shape, generator=g)
    mine, total = A.clip_grad_norm([p.grad.numpy().copy() for p in ps], max_norm)
    ref_total = torch.nn.utils.clip_grad_norm_(ps, max_norm)
    assert abs(total - ref_total.item()) < 1e-5 * ref_total.item()
    for m, p in zip(mine, ps):
        assert _rel_l2(m, p.grad.numpy()) < 1e-6


def _c_layout(struct, fields):
    prog = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {',
            f'  printf("sizeof %zu\\n", sizeof({struct}));']
    prog += [f'  printf("{f} %zu\\n", offsetof({struct}, {f}));' for f in fields]
    prog += ['  return 0;', '}']
    with tempfile.TemporaryDirectory() as d:
        c, exe = os.path.join(d, "l.c"), os.path.join(d, "l")
        with open(c, "w") as fh:
            fh.write("\n".join(prog))
        subprocess.run(["gcc", "-std=c99", "-o", exe, c], check=True)
        out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()
    return dict(zip(out[0::2], map(int, out[1::2])))


def test_multi_tensor_struct_layouts_match_the_c_compiler():
    from wav2vec2forbrain_b200 import optim
    got = _c_layout("b2t_adam_group", [f[0] for f in optim.AdamGroup._fields_])
    assert got["sizeof"] == ctypes.sizeof(optim.AdamGroup)
    for f, _ in optim.AdamGroup._fields_:
        assert got[f] == getattr(optim.AdamGroup, f).offset
    got = _c_layout("b2t_mt_tensor", ["p", "g", "m", "v", "n", "group"])
    assert got["sizeof"] == optim._TENSOR_DT.itemsize
    for f in ("p", "g", "m", "v", "n", "group"):
        assert got[f] == optim._TENSOR_DT.fields[f][1]
    src = open(HEADER).read()
    assert f"#define B2T_MT_CHUNK {optim.MT_CHUNK}" in src and f"#define B2T_ADAM_MAX_GROUPS {optim.MAX_GROUPS}" in src


# ---------------------------------------------------------------------------------------------------------
# n3
# ---------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def hf_tokenizer():
    """The reference's tokenizer class (b2t_experiment.py:27 AutoTokenizer -> Wav2Vec2CTCTokenizer) from a local vocab."""
    from transformers import Wav2Vec2CTCTokenizer
    d = tempfile.mkdtemp()
    with open(os.path.join(d, "vocab.json"), "w") as f:
        json.dump({t: i for i, t in enumerate(A.VOCAB)}, f)
    return Wav2Vec2CTCTokenizer(os.path.join(d, "vocab.json"), unk_token="<unk>", pad_token="<pad>", bos_token="<s>",
                                eos_token="</s>", word_delimiter_token="|", clean_up_tokenization_spaces=True)


def _random_ids(rng, n, T, special_p=0.05):
    """argmax-like id rows: runs of repeats, many blanks, delimiters, apostrophes, a few special tokens"""
    rows = []
    for _ in range(n):
        row = []
        while len(row) < T:
            r = rng.random()
            if r < 0.35:
                tok = 0
            elif r < 0.50:
                tok = 4
            elif r < 0.58:
                tok = 27
            elif r < 0.58 + special_p:
                tok = int(rng.integers(1, 4))
            else:
                tok = int(rng.integers(5, 32))
            row += [tok] * int(rng.integers(1, 4))
        rows.append(row[:T])
    return np.asarray(rows, dtype=np.int64)


def test_oracle_decode_matches_the_transformers_tokenizer(hf_tokenizer):
    rng = np.random.default_rng(3)
    ids = _random_ids(rng, 200, 60)
    ref_group = hf_tokenizer.batch_decode(ids, group_tokens=True)
    ref_plain = hf_tokenizer.batch_decode(ids, group_tokens=False)
    for b in range(ids.shape[0]):
        assert A.decode_string(ids[b].tolist(), True) == ref_group[b]
        assert A.decode_string(ids[b].tolist(), False) == ref_plain[b]
    # the apostrophe clean-up rule is hit by the data above, and only that rule can be
    assert any(" ' " in A.decode_string(r.tolist(), True, cleanup=False) for r in ids)


def test_label_tokenisation_matches_the_transformers_tokenizer(hf_tokenizer):
    from wav2vec2forbrain_b200.pipeline import tokenize_label
    texts = ["Hello world", "don't stop -- now!", "it's  a: \"test\"?", "naïve café", ""]
    for t in texts:
        import re
        clean = re.sub(r'[\,\?\.\!\-\;\:"]', "", t.upper())
        ref = hf_tokenizer(clean).input_ids
        assert A.tokenize_label(t) == ref
        assert tokenize_label(t) == ref


def test_levenshtein_known_answers_and_properties():
    assert A.levenshtein("kitten", "sitting") == 3
    assert A.levenshtein("", "abc") == 3 and A.levenshtein("abc", "") == 3 and A.levenshtein("", "") == 0
    assert A.levenshtein("flaw", "lawn") == 2
    assert A.levenshtein("the cat sat".split(), "the cat sat on".split()) == 1
    rng = np.random.default_rng(0)
    for _ in range(50):
        a = rng.integers(0, 4, rng.integers(0, 12)).tolist()
        b = rng.integers(0, 4, rng.integers(0, 12)).tolist()
        c = rng.integers(0, 4, rng.integers(0, 12)).tolist()
        dab = A.levenshtein(a, b)
        assert dab == A.levenshtein(b, a)
        assert abs(len(a) - len(b)) <= dab <= max(len(a), len(b))
        assert A.levenshtein(a, c) <= dab + A.levenshtein(b, c)


def test_error_counts_definition():
    """WER / CER as evaluator.py:163-229 defines them, on a hand-checked pair."""
    v = {t: i for i, t in enumerate(A.VOCAB)}
    enc = lambda s: [v["|" if c == " " else c] for c in s]
    pred = np.asarray([[0] + enc("THE CAT") + [0, 0] + [v["</s>"]] + enc("X")], dtype=np.int64)
    tgt = np.asarray([enc("THE CART SAT") + [0]], dtype=np.int64)
    counts, ps, ls = A.error_counts(pred, tgt)
    assert ps == ["THE CAT</s>"] and ls == ["THE CART SAT"]
    assert counts[0].tolist() == [A.levenshtein("THE CART SAT", "THE CAT</s>"), 12, 2, 3]


# ---------------------------------------------------------------------------------------------------------
# n4
# ---------------------------------------------------------------------------------------------------------
def test_oracle_block_zscore_and_collate():
    rng = np.random.default_rng(5)
    feats = [rng.normal(2.0, 3.0, (int(rng.integers(20, 60)), 8)) for _ in range(7)]
    blocks = [np.asarray([0, 1, 2]), np.asarray([3, 4, 5, 6])]
    z = A.block_zscore(feats, blocks)
    cat = np.concatenate(z[:3])
    assert np.allclose(cat.mean(0), 0, atol=1e-12) and np.allclose(cat.std(0), 1, atol=1e-6)
    labels = [A.tokenize_label(s) for s in ("ab c", "hello", "x", "y z", "", "q", "it's")]
    c = A.collate([f.astype(np.float32) for f in z], labels, list(range(7)), [1, 4, 6])
    assert c["input"].shape == (3, max(feats[i].shape[0] for i in (1, 4, 6)), 8)
    assert c["target"].shape[1] == 5 and c["target_lens"].tolist() == [5, 0, 4]
    assert np.all(c["input"][1, feats[4].shape[0]:] == 0)


def test_bucketed_batches_cover_every_trial_once_and_cut_padding():
    from wav2vec2forbrain_b200.pipeline import bucketed_batches
    L = np.random.default_rng(0).integers(100, 600, 1000)

    def padding(batches):
        tot = sum(int(L[b].max()) * len(b) for b in batches)
        return (tot - sum(int(L[b].sum()) for b in batches)) / tot
    plain = bucketed_batches(L, 32, 42, 0, None)
    buck = bucketed_batches(L, 32, 42, 0, 32)
    for bs in (plain, buck):
        assert sorted(np.concatenate(bs).tolist()) == list(range(1000))
        assert all(len(b) <= 32 for b in bs)
    assert padding(buck) < 0.25 * padding(plain)
    # same seed/epoch -> same order on every rank; another epoch -> another order
    again = bucketed_batches(L, 32, 42, 0, 32)
    assert all(np.array_equal(a, b) for a, b in zip(buck, again))
    other = bucketed_batches(L, 32, 42, 1, 32)
    assert not all(np.array_equal(a, b) for a, b in zip(buck, other))


def test_vocab_table_layout():
    from wav2vec2forbrain_b200.evaluator import DEFAULT_VOCAB, vocab_table
    assert DEFAULT_VOCAB == A.VOCAB
    t = vocab_table(DEFAULT_VOCAB)
    assert t.shape == (32, 8) and t[2, 0] == 4 and bytes(t[2, 1:5].tolist()) == b"</s>" and t[5, 0] == 1
