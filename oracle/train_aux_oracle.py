"""CPU oracle for the rows either side of the model step.

TEST INFRASTRUCTURE ONLY (same rule as oracle/b2t_oracle.py): imported by tests/ only, never by the product.

Restates, in numpy / plain Python, what the reference executes through third-party code:
  * torch.optim.Adam's single-tensor update and torch.nn.utils.clip_grad_norm_
    (src/experiments/experiment.py:25-28, src/experiments/b2t_gru_w2v_experiment.py:138-145,
    src/train/train_loop.py:72-78; pytorch==2.1.1 torch/optim/adam.py `_single_tensor_adam`)
  * the evaluator's string metrics: Wav2Vec2CTCTokenizer.batch_decode rules
    (transformers==4.35.2 tokenization_wav2vec2.py convert_tokens_to_string / _decode / clean_up_tokenization),
    torcheval WordErrorRate and edit_distance.SequenceMatcher.distance (src/train/evaluator.py:108-120,163-229)
  * block-wise z-scoring and the collate function (src/datasets/preprocessing.py:60-72,100-113,
    src/datasets/brain2text.py:178-211, src/util/nn_helper.py:52-57)
Pinning (tests/test_train_aux_cpu.py): Adam / clip against torch.optim.Adam / torch.nn.utils.clip_grad_norm_ run in
the test; the decode rules against the installed transformers Wav2Vec2CTCTokenizer built from a local vocabulary
(the reference's own tokenizer class); Levenshtein against known answers; torcheval / edit_distance are not installed
in this image, their published algorithm (unit-cost Levenshtein over words / characters) is restated.
"""
from __future__ import annotations

import math
import re
from typing import List, Optional, Sequence, Tuple

import numpy as np

# facebook/wav2vec2-base-960h vocabulary
VOCAB = ["<pad>", "<s>", "</s>", "<unk>", "|", "E", "T", "A", "O", "N", "I", "H", "S", "R", "D", "L", "U", "M", "W", "C", "F",
         "G", "Y", "P", "B", "V", "K", "'", "X", "J", "Q", "Z"]
PAD_ID, BOS_ID, EOS_ID, UNK_ID, DELIM_ID = 0, 1, 2, 3, 4


# ---------------------------------------------------------------------------------------------------------
# n1: Adam + gradient clipping
# ---------------------------------------------------------------------------------------------------------
def adam_step(p, g, m, v, step: int, lr: float, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
    """One torch.optim.Adam update (amsgrad=False, maximize=False) in fp32, operation order of
    `_single_tensor_adam`.  Returns (p, m, v)."""
    f = np.float32
    b1, b2 = betas
    g = g.astype(f) + f(weight_decay) * p.astype(f)
    m = m + (g - m) * f(1.0 - b1)
    v = v * f(b2) + f(1.0 - b2) * g * g
    bc1 = 1.0 - b1 ** step
    bc2 = 1.0 - b2 ** step
    step_size = lr / bc1
    denom = np.sqrt(v) / f(math.sqrt(bc2)) + f(eps)
    p = p - f(step_size) * (m / denom)
    return p.astype(f), m.astype(f), v.astype(f)


def clip_grad_norm(grads: Sequence[np.ndarray], max_norm: float) -> Tuple[List[np.ndarray], float]:
    """torch.nn.utils.clip_grad_norm_ with norm_type=2: -> (scaled grads, total norm)."""
    total = math.sqrt(sum(float((g.astype(np.float64) ** 2).sum()) for g in grads))
    coef = min(1.0, max_norm / (total + 1e-6))
    return [(g * np.float32(coef)).astype(np.float32) for g in grads], total


# ---------------------------------------------------------------------------------------------------------
# n3: greedy-decode strings, WER / CER
# ---------------------------------------------------------------------------------------------------------
def decode_string(ids: Sequence[int], group_tokens: bool, vocab: Sequence[str] = VOCAB, cleanup: bool = True) -> str:
    """Wav2Vec2CTCTokenizer._decode(ids, skip_special_tokens=False, group_tokens=...): group repeats, drop <pad>,
    '|' -> ' ', join, strip, clean_up_tokenization."""
    toks = [vocab[i] for i in ids]
    if group_tokens:
        toks = [t for k, t in enumerate(toks) if k == 0 or t != toks[k - 1]]
    toks = [t for t in toks if t != vocab[PAD_ID]]
    s = "".join(" " if t == vocab[DELIM_ID] else t for t in toks).strip()
    if cleanup:
        for a, b in ((" .", "."), (" ?", "?"), (" !", "!"), (" ,", ","), (" ' ", "'"), (" n't", "n't"), (" 'm", "'m"),
                     (" 's", "'s"), (" 've", "'ve"), (" 're", "'re")):
            s = s.replace(a, b)
    return s


def cut_after_eos(s: str, eos: str = "</s>") -> str:
    """evaluator.py:169-176"""
    i = s.find(eos)
    return s[:i + len(eos)] if i != -1 else s


def levenshtein(a: Sequence, b: Sequence) -> int:
    """unit-cost insert / delete / substitute edit distance"""
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


def error_counts(pred_ids: np.ndarray, targets: np.ndarray, cut_eos: bool = True, cleanup: bool = True):
    """-> (per-trial [B,4] = char distance, target chars, word distance, target words; predicted strings; label strings)
    following evaluator.py:108-120 (decode) and :163-229 (metrics)."""
    out, ps, ls = [], [], []
    for b in range(pred_ids.shape[0]):
        p = decode_string(pred_ids[b].tolist(), True, cleanup=cleanup)
        if cut_eos:
            p = cut_after_eos(p)
        t = decode_string(targets[b].tolist(), False, cleanup=cleanup)
        out.append((levenshtein(t, p), len(t), levenshtein(p.split(), t.split()), len(t.split())))
        ps.append(p)
        ls.append(t)
    return np.asarray(out, dtype=np.int64), ps, ls


# ---------------------------------------------------------------------------------------------------------
# n4: input pipeline
# ---------------------------------------------------------------------------------------------------------
def block_zscore(features: List[np.ndarray], blocks: List[np.ndarray]) -> List[Optional[np.ndarray]]:
    """preprocessing.py:60-72 / :100-113: per block, mean / population std over all frames of the block's trials,
    (x - mean) / (std + 1e-8).  Returns the normalised trial arrays indexed like `features` (None = in no block)."""
    out: List[Optional[np.ndarray]] = [None] * len(features)
    for idx in blocks:
        cat = np.concatenate([features[i] for i in range(idx[0], idx[-1] + 1)], axis=0)
        mean = np.mean(cat, axis=0, keepdims=True)
        std = np.std(cat, axis=0, keepdims=True)
        for i in idx:
            out[i] = (features[i] - mean) / (std + 1e-8)
    return out


def tokenize_label(text: str, remove_punctuation: bool = True, vocab: Sequence[str] = VOCAB) -> List[int]:
    """brain2text.py:137 (.upper()), :192-197 (punctuation regex), then Wav2Vec2CTCTokenizer character tokenisation:
    ' ' -> '|', characters outside the vocabulary -> <unk>."""
    text = text.upper()
    if remove_punctuation:
        text = re.sub(r'[\,\?\.\!\-\;\:"]', "", text)
    table = {t: i for i, t in enumerate(vocab)}
    return [table.get("|" if c == " " else c, UNK_ID) for c in text]


def calc_seq_len(ids: Sequence[int]) -> int:
    """nn_helper.py:52-57: index after the last id > 0."""
    for j in range(len(ids) - 1, -1, -1):
        if ids[j] > 0:
            return j + 1
    return 0


def collate(features: List[np.ndarray], labels: List[List[int]], days: List[int], idx: Sequence[int]):
    """brain2text.py:178-211 for the trials `idx`: zero-padded inputs [B, Tmax, C], labels padded with 0 to the longest,
    day_idxs, input_lens, target_lens."""
    tmax = max(features[i].shape[0] for i in idx)
    smax = max(len(labels[i]) for i in idx)
    C = features[idx[0]].shape[1]
    x = np.zeros((len(idx), tmax, C), dtype=np.float32)
    tg = np.zeros((len(idx), smax), dtype=np.int64)
    for r, i in enumerate(idx):
        x[r, :features[i].shape[0]] = features[i]
        tg[r, :len(labels[i])] = labels[i]
    return {"input": x, "target": tg, "day_idxs": np.asarray([days[i] for i in idx], dtype=np.int64),
            "input_lens": np.asarray([features[i].shape[0] for i in idx], dtype=np.int64),
            "target_lens": np.asarray([calc_seq_len(row) for row in tg], dtype=np.int64)}
