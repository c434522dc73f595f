"""GPU evaluator path: greedy CTC decode + WER / CER on the device.

The reference evaluates EVERY batch — training batches included (src/train/train_loop.py:79) — on the host:
`logits.argmax(-1).cpu().numpy()` -> `tokenizer.batch_decode` -> torcheval WordErrorRate + edit_distance character
distance in Python (src/train/evaluator.py:108-120,163-229), plus three `.item()` synchronisations.  At B200 step times
that host work is the end-to-end bottleneck.  Here `track_batch` enqueues two kernels
(b2t_greedy_decode, b2t_ctc_error_counts) and keeps the four error counts and the loss on the device; the host reads
them once, in `evaluate()` at the end of the epoch (or when somebody asks for the running loss).

Metric definitions are the reference's: per batch `word_error_rate` = sum of word edit distances / sum of target
words, `char_error_rate` = sum of character edit distances / sum of target characters, over strings decoded by the
tokenizer's rules (group repeats, drop <pad>, '|' -> ' ', strip, clean-up; predictions cut after "</s>" as
EvaluatorWithW2vLMDecoder does).  Strings themselves (`decode_predictions`) are only materialised on request.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib, ops
from ._lib import c_int, c_int64, check, lib, ptr, stream

# facebook/wav2vec2-base-960h vocabulary; `|` is the word delimiter, <pad> the CTC blank
DEFAULT_VOCAB = ["<pad>", "<s>", "</s>", "<unk>", "|", "E", "T", "A", "O", "N", "I", "H", "S", "R", "D", "L", "U", "M", "W",
                 "C", "F", "G", "Y", "P", "B", "V", "K", "'", "X", "J", "Q", "Z"]


def vocab_table(vocab: Sequence[str]) -> torch.Tensor:
    """[V, 8] uint8: byte 0 = length, bytes 1.. = the token's characters (b2t_ctc_error_counts)."""
    t = torch.zeros(len(vocab), 8, dtype=torch.uint8)
    for i, tok in enumerate(vocab):
        b = tok.encode("ascii")
        if len(b) > 7:
            raise ValueError(f"token {tok!r} is longer than 7 characters")
        t[i, 0] = len(b)
        t[i, 1:1 + len(b)] = torch.tensor(list(b), dtype=torch.uint8)
    return t


def ctc_error_counts(argmax_ids: torch.Tensor, targets: torch.Tensor, table: torch.Tensor, counts: torch.Tensor,
                     pad_id: int = 0, delim_id: int = 4, eos_id: int = 2, cleanup: bool = True,
                     per_trial: Optional[torch.Tensor] = None) -> None:
    """counts [4] int64 += (char distance, target chars, word distance, target words) of the batch."""
    _lib.ensure_device()
    B, T = argmax_ids.shape
    assert argmax_ids.dtype == torch.int32 and argmax_ids.stride(1) == 1
    assert targets.dtype == torch.int64 and targets.is_contiguous() and targets.shape[0] == B
    assert counts.dtype == torch.int64 and counts.numel() >= 4 and table.dtype == torch.uint8
    check(lib().b2t_ctc_error_counts(ptr(argmax_ids), c_int(T), c_int64(argmax_ids.stride(0)), ptr(targets),
                                     c_int(targets.shape[1]), ptr(table), c_int(table.shape[0]), c_int(pad_id),
                                     c_int(delim_id), c_int(eos_id), c_int(1 if cleanup else 0), ptr(counts), ptr(per_trial),
                                     c_int(B), stream()), "b2t_ctc_error_counts")
    ops._count()


class GpuEvaluator:
    """`track_batch(predictions, sample)` / `evaluate()` with the reference Evaluator's meaning
    (src/train/evaluator.py:24-47), metrics computed and accumulated on the device."""

    def __init__(self, mode: str = "train", vocab: Sequence[str] = DEFAULT_VOCAB, pad_id: int = 0, delim_id: int = 4,
                 eos_id: Optional[int] = 2, cleanup: bool = True):
        self.mode = mode
        self.vocab = list(vocab)
        self.pad_id, self.delim_id = pad_id, delim_id
        self.eos_id = -1 if eos_id is None else eos_id
        self.cleanup = cleanup
        self._table: Optional[torch.Tensor] = None
        self._counts: List[torch.Tensor] = []      # per batch [4] int64 (device)
        self._losses: List[torch.Tensor] = []      # per batch 0-dim (device)
        self._extra: List[dict] = []
        self.running_loss = 0.0
        self._n_host = 0

    # ---- per batch: two kernel launches, no synchronisation -------------------------------------------------
    def track_batch(self, predictions, sample) -> None:
        logits = predictions.logits
        if self._table is None or self._table.device != logits.device:
            self._table = vocab_table(self.vocab).to(logits.device)
        am, _, _ = ops.greedy_decode(logits.detach())
        counts = torch.zeros(4, dtype=torch.int64, device=logits.device)
        if sample.target is not None:
            ctc_error_counts(am, sample.target.to(logits.device).contiguous(), self._table, counts, self.pad_id,
                             self.delim_id, self.eos_id, self.cleanup)
        assert predictions.loss is not None, "Loss is None. Make sure to set loss in ModelOutput"
        self._counts.append(counts)
        self._losses.append(predictions.loss.detach().reshape(()))
        # metrics the model put there itself (ctc_loss) stay lazy device values until evaluate()
        self._extra.append(list(dict.items(predictions.metrics)) if isinstance(predictions.metrics, dict) else [])

    # ---- host reads -----------------------------------------------------------------------------------------
    def _materialise(self):
        if not self._counts:
            return torch.zeros(0, 4, dtype=torch.int64), torch.zeros(0)
        counts = torch.stack(self._counts).cpu()
        losses = torch.stack([l.float() for l in self._losses]).cpu()
        return counts, losses

    def get_latest_loss(self) -> float:
        return float(self._losses[-1].item()) if self._losses else float("nan")

    def get_running_loss(self) -> float:
        return float(torch.stack([l.float() for l in self._losses]).mean().item()) if self._losses else float("nan")

    def evaluate(self) -> dict:
        """One device->host copy for the whole epoch.  -> {"batches": [{"loss", "metrics": {...}}], "average": {...}}
        with the reference's per-batch metric names; the epoch average is the mean of the per-batch values, as
        SingleEpochHistory.get_average computes it (src/train/history.py:36-60)."""
        counts, losses = self._materialise()
        batches = []
        for i in range(counts.shape[0]):
            cd, tc, wd, tw = (int(v) for v in counts[i])
            m = {}
            for k, v in self._extra[i]:
                m[k] = float(v.item()) if isinstance(v, torch.Tensor) else v
            if tw > 0 or tc > 0:
                m["word_error_rate"] = wd / tw if tw > 0 else float("nan")
                if tc > 0:
                    m["char_error_rate"] = cd / tc
            batches.append({"loss": float(losses[i]), "metrics": m})
        avg: Dict[str, float] = {}
        if batches:
            keys = set().union(*(b["metrics"].keys() for b in batches))
            for k in keys:
                vals = [b["metrics"][k] for b in batches if k in b["metrics"]]
                avg[k] = sum(vals) / len(vals)
            avg["loss"] = sum(b["loss"] for b in batches) / len(batches)
        tot = counts.sum(0) if counts.numel() else torch.zeros(4, dtype=torch.int64)
        return {"batches": batches, "average": avg,
                "totals": {"char_distance": int(tot[0]), "target_chars": int(tot[1]), "word_distance": int(tot[2]),
                           "target_words": int(tot[3])}}

    def clean_up(self) -> None:
        self._counts, self._losses, self._extra = [], [], []

    # ---- strings, on request only (test-mode prediction dumps; evaluator.py:108-120) ----------------------------
    def decode_predictions(self, predictions, sample=None):
        am, ids, lens = ops.greedy_decode(predictions.logits.detach())
        ids_c, lens_c = ids.cpu(), lens.cpu()
        preds = [self._ids_to_string(ids_c[b, :lens_c[b]].tolist(), cut=True) for b in range(ids_c.shape[0])]
        labels = None
        if sample is not None and sample.target is not None:
            labels = [self._ids_to_string([t for t in row if t != self.pad_id], cut=False) for row in sample.target.cpu().tolist()]
        return preds, labels

    def _ids_to_string(self, ids: List[int], cut: bool) -> str:
        s = "".join(" " if i == self.delim_id else self.vocab[i] for i in ids).strip()
        if self.cleanup:
            s = s.replace(" ' ", "'")
        if cut and self.eos_id >= 0:
            e = self.vocab[self.eos_id]
            k = s.find(e)
            if k != -1:
                s = s[:k + len(e)]
        return s
