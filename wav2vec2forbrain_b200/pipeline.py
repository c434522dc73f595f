"""Input pipeline: device-resident trials, block z-scoring and collation on the device,
length-bucketed batches, data-parallel sharding.

Replaces, for the hot path only, what the reference does on one CPU thread per batch (num_workers=0):
`Brain2TextDataset.__init__` block-wise feature normalisation (src/datasets/preprocessing.py:60-72,100-113 via
src/datasets/brain2text.py:111-135), label clean-up + tokenisation and the collate function
(src/datasets/brain2text.py:178-211), followed by `batch.cuda()` (src/train/train_loop.py:42).

Layout for 180 GB of HBM: the whole split lives on the device as ONE flat fp32 array `[sum_T, 256]` (the full
Brain-to-Text '24 training split is ~2.7 GB) with int64 trial offsets; labels are a padded int64 table.  A batch is
built by one kernel launch from an index vector — no per-step host tensor work and no H2D copy besides the B indices
(pinned, asynchronous).  Batches are formed from length buckets so that padded frames (which the model computes at
full cost: no attention mask, no packed GRU — SURVEY.md fact 4) are few; `bucket_width=None` reproduces the reference's
plain shuffled batches.  Rank r of N takes batches r::N of the identically seeded epoch order.

The .mat reader and the tokenizer stay the reference's (out of scope): `TrialStore.from_trials`
takes the per-trial arrays and label strings they produce.
"""
from __future__ import annotations

import ctypes
import re
from typing import Iterator, List, Optional, Sequence

import numpy as np
import torch

from . import _lib, ops
from ._lib import c_int, check, lib, ptr, stream
from .evaluator import DEFAULT_VOCAB
from .models import B2tSampleBatch

_PUNCT = re.compile(r'[\,\?\.\!\-\;\:"]')      # brain2text.py:194


def tokenize_label(text: str, vocab: Sequence[str] = DEFAULT_VOCAB, remove_punctuation: bool = True, unk_id: int = 3) -> List[int]:
    """brain2text.py:137,192-197 + Wav2Vec2CTCTokenizer's character tokenisation (' ' -> '|', unknown -> <unk>)."""
    text = text.upper()
    if remove_punctuation:
        text = _PUNCT.sub("", text)
    table = {t: i for i, t in enumerate(vocab)}
    return [table.get("|" if c == " " else c, unk_id) for c in text]


class TrialStore:
    """All trials of a split, resident on one device."""

    def __init__(self, frames: torch.Tensor, offsets: torch.Tensor, labels: torch.Tensor, label_lens: torch.Tensor,
                 days: torch.Tensor, lengths_host: np.ndarray, label_lens_host: np.ndarray):
        self.frames, self.offsets, self.labels, self.label_lens, self.days = frames, offsets, labels, label_lens, days
        self.lengths_host, self.label_lens_host = lengths_host, label_lens_host
        self.device = frames.device

    def __len__(self) -> int:
        return len(self.lengths_host)

    @classmethod
    def from_trials(cls, features: Sequence[np.ndarray], transcriptions: Sequence[str], day_idxs: Sequence[int],
                    device, blocks: Optional[Sequence[Sequence[int]]] = None, vocab: Sequence[str] = DEFAULT_VOCAB,
                    remove_punctuation: bool = True) -> "TrialStore":
        """features[i]: [T_i, C] float array of trial i (raw when `blocks` is given: each entry of `blocks` is the list
        of trial indices of one recording block and the features are z-scored per block ON THE DEVICE; already
        normalised otherwise)."""
        lengths = np.asarray([f.shape[0] for f in features], dtype=np.int64)
        C = features[0].shape[1]
        offs = np.zeros(len(features) + 1, dtype=np.int64)
        np.cumsum(lengths, out=offs[1:])
        host = torch.empty(int(offs[-1]), C, dtype=torch.float32, pin_memory=torch.cuda.is_available())
        hv = host.numpy()
        for i, f in enumerate(features):
            hv[offs[i]:offs[i + 1]] = f
        frames = host.to(device, non_blocking=True)
        offsets = torch.from_numpy(offs).to(device)
        ids = [tokenize_label(t, vocab, remove_punctuation) for t in transcriptions]
        smax = max(1, max(len(x) for x in ids))
        lab = np.zeros((len(ids), smax), dtype=np.int64)
        for i, x in enumerate(ids):
            lab[i, :len(x)] = x
        # nn_helper.calc_seq_len (:52-57): index after the last id > 0
        tl = np.asarray([max([j + 1 for j, v in enumerate(row) if v > 0], default=0) for row in lab], dtype=np.int64)
        store = cls(frames, offsets, torch.from_numpy(lab).to(device), torch.from_numpy(tl).to(device),
                    torch.as_tensor(np.asarray(day_idxs, dtype=np.int64)).to(device), lengths, tl)
        if blocks is not None:
            store.zscore_blocks(blocks)
        return store

    def zscore_blocks(self, blocks: Sequence[Sequence[int]], eps: float = 1e-8) -> None:
        """In-place block-wise z-scoring (preprocessing.py:60-72): a block's statistics span the trials
        blocks[b][0] .. blocks[b][-1] (the reference concatenates that RANGE), and are applied to the listed trials."""
        _lib.ensure_device()
        n = len(self)
        stat_block = np.full(n, -1, dtype=np.int32)
        for b, idx in enumerate(blocks):
            stat_block[int(idx[0]):int(idx[-1]) + 1] = b
        apply_block = np.full(n, -1, dtype=np.int32)
        for b, idx in enumerate(blocks):
            apply_block[np.asarray(idx, dtype=np.int64)] = b
        if not np.array_equal(stat_block, apply_block):
            raise NotImplementedError("blocks whose index range holds trials of another block")
        tb = torch.from_numpy(apply_block).to(self.device)
        C = self.frames.shape[1]
        ws = torch.empty(3 * len(blocks) * C, dtype=torch.float64, device=self.device)
        check(lib().b2t_block_zscore(ptr(self.frames), ptr(self.frames), ptr(self.offsets), ptr(tb), c_int(n),
                                     c_int(len(blocks)), c_int(C), ptr(ws), ctypes.c_double(eps), stream()),
              "b2t_block_zscore")
        ops._count(3)

    def collate(self, idx_dev: torch.Tensor, t_max: int, s_max: int) -> B2tSampleBatch:
        """The reference's collate (brain2text.py:178-211) for the trials idx_dev (int32, device) as one launch."""
        _lib.ensure_device()
        B = idx_dev.numel()
        C = self.frames.shape[1]
        dev = self.device
        x = torch.empty(B, t_max, C, dtype=torch.float32, device=dev)
        tg = torch.empty(B, s_max, dtype=torch.int64, device=dev)
        meta = torch.empty(3, B, dtype=torch.int64, device=dev)
        check(lib().b2t_collate(ptr(self.frames), ptr(self.offsets), ptr(idx_dev), c_int(B), c_int(C), c_int(t_max), ptr(x),
                                ptr(self.labels), c_int(self.labels.shape[1]), ptr(self.label_lens), ptr(self.days), ptr(tg),
                                c_int(s_max), ptr(meta[0]), ptr(meta[1]), ptr(meta[2]), stream()), "b2t_collate")
        ops._count()
        batch = B2tSampleBatch(x, tg)
        batch.input_lens, batch.target_lens, batch.day_idxs = meta[0], meta[1], meta[2]
        return batch


def bucketed_batches(lengths: np.ndarray, batch_size: int, seed: int, epoch: int, bucket_width: Optional[int] = 32,
                     shuffle: bool = True, drop_last: bool = False) -> List[np.ndarray]:
    """Epoch order of batches (lists of trial indices).  bucket_width=None: the reference's DataLoader(shuffle=True)
    semantics (random permutation cut into batches).  Otherwise trials are grouped by ceil(len / bucket_width), shuffled
    inside each group, cut into batches, the ragged tails of all groups are merged (sorted by length) into batches of
    their own, and the batches are shuffled: every trial is still seen exactly once per epoch and padding per full
    batch is bounded by bucket_width frames."""
    rng = np.random.default_rng([seed, epoch])
    n = len(lengths)
    if bucket_width is None:
        order = rng.permutation(n) if shuffle else np.arange(n)
        batches = [order[i:i + batch_size] for i in range(0, n, batch_size)]
    else:
        key = (np.asarray(lengths) + bucket_width - 1) // bucket_width
        batches, tails = [], []
        for k in np.unique(key):
            members = np.flatnonzero(key == k)
            if shuffle:
                members = rng.permutation(members)
            full = len(members) // batch_size * batch_size
            batches += [members[i:i + batch_size] for i in range(0, full, batch_size)]
            if full < len(members):
                tails.append(members[full:])
        if tails:      # np.unique returns the keys sorted, so neighbouring tails have neighbouring lengths
            rest = np.concatenate(tails)
            batches += [rest[i:i + batch_size] for i in range(0, len(rest), batch_size)]
        if shuffle:
            batches = [batches[i] for i in rng.permutation(len(batches))]
    if drop_last:
        batches = [b for b in batches if len(b) == batch_size]
    return batches


class DeviceLoader:
    """for batch in loader: batch is a B2tSampleBatch already on the device (train_loop.py:41-42's `.cuda()` is then a
    no-op).  The index vectors of the whole epoch are staged in one pinned buffer and copied once."""

    def __init__(self, store: TrialStore, batch_size: int, seed: int = 42, bucket_width: Optional[int] = 32,
                 shuffle: bool = True, drop_last: bool = False, rank: int = 0, world: int = 1):
        self.store, self.batch_size, self.seed = store, batch_size, seed
        self.bucket_width, self.shuffle, self.drop_last = bucket_width, shuffle, drop_last
        self.rank, self.world = rank, world
        self.epoch = 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def epoch_batches(self) -> List[np.ndarray]:
        b = bucketed_batches(self.store.lengths_host, self.batch_size, self.seed, self.epoch, self.bucket_width, self.shuffle,
                             self.drop_last)
        if self.world > 1:
            b = b[:len(b) // self.world * self.world]     # every rank steps the same number of times (all-reduce)
            b = b[self.rank::self.world]
        return b

    def __len__(self) -> int:
        return len(self.epoch_batches())

    def padded_fraction(self) -> float:
        """frames of padding / frames computed, over one epoch (what bucketing buys)."""
        L = self.store.lengths_host
        tot = pad = 0
        for b in self.epoch_batches():
            t = int(L[b].max())
            tot += t * len(b)
            pad += t * len(b) - int(L[b].sum())
        return pad / max(tot, 1)

    def __iter__(self) -> Iterator[B2tSampleBatch]:
        batches = self.epoch_batches()
        self.epoch += 1
        if not batches:
            return
        flat = np.concatenate(batches).astype(np.int32)
        host = torch.from_numpy(flat)
        if torch.cuda.is_available():
            host = host.pin_memory()
        dev = host.to(self.store.device, non_blocking=True)
        L, S = self.store.lengths_host, self.store.label_lens_host
        off = 0
        for b in batches:
            idx = dev[off:off + len(b)]
            off += len(b)
            yield self.store.collate(idx, int(L[b].max()), max(1, int(S[b].max())))
