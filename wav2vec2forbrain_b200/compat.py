"""Drop-in glue: run the reference's own `run.py` / `Experiment` / `Trainer` code, unchanged, on the B200 model.

    import wav2vec2forbrain_b200.compat as compat
    compat.install("/path/to/Wav2Vec2ForBrain")        # before anything of `src.*` is imported
    from src.args.argparsing import get_experiment_from_args
    get_experiment_from_args().run()                    # == the reference's run.py:1-5

(`run_b200.py` at the repository root is exactly that.)  Nothing under the reference checkout is modified.  What
`install` does, in order:

 1. environment shims for the box this runs on — pydantic 2 exposes the v1 API the reference is written against
    (`src/args/argparsing.py:42-63` uses `__fields__`, `field.type_`, `outer_type_`) as `pydantic.v1`; modules the
    reference imports at module scope but that are not installed offline (`torcheval`, `edit_distance`, `pyctcdecode`,
    `g2p_en`, `git`, `matplotlib`, `wandb`) get small stand-ins — the two the hot path's callers really use
    (`torcheval.metrics.WordErrorRate`, `edit_distance.SequenceMatcher.distance`, `src/train/evaluator.py:1,14,163-229`)
    are working implementations, the rest raise when used;
 2. offline factories: `AutoTokenizer.from_pretrained` falls back to a local `Wav2Vec2CTCTokenizer` with the
    wav2vec2-base-960h vocabulary when the hub / cache is unreachable;
 3. THE SWAP: the names `bfe_w_preprocessing_from_config`, `W2VBrainEncoderModel`,
    `W2VConformerBrainEncoderModel` inside `src.experiments.b2t_gru_w2v_experiment` /
    `src.experiments.b2t_gru_w2v_conformer_experiment` are rebound to this package's classes, so the reference's
    `_create_model` (`b2t_gru_w2v_experiment.py:95-107`, `b2t_gru_w2v_conformer_experiment.py:78-85`) builds the
    B200 model; `unfreeze_strategy == "brain_encoder"` additionally switches the w2v weight gradients off;
 4. optional accelerations of the rows next to the model, all flag-free: `Optimizers["adam"]` ->
    `optim.FusedAdam`, `torch.nn.utils.clip_grad_norm_` -> the fused clip when every gradient is a CUDA fp32 tensor,
    `create_evaluator` -> the GPU evaluator adapter for train / val epochs;
 5. data parallel when launched under torchrun (WORLD_SIZE > 1): process group, `dist.DataParallel` around the
    created model, every rank keeps rows rank::world of each (identically shuffled) global batch.
"""
from __future__ import annotations

import importlib
import importlib.machinery
import os
import sys
import types
from typing import Optional

import torch

_INSTALLED = False


# -----------------------------------------------------------------------------------------------------------
# 1. environment shims
# -----------------------------------------------------------------------------------------------------------
def _stub(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, None)     # transformers' availability probes read __spec__
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _missing(name: str) -> bool:
    if name in sys.modules:
        return False
    try:
        return importlib.util.find_spec(name) is None
    except (ImportError, ValueError):
        return True


def _levenshtein(a, b) -> int:
    prev = list(range(len(b) + 1))
    for i, x in enumerate(a, 1):
        cur = [i]
        for j, y in enumerate(b, 1):
            cur.append(min(prev[j] + 1, cur[j - 1] + 1, prev[j - 1] + (x != y)))
        prev = cur
    return prev[-1]


class _WordErrorRate:
    """torcheval.metrics.WordErrorRate: sum of word-level edit distances / sum of target words."""

    def __init__(self):
        self.errors, self.total = 0.0, 0.0

    def update(self, input, target):
        if isinstance(input, str):
            input, target = [input], [target]
        for p, t in zip(input, target):
            self.errors += _levenshtein(p.split(), t.split())
            self.total += len(t.split())
        return self

    def compute(self):
        return torch.tensor(self.errors / self.total if self.total > 0 else float("nan"), dtype=torch.float64)


class _SequenceMatcher:
    """edit_distance.SequenceMatcher: only .distance() (unit-cost Levenshtein) is used (evaluator.py:231-242)."""

    def __init__(self, a=None, b=None):
        self.a, self.b = a or "", b or ""

    def distance(self):
        return _levenshtein(self.a, self.b)


def _unavailable(what: str):
    def raiser(*a, **k):
        raise RuntimeError(f"{what} is not installed on this box (stand-in from wav2vec2forbrain_b200.compat)")
    return raiser


def _install_shims() -> None:
    # third-party packages that read the real pydantic must be imported BEFORE the alias goes in
    import transformers  # noqa: F401
    if _missing("wandb"):
        class _Run:
            def __enter__(self):
                return self

            def __exit__(self, *exc):
                return False

            def log_artifact(self, *a, **k):
                pass
        w = _stub("wandb", login=lambda *a, **k: None, log=lambda *a, **k: None, watch=lambda *a, **k: None,
                  Artifact=lambda *a, **k: types.SimpleNamespace(add_dir=lambda *a, **k: None), run=None)

        def _init(*a, **k):
            w.run = _Run()
            return w.run
        w.init = _init
        sdk = _stub("wandb.sdk")
        wr = _stub("wandb.sdk.wandb_run", Run=_Run)
        w.sdk, sdk.wandb_run = sdk, wr
    else:
        import wandb  # noqa: F401
    import pydantic
    if int(pydantic.VERSION.split(".")[0]) >= 2:
        import pydantic.v1 as pv1
        sys.modules["pydantic"] = pv1
    if _missing("torcheval"):
        te = _stub("torcheval")
        te.metrics = _stub("torcheval.metrics", WordErrorRate=_WordErrorRate)
    if _missing("edit_distance"):
        _stub("edit_distance", SequenceMatcher=_SequenceMatcher)
    if _missing("pyctcdecode"):
        pc = _stub("pyctcdecode", build_ctcdecoder=_unavailable("pyctcdecode"))
        # pyctcdecode/constants.py values (defaults of the lm_decode_* flags, b2t_gru_w2v_experiment.py:5-9,68-70)
        pc.constants = _stub("pyctcdecode.constants", DEFAULT_BEAM_WIDTH=100, DEFAULT_MIN_TOKEN_LOGP=-5.0,
                             DEFAULT_PRUNE_LOGP=-10.0)
    if _missing("g2p_en"):
        _stub("g2p_en", G2p=_unavailable("g2p_en"))
    if _missing("git"):
        import typing
        _stub("git", Optional=typing.Optional)       # `from git import Optional` (experiment.py:3)
    if _missing("matplotlib"):
        mp = _stub("matplotlib", use=lambda *a, **k: None)
        mp.pyplot = _stub("matplotlib.pyplot", **{n: _unavailable("matplotlib") for n in
                                                  ("figure", "subplots", "plot", "savefig", "close", "legend", "show")})
        mp.figure = _stub("matplotlib.figure", Figure=object)
        mp.axes = _stub("matplotlib.axes", Axes=object)


def _install_scheduler_shim() -> None:
    """torch >= 2.7 removed the `verbose=` argument of the LR schedulers that the reference (pinned to torch 2.1.1,
    environment.yaml:245) passes (src/util/warmup_scheduler.py:49-56, src/experiments/experiment.py:357-364)."""
    import inspect
    from torch.optim import lr_scheduler
    for name in ("LambdaLR", "StepLR", "ExponentialLR", "MultiStepLR", "CosineAnnealingLR", "ReduceLROnPlateau", "LinearLR"):
        cls = getattr(lr_scheduler, name, None)
        if cls is None or "verbose" in inspect.signature(cls.__init__).parameters or getattr(cls, "_b2t_verbose", False):
            continue
        real = cls.__init__

        def __init__(self, *a, _real=real, **k):
            k.pop("verbose", None)
            _real(self, *a, **k)
        cls.__init__ = __init__
        cls._b2t_verbose = True


# -----------------------------------------------------------------------------------------------------------
# 2. offline tokenizer
# -----------------------------------------------------------------------------------------------------------
def local_tokenizer(cache_dir: Optional[str] = None):
    """Wav2Vec2CTCTokenizer with the facebook/wav2vec2-base-960h vocabulary, built locally."""
    import json
    import tempfile
    from transformers import Wav2Vec2CTCTokenizer
    from .evaluator import DEFAULT_VOCAB
    d = tempfile.mkdtemp(prefix="b2t_vocab_", dir=cache_dir if cache_dir and os.path.isdir(cache_dir) else None)
    path = os.path.join(d, "vocab.json")
    with open(path, "w") as f:
        json.dump({t: i for i, t in enumerate(DEFAULT_VOCAB)}, f)
    return Wav2Vec2CTCTokenizer(path, unk_token="<unk>", pad_token="<pad>", bos_token="<s>", eos_token="</s>",
                                word_delimiter_token="|", do_lower_case=False, clean_up_tokenization_spaces=True)


def _install_offline_tokenizer() -> None:
    import transformers
    real = transformers.AutoTokenizer.from_pretrained

    def from_pretrained(name, *a, **kw):
        try:
            return real(name, *a, **kw)
        except Exception:
            if "wav2vec2" in str(name):
                return local_tokenizer(kw.get("cache_dir"))
            raise
    transformers.AutoTokenizer.from_pretrained = staticmethod(from_pretrained)


# -----------------------------------------------------------------------------------------------------------
# 3.-5. the swap, accelerations, data parallel
# -----------------------------------------------------------------------------------------------------------
def _wrap_create_model(cls, fused: bool) -> None:
    from . import dist as bdist
    real = cls._create_model

    def _create_model(self):
        model = real(self)
        if getattr(self.config, "unfreeze_strategy", None) == "brain_encoder":
            # only the brain encoder is optimised (b2t_gru_w2v_experiment.py:129-133): dgrad-only w2v backward
            model.runtime.skip_w2v_wgrad = True
        wn, co = getattr(self.config, "whiteNoiseSD", 0.0), getattr(self.config, "constantOffsetSD", 0.0)
        if os.environ.get("B2T_APPLY_AUGMENTATION") == "1":
            model.runtime.augment = (float(wn), float(co))      # the reference discards its augmentation (fact 3)
        _, _, world = bdist.env_rank_world()
        if world > 1:
            bdist.init_process_group()
            model._dp = bdist.DataParallel(model.cuda())
            bdist.tune_gc()      # a full garbage collection in one rank stalls every rank at the next collective
        return model
    cls._create_model = _create_model


def _shard_dataloaders(exp_cls) -> None:
    """WORLD_SIZE > 1: the reference's DataLoader draws global batches of batch_size * world with the SAME shuffle on
    every rank (same torch seed, experiment.py:34); rank r collates rows r::world of each."""
    from torch.utils.data import DataLoader
    from . import dist as bdist
    rank, _, world = bdist.env_rank_world()
    if world <= 1:
        return
    real = exp_cls._create_dataloader

    def _create_dataloader(self, split):
        dl = real(self, split)
        inner = dl.collate_fn
        if dl.batch_sampler is not None and getattr(dl, "batch_size", None) is None:
            return dl          # custom batch sampler (--day_batches): left alone, every rank sees every batch
        return DataLoader(dl.dataset, batch_size=dl.batch_size * world, shuffle=split == "train",
                          collate_fn=lambda samples: inner(samples[rank::world] or samples[:1]))
    exp_cls._create_dataloader = _create_dataloader


def make_gpu_evaluator_class():
    """Adapter with the reference's Evaluator interface (src/train/evaluator.py:24-47) over evaluator.GpuEvaluator:
    per-batch WER / CER / loss stay on the device, `evaluate()` builds the reference's SingleEpochHistory from one
    device->host copy."""
    from src.train.evaluator import Evaluator
    from src.train.history import MetricEntry, SingleEpochHistory
    from .evaluator import GpuEvaluator

    class GpuEvaluatorAdapter(Evaluator):
        def __init__(self, tokenizer, mode, track_non_test_predictions: bool = False):
            super().__init__(mode, track_non_test_predictions)
            vocab = tokenizer.convert_ids_to_tokens(list(range(tokenizer.vocab_size)))
            self.impl = GpuEvaluator(mode, vocab=vocab, pad_id=tokenizer.pad_token_id,
                                     delim_id=tokenizer.convert_tokens_to_ids(tokenizer.word_delimiter_token),
                                     eos_id=tokenizer.eos_token_id,
                                     cleanup=bool(getattr(tokenizer, "clean_up_tokenization_spaces", True)))

        def track_batch(self, predictions, sample):      # no .item(): nothing here synchronises with the device
            self.impl.track_batch(predictions, sample)
            self.n_losses += 1

        def _track_batch(self, predictions, sample):      # pragma: no cover - track_batch is overridden
            self.impl.track_batch(predictions, sample)

        def get_running_loss(self):
            return self.impl.get_running_loss()

        def get_latest_loss(self):
            return self.impl.get_latest_loss()

        def evaluate(self):
            res = self.impl.evaluate()
            hist = SingleEpochHistory()
            for b in res["batches"]:
                hist.add_batch_metric(MetricEntry(dict(b["metrics"]), b["loss"]), None)
            return hist

        def clean_up(self):
            self.impl.clean_up()
    return GpuEvaluatorAdapter


def _install_fused_clip() -> None:
    from . import optim
    real = torch.nn.utils.clip_grad_norm_
    if getattr(real, "_b2t_fused", False):
        return

    def clip_grad_norm_(parameters, max_norm, norm_type=2.0, error_if_nonfinite=False, foreach=None):
        params = [parameters] if isinstance(parameters, torch.Tensor) else list(parameters)
        grads = [p.grad for p in params if p.grad is not None]
        if grads and float(norm_type) == 2.0 and all(g.is_cuda and g.dtype == torch.float32 and g.is_contiguous() for g in grads):
            return optim.clip_grad_norm_(params, max_norm, norm_type, error_if_nonfinite)
        return real(params, max_norm, norm_type, error_if_nonfinite, foreach)
    clip_grad_norm_._b2t_fused = True
    torch.nn.utils.clip_grad_norm_ = clip_grad_norm_


def install(reference_root: Optional[str] = None, fused_optimizer: bool = True, gpu_evaluator: bool = True) -> None:
    """See the module docstring.  `reference_root`: checkout of yuanhao-chen-nyoeghau/Wav2Vec2ForBrain (default: env
    B2T_REFERENCE, else the current directory); it is put on sys.path, never written to."""
    global _INSTALLED
    if _INSTALLED:
        return
    root = reference_root or os.environ.get("B2T_REFERENCE") or os.getcwd()
    if not os.path.isdir(os.path.join(root, "src", "experiments")):
        raise RuntimeError(f"{root} is not a Wav2Vec2ForBrain checkout (no src/experiments); pass reference_root or set "
                           "B2T_REFERENCE")
    if root not in sys.path:
        sys.path.insert(0, root)
    _install_shims()
    _install_scheduler_shim()
    _install_offline_tokenizer()
    from . import models, optim
    import src.experiments.b2t_gru_w2v_experiment as e_base
    import src.experiments.b2t_gru_w2v_conformer_experiment as e_conf
    import src.experiments.experiment as e_exp
    import src.experiments.b2t_experiment as e_b2t
    # ---- the swap: the reference's _create_model now builds the B200 model ----
    e_base.bfe_w_preprocessing_from_config = models.bfe_w_preprocessing_from_config
    e_base.W2VBrainEncoderModel = models.W2VBrainEncoderModel
    e_conf.bfe_w_preprocessing_from_config = models.bfe_w_preprocessing_from_config
    e_conf.W2VConformerBrainEncoderModel = models.W2VConformerBrainEncoderModel
    for cls in (e_base.B2TGruAndW2VExperiment, e_conf.B2TGruAndW2VConformerExperiment):
        _wrap_create_model(cls, fused_optimizer)
    _shard_dataloaders(e_b2t.B2TExperiment)
    if fused_optimizer:
        e_exp.Optimizers["adam"] = optim.FusedAdam
        _install_fused_clip()
    if gpu_evaluator:
        adapter = make_gpu_evaluator_class()
        for cls in (e_base.B2TGruAndW2VExperiment, e_conf.B2TGruAndW2VConformerExperiment):
            real = cls.create_evaluator

            def create_evaluator(self, mode, track_non_test_predictions=False, _real=real):
                # test mode keeps the reference evaluator (prediction dumps, optional LM beam decode on the host)
                if mode == "test" or track_non_test_predictions:
                    return _real(self, mode, track_non_test_predictions)
                return adapter(self.tokenizer, mode, track_non_test_predictions)
            cls.create_evaluator = create_evaluator
    _INSTALLED = True


def install_synthetic_dataset(n_trials: int = 64, t_range=(200, 500), seed: int = 0) -> None:
    """Offline runs (the Brain-to-Text '24 .mat files are a private download): `B2TExperiment._create_dataset` returns a
    synthetic split — z-scored-like N(0,1) features [T,256], random upper-case sentences — through the reference's OWN
    Brain2TextDataset.__getitem__ / get_collate_fn (src/datasets/brain2text.py:149-213); only __init__ (the .mat
    reader) is bypassed.  Also selected by env B2T_SYNTHETIC_DATA=<n_trials> in run_b200.py."""
    import numpy as np
    import src.experiments.b2t_experiment as e_b2t
    from src.datasets.brain2text import B2tSample, Brain2TextDataset

    class SyntheticBrain2TextDataset(Brain2TextDataset):
        def __init__(self, config, yaml_config, split="train", tokenizer=None):
            torch.utils.data.Dataset.__init__(self)
            self.config, self.tokenizer = config, tokenizer
            rng = np.random.default_rng(seed + {"train": 0, "val": 1, "test": 2}[split])
            words = ["THE", "BRAIN", "SPEAKS", "IT'S", "A", "TEST", "OF", "WORDS", "AND", "SIGNALS", "DON'T", "STOP"]
            self.samples = []
            n = n_trials if split == "train" else max(2, n_trials // 4)
            for _ in range(n):
                T = int(rng.integers(t_range[0], t_range[1] + 1))
                s = B2tSample(torch.from_numpy(rng.standard_normal((T, 256)).astype(np.float32)),
                              " ".join(rng.choice(words, int(rng.integers(2, 7)))))
                s.day_idx = int(rng.integers(0, 24))
                self.samples.append(s)

    e_b2t.B2TExperiment._create_dataset = lambda self, split="train": SyntheticBrain2TextDataset(
        self.config, self.yaml_config, split, self.tokenizer)
