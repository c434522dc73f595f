"""Subprocess body of tests/test_dropin_cpu.py (compat.install rebinds sys.modules["pydantic"] and reference module
globals, so it gets a process of its own).  TEST INFRASTRUCTURE: runs the reference's OWN argument parser, Experiment
and Trainer code from /root/reference on the product model, with tests/abi_emulator.py standing in for the CUDA
library (this container has no GPU) and `.cuda()` made an identity."""
import json
import math
import os
import sys
import tempfile

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
REF = os.environ.get("B2T_REFERENCE", "/root/reference")
kind = sys.argv[1]                      # "base" | "conformer"
sys.argv = [sys.argv[0]]

# ---- CPU harness (SURVEY.md §8c patch 2; on a GPU box none of this block is needed) ----
torch.Tensor.cuda = lambda self, *a, **k: self
torch.nn.Module.cuda = lambda self, *a, **k: self
from tests import abi_emulator  # noqa: E402
from wav2vec2forbrain_b200 import compat, engine, models  # noqa: E402
from wav2vec2forbrain_b200.config import W2VConfig  # noqa: E402
engine.ops = abi_emulator
models._EMULATED = True
_real_load = torch.load
torch.load = lambda f, *a, **k: _real_load(f, *a, **{**k, "map_location": "cpu"})

compat.install(REF, fused_optimizer=False, gpu_evaluator=False)     # the fused optimizer / evaluator kernels need the GPU
compat.install_synthetic_dataset(n_trials=8, t_range=(80, 120))

# a tiny architecture behind the checkpoint names the reference's flags and tables know
TINY = dict(hidden=48, layers=2, heads=4, ffn=96, vocab=32, pos_conv_k=16, pos_conv_groups=4)
if kind == "base":
    ck, exp_type = "facebook/wav2vec2-base-960h", "b2p2t_gru+w2v"
    models.register_architecture(ck, lambda **kw: W2VConfig(**dict(dict(kind="base", **TINY), **kw)), 48)
else:
    ck, exp_type = "facebook/wav2vec2-conformer-rope-large-960h-ft", "b2p2t_gru+w2v_conformer"
    models.register_architecture(ck, lambda **kw: W2VConfig(**dict(dict(kind="conformer", hidden_act="swish", dw_kernel=7, **TINY), **kw)), 48)

work = tempfile.mkdtemp(prefix="b2t_dropin_")
os.chdir(work)
with open("config.yaml", "w") as f:
    f.write(f"cache_dir: {work}\nfig_dir: {work}\nn3gram_lm_model_path: none\nn5gram_lm_model_path: none\n"
            f"dataset_splits_dir: {work}\nwandb_api_key: none\ntimit_dataset_splits_dir: {work}\nelevenlabs_api_key: none\n")

# ---- real run.py flags (README.md:34 / sweeps/*.yaml:12-30 flag set, sizes shrunk) ----
sys.argv += [f"--experiment_type={exp_type}", "--encoder_gru_hidden_size=32", "--encoder_num_gru_layers=2",
             "--encoder_fc_hidden_sizes=[24]", "--gaussian_smooth_width=1.5", "--encoder_dropout=0.45", "--whiteNoiseSD=0.02",
             "--constantOffsetSD=0.25", "--weight_decay=1e-4", "--learning_rate=1e-3", "--w2v_learning_rate=1e-4",
             "--unfreeze_strategy=brain_encoder+w2v", "--batch_size=4", "--epochs=1", "--gradient_clipping=1.0",
             "--use_prefix_beam_search=false", f"--wav2vec_checkpoint={ck}", f"--tokenizer_checkpoint={ck}",
             "--return_best_model=false"]
if kind == "base":
    sys.argv.append("--w2v_skip_loading_weights=true")

from src.args.argparsing import get_experiment_from_args  # noqa: E402
from src.train.train_loop import Trainer  # noqa: E402

experiment = get_experiment_from_args()                 # run.py:4
model = experiment.model
assert type(model).__module__ == "wav2vec2forbrain_b200.models", type(model)
assert type(model).__name__ == ("W2VBrainEncoderModel" if kind == "base" else "W2VConformerBrainEncoderModel")
trainer = Trainer(experiment)                           # experiment.run() -> Trainer(self), experiment.py:100
assert len(trainer.optimizer.param_groups) == 2 and trainer.optimizer.param_groups[1]["lr"] == 1e-4
before = {k: v.detach().clone() for k, v in model.named_parameters()}
hist = trainer._train_epoch(experiment.dataloader_train)          # train_loop.py:37-87, unmodified
avg = hist.get_average()
assert math.isfinite(avg.loss) and "word_error_rate" in avg.metrics and "ctc_loss" in avg.metrics
moved = [k for k, v in model.named_parameters() if not torch.equal(v, before[k])]
unused = [k for k in before if "inpLayer" in k or k.endswith("hidden_start") or ("conformer" in k and "pos_conv_embed" in k)]
assert moved and not set(moved) & set(unused), "Adam must not move parameters that never get a gradient"
val = trainer._evaluate_epoch("val")                              # train_loop.py:89-109
assert math.isfinite(val.get_average().loss)

# ---- checkpoint interchange with a model built by the reference's OWN classes (src/model/*, untouched by the swap) ----
import transformers  # noqa: E402
from src.model import brain_feature_extractor as ref_bfe  # noqa: E402
ref_bfe.PRETRAINED_LATENT_SIZES[ck] = 48
if kind == "base":
    from src.model.w2v_custom_feat_extractor import W2VBrainEncoderModel as RefModel
    transformers.Wav2Vec2Config.from_pretrained = classmethod(lambda cls, name, **kw: transformers.Wav2Vec2Config(
        hidden_size=48, num_hidden_layers=2, num_attention_heads=4, intermediate_size=96, vocab_size=32,
        num_conv_pos_embeddings=16, num_conv_pos_embedding_groups=4, pad_token_id=0, **kw))
    ref = RefModel(experiment.config, ref_bfe.bfe_w_preprocessing_from_config(experiment.config, None, ck), ck, None, True)
else:
    from src.model import w2v_conformer_custom_feat_extractor as cm
    transformers.Wav2Vec2ConformerConfig.from_pretrained = classmethod(
        lambda cls, name, **kw: transformers.Wav2Vec2ConformerConfig(
            hidden_size=48, num_hidden_layers=2, num_attention_heads=4, intermediate_size=96, vocab_size=32,
            num_conv_pos_embeddings=16, num_conv_pos_embedding_groups=4, pad_token_id=0, hidden_act="swish",
            position_embeddings_type="rotary", conv_depthwise_kernel_size=7, **kw))
    cm.Wav2Vec2ConformerWithoutFeatExtrForCTC.from_pretrained = classmethod(lambda cls, name, config=None, **kw: cls(config))
    ref = cm.W2VConformerBrainEncoderModel(ref_bfe.bfe_w_preprocessing_from_config(experiment.config, None, ck), ck)
ref.w2v_encoder.config._attn_implementation = "eager"
path = os.path.join(work, "reference_model.pt")
torch.save(ref.state_dict(), path)                                  # experiment.py:137-141 store_trained_model
res = model.load_state_dict(torch.load(path), strict=True)          # experiment.py:72-75 --from_checkpoint
assert not res.missing_keys and not res.unexpected_keys
res = ref.load_state_dict(model.state_dict(), strict=True)          # and the product's checkpoint loads into the reference
assert not res.missing_keys and not res.unexpected_keys
batch = next(iter(experiment.dataloader_val))
model.eval(), ref.eval()
with torch.no_grad():
    a, b = model.forward(batch), ref.forward(batch)
err = ((a.logits - b.logits).norm() / b.logits.norm()).item()
assert err < 2e-4, err
assert abs(a.loss.item() - b.loss.item()) < 2e-4 * abs(b.loss.item())
assert torch.equal(a.logits.argmax(-1), b.logits.argmax(-1))
print(json.dumps({"kind": kind, "train_loss": avg.loss, "val_loss": val.get_average().loss, "moved": len(moved),
                  "logits_rel_l2_vs_reference_classes": err, "wer": avg.metrics["word_error_rate"]}))
