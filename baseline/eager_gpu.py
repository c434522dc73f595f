"""The same-box bar of SURVEY.md §2.1 / §8(d): the reference's training step executed by PyTorch-eager LIBRARY
kernels (cuDNN GRU, cuBLAS GEMMs, ATen conv / LayerNorm / softmax / CTC, transformers' own encoder classes)
on the B200 the product runs on.

BENCH INFRASTRUCTURE ONLY — not part of the product (nothing under wav2vec2forbrain_b200/ imports it) and not
the `--impl reference` arm (which stays the CPU path).  `bench.py` times it next to the product and prints it as
`gpu_eager_baseline`.  /root/reference does not exist on the GPU box, so the reference's thin model wrappers are
restated here with stock torch.nn modules issuing the same library calls, and the encoder arithmetic comes from
the installed `transformers` classes the reference itself instantiates:

  stage 1  smooth -> day affine -> softsign -> unfold        src/model/b2p2t_model.py:138-167
  stage 2  nn.GRU (bidirectional, dropout between layers) + Linear/GELU stack
                                                             src/model/brain_feature_extractor.py:39-67
  stage 3  transformers Wav2Vec2Encoder / Wav2Vec2ConformerEncoder, attention_mask=None, eager attention,
           dropout(final_dropout) + lm_head                  src/model/w2v_custom_feat_extractor.py:144-191,
                                                             src/model/w2v_conformer_custom_feat_extractor.py:62-112
  stage 4  log_softmax + nn.CTCLoss(blank=0, mean, zero_infinity) + ctc_loss.item()
                                                             src/model/w2v_custom_feat_extractor.py:59,81-94
  step     zero_grad -> forward -> backward -> torch.optim.Adam(2 groups, weight_decay, eps).step()
                                                             src/train/train_loop.py:44-78,
                                                             src/experiments/b2t_gru_w2v_experiment.py:109-145
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F
from torch import nn


class EagerBrainEncoder(nn.Module):
    def __init__(self, out_size: int, hidden: int = 512, layers: int = 3, fc_hidden=(256,), sigma: float = 1.5,
                 dropout: float = 0.45, klen: int = 32, stride: int = 4, n_days: int = 24, dim: int = 256):
        super().__init__()
        grid = torch.arange(20, dtype=torch.float32)
        k = torch.exp(-(((grid - 9.5) / sigma) ** 2) / 2) / (sigma * math.sqrt(2 * math.pi))
        self.register_buffer("taps", (k / k.sum()).view(1, 1, -1).repeat(dim, 1, 1))
        self.day_weights = nn.Parameter(torch.eye(dim).repeat(n_days, 1, 1))
        self.day_bias = nn.Parameter(torch.zeros(n_days, 1, dim))
        self.unfold = nn.Unfold((klen, 1), dilation=1, padding=0, stride=stride)
        self.klen, self.stride, self.dim = klen, stride, dim
        self.gru = nn.GRU(dim * klen, hidden, layers, batch_first=True, bidirectional=True, dropout=dropout)
        sizes = [2 * hidden] + list(fc_hidden) + [out_size]
        fc = []
        for i in range(len(sizes) - 1):
            fc.append(nn.Linear(sizes[i], sizes[i + 1]))
            if i < len(sizes) - 2:
                fc.append(nn.GELU())
        self.fc = nn.Sequential(*fc)

    def forward(self, x, day_idxs, input_lens):
        x = F.conv1d(x.permute(0, 2, 1), self.taps, padding="same", groups=self.dim).permute(0, 2, 1)
        w = torch.index_select(self.day_weights, 0, day_idxs)
        x = torch.einsum("btd,bdk->btk", x, w) + torch.index_select(self.day_bias, 0, day_idxs)
        x = F.softsign(x)
        x = self.unfold(x.permute(0, 2, 1).unsqueeze(3)).permute(0, 2, 1)
        lens = ((input_lens - self.klen) / self.stride).to(torch.int32)
        h, _ = self.gru(x)
        return self.fc(h), lens


class EagerModel(nn.Module):
    def __init__(self, workload: str):
        super().__init__()
        if workload == "conformer":
            from transformers import Wav2Vec2ConformerConfig
            from transformers.models.wav2vec2_conformer.modeling_wav2vec2_conformer import Wav2Vec2ConformerEncoder
            cfg = Wav2Vec2ConformerConfig(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16,
                                          intermediate_size=4096, hidden_act="swish", position_embeddings_type="rotary",
                                          rotary_embedding_base=10000, conv_depthwise_kernel_size=31, vocab_size=32,
                                          pad_token_id=0)
            cfg._attn_implementation = "eager"
            self.encoder = Wav2Vec2ConformerEncoder(cfg)
            width = 1024
        else:
            from transformers import Wav2Vec2Config
            from transformers.models.wav2vec2.modeling_wav2vec2 import Wav2Vec2Encoder
            cfg = Wav2Vec2Config(vocab_size=32, pad_token_id=0)      # library defaults == wav2vec2-base-960h architecture
            cfg._attn_implementation = "eager"
            self.encoder = Wav2Vec2Encoder(cfg)
            width = 768
        self.cfg = cfg
        self.brain_encoder = EagerBrainEncoder(width)
        self.dropout = nn.Dropout(cfg.final_dropout)
        self.lm_head = nn.Linear(width, 32)
        self.loss = nn.CTCLoss(blank=0, reduction="mean", zero_infinity=True)

    def w2v_parameters(self):
        return list(self.encoder.parameters()) + list(self.lm_head.parameters())

    def forward(self, x, day_idxs, input_lens, targets, target_lens):
        feats, lens = self.brain_encoder(x, day_idxs, input_lens)
        h = self.encoder(feats, attention_mask=None, output_attentions=False, output_hidden_states=False,
                         return_dict=True)[0]
        logits = self.lm_head(self.dropout(h))
        tg = torch.where(targets < 1, torch.tensor(-100, device=targets.device), targets)
        lp = F.log_softmax(logits.float(), dim=-1).transpose(0, 1)
        loss = self.loss(lp, tg, lens.long(), target_lens.long())
        return logits, loss


def make_step(workload: str, batch: dict, device, autocast_bf16: bool):
    """-> step() running one reference-style training step on `device`; batch = dict of device tensors."""
    torch.manual_seed(42)
    model = EagerModel("conformer" if workload == "conformer" else "base").to(device).train()
    groups = [{"params": list(model.brain_encoder.parameters())}]
    if workload != "frozen":
        groups.append({"params": model.w2v_parameters(), "lr": 1e-5})
    opt = torch.optim.Adam(groups, lr=1e-4, weight_decay=1e-4, eps=1e-8)

    def step():
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast_bf16):
            logits, loss = model(batch["input"], batch["day_idxs"], batch["input_lens"], batch["target"],
                                 batch["target_lens"])
        _ = loss.item()          # w2v_custom_feat_extractor.py:94: the reference syncs on the loss inside forward
        loss.backward()          # frozen: weight gradients of w2v are still computed (train_loop.py:66)
        opt.step()
        return loss
    return step, model


def time_eager(workload: str, batch: dict, device, autocast_bf16: bool, warmup: int = 20, steps: int = 50,
               budget_s: float = 60.0):
    """CUDA-event time of `steps` reference-style steps after `warmup`; the step count shrinks (never below 10 timed
    after 5 warm-up) if one step is so slow that the default bench run would not finish within minutes."""
    import time
    step, model = make_step(workload, batch, device, autocast_bf16)
    t0 = time.perf_counter()
    step()
    torch.cuda.synchronize()
    step()
    torch.cuda.synchronize()
    per = (time.perf_counter() - t0) / 2
    if per * (warmup + steps) > budget_s:
        scale = budget_s / (per * (warmup + steps))
        warmup, steps = max(5, int(warmup * scale)), max(10, int(steps * scale))
    for _ in range(warmup):
        step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    B = batch["input"].shape[0]
    del step, model
    torch.cuda.empty_cache()
    return {"value": B / (ms / 1e3), "unit": "trials/s", "ms_per_step": ms, "warmup": warmup, "steps": steps}
