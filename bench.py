#!/usr/bin/env python
"""Benchmark of the b2p2t_gru+w2v training hot path (BASELINE.json metric: train trials/sec).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload base|frozen|conformer]

One "step" = the reference's per-batch training step (src/train/train_loop.py:44-78):
optimizer.zero_grad -> model.forward(batch) -> loss.backward() -> [grad all-reduce] -> Adam.step,
on one synthetic batch of the shape BASELINE.json names (256-ch features, T=500, 40-char targets).

Default workload (no --workload flag) = the north-star configuration: b2p2t_gru+w2v_conformer, GRU 3x512
bidirectional + FC[256] + wav2vec2-conformer-rope-large architecture (random init), batch 32 per GPU —
BASELINE.json configs[3] at N=1 (the largest single-GPU config) and configs[4] at N>1 (weak scaling, NCCL
gradient all-reduce overlapped with backward).  `--workload base|frozen` select configs[1] / configs[2]; their
lines are kept under profiles/.

`--impl reference` times the reference's own CPU implementation of the same step (the oracle port,
oracle/b2t_oracle.py — the reference is pure PyTorch and /root/reference does not exist on the GPU
box) on the host cores, on a bounded sample (8 trials per step, printed as sample_trials_per_step) of the
same workload.

Besides the contract keys the N=1 line carries `inference` (eval + no_grad forward + greedy decode, trials/s)
and `gpu_eager_baseline` (baseline/eager_gpu.py: the reference's step executed by PyTorch-eager library kernels
on the same GPU, fp32 and bf16-autocast — the same-box bar of SURVEY.md §2.1; not the product, not the
`--impl reference` arm); N>1 lines carry `comm` (exposed gradient all-reduce milliseconds per step).
"""
from __future__ import annotations

import argparse
import datetime
import json
import os
import subprocess
import sys
import tempfile
import time

import torch

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

T_FRAMES, S_TARGET, N_CH = 500, 40, 256
PER_GPU_BATCH = {"base": 64, "frozen": 128, "conformer": 32}      # BASELINE.json configs[1], [2], [3]/[4]
REFERENCE_SAMPLE_BATCH = 8                                        # trials per step of the CPU reference arm

# algorithmic forward GFLOP per trial at T=500; training = 3x forward
GF_TRAIN = {"base": 91.18, "frozen": 69.50, "conformer": 440.66}
GF_INFER = {"base": 30.39, "frozen": 30.39, "conformer": 146.89}


def read_peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops_sustained": p["bf16_tflops_sustained"], "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def make_host_batch(B, seed=1234, pin=False):
    g = torch.Generator().manual_seed(seed)
    b = {
        "input": torch.randn(B, T_FRAMES, N_CH, generator=g),
        "day_idxs": torch.randint(0, 24, (B,), generator=g),
        "target": torch.randint(4, 32, (B, S_TARGET), generator=g),
        "input_lens": torch.full((B,), T_FRAMES, dtype=torch.long),
        "target_lens": torch.full((B,), S_TARGET, dtype=torch.long),
    }
    if pin:
        b = {k: v.pin_memory() for k, v in b.items()}
    return b


# ---------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle port on the host cores
# ---------------------------------------------------------------------------------------------
def oracle_step_fn(workload: str, B: int):
    from oracle import b2t_oracle as O
    be = O.BrainEncoderCfg(out_size=1024 if workload == "conformer" else 768, dropout=0.45)
    w2v = O.conformer_large_cfg() if workload == "conformer" else O.base_cfg()
    w2v.hidden_dropout = w2v.attention_dropout = w2v.activation_dropout = w2v.final_dropout = 0.1
    sd = O.make_state_dict(be, w2v, seed=42)
    params = {k: (v.clone().requires_grad_(not O.is_buffer(k)) if v.is_floating_point() else v) for k, v in sd.items()}
    brain = [v for k, v in params.items() if k.startswith("brain_encoder.") and v.requires_grad]
    w2vp = [v for k, v in params.items() if k.startswith("w2v_encoder.") and v.requires_grad]
    groups = [{"params": brain, "lr": 1e-4}]
    if workload != "frozen":
        groups.append({"params": w2vp, "lr": 1e-5})
    opt = torch.optim.Adam(groups, weight_decay=1e-4, eps=1e-8)   # base_args.py:113,121; experiment eps
    batch = make_host_batch(B)

    def step():
        opt.zero_grad()
        out = O.forward(batch, params, be, w2v, training=True, gru_impl="aten")
        out.loss.backward()
        opt.step()
        return out.loss.item()
    return step


def time_oracle(workload: str, steps: int, warmup: int, B: int = 8):
    torch.set_num_threads(os.cpu_count() or 1)
    step = oracle_step_fn(workload, B)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    return B * steps / dt, dt / steps * 1e3


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = REFERENCE_SAMPLE_BATCH
    steps = max(1, min(args.steps, 20))
    warmup = max(1, min(args.warmup, 3))
    tps, ms = time_oracle(args.workload, steps, warmup, B)
    cores = os.cpu_count() or 1
    # `config` names the workload of our arm (the contract: same config / metric / unit); what one step of THIS arm
    # really ran is a bounded sample of that workload's per-GPU batch, stated in `sample_trials_per_step`,
    # `ms_per_step` (of that sample) and cpu_baseline.sample (CPU trials/s is batch-independent to a few %)
    cfg = workload_config(args.workload, args.gpus, PER_GPU_BATCH[args.workload])
    line = {
        "impl": "reference", "metric": "train trials/sec", "value": tps, "unit": "trials/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg, "sample_trials_per_step": B,
        "cpu_baseline": {"value": tps, "unit": "trials/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} steps of {B} trials (T=500) of the same workload after {warmup} warm-up, "
                                   "torch CPU fp32, all host threads"},
        "e2e": {"value": tps, "unit": "trials/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def workload_config(workload: str, n_gpus: int, per_gpu_batch: int):
    names = {
        "base": "b2p2t_gru+w2v full fine-tuning, wav2vec2-base arch (BASELINE.json configs[1])",
        "frozen": "b2p2t_gru+w2v frozen_w2v (BASELINE.json configs[2])",
        "conformer": "b2p2t_gru+w2v_conformer, wav2vec2-conformer-rope-large arch, batch 32/GPU (BASELINE.json "
                     + ("configs[3])" if n_gpus == 1 else "configs[4], data parallel)"),
    }
    return {"workload": names[workload], "per_gpu_batch": per_gpu_batch, "global_batch": per_gpu_batch * n_gpus,
            "T": T_FRAMES, "channels": N_CH, "target_len": S_TARGET, "gru": "3x512 bidirectional", "fc": [256],
            "parallelism": f"dp{n_gpus}", "l2": "per-step working set (weights+activations, several GB) exceeds the 126 MB L2"}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def build_model(workload: str, precision: str, optimizer: str = "fused"):
    from wav2vec2forbrain_b200 import models
    a = models.BrainEncoderArgs(encoder_gru_hidden_size=512, encoder_num_gru_layers=3, encoder_fc_hidden_sizes=(256,),
                                gaussian_smooth_width=1.5, encoder_dropout=0.45)
    torch.manual_seed(42)
    if workload == "conformer":
        ckpt = "facebook/wav2vec2-conformer-rope-large-960h-ft"
        brain = models.bfe_w_preprocessing_from_config(a, None, ckpt, precision=precision)
        model = models.W2VConformerBrainEncoderModel(brain, ckpt, precision=precision)
    else:
        ckpt = "facebook/wav2vec2-base-960h"
        brain = models.bfe_w_preprocessing_from_config(a, None, ckpt, precision=precision)
        model = models.W2VBrainEncoderModel(a, brain, ckpt, None, True, precision=precision)
    model = model.cuda()
    model.train()
    brain_p = list(model.brain_encoder.parameters())
    groups = [{"params": brain_p, "lr": 1e-4}]
    if workload != "frozen":
        groups.append({"params": list(model.w2v_encoder.parameters()), "lr": 1e-5})   # b2t_gru_w2v_experiment.py:111-145
    else:
        # unfreeze_strategy == "brain_encoder": nobody reads the w2v weight gradients -> dgrad only
        model.runtime.skip_w2v_wgrad = True
    # weight_decay / eps as create_optimizer passes them (b2t_gru_w2v_experiment.py:138-145, base_args.py:113,121)
    if optimizer == "torch":
        opt = torch.optim.Adam(groups, weight_decay=1e-4, eps=1e-8, fused=True)
    else:
        from wav2vec2forbrain_b200 import optim as boptim
        opt = boptim.FusedAdam(groups, weight_decay=1e-4, eps=1e-8)
    return model, opt


def to_device_batch(models, hb, device):
    b = models.B2tSampleBatch(hb["input"].to(device, non_blocking=True), hb["target"].to(device, non_blocking=True))
    b.day_idxs = hb["day_idxs"].to(device, non_blocking=True)
    b.input_lens = hb["input_lens"].to(device, non_blocking=True)
    b.target_lens = hb["target_lens"].to(device, non_blocking=True)
    return b


class NvmlSampler:
    """In-process NVML (what nvidia-smi reads) from a thread, two queries every 250 ms while the timed region runs.
    Preferred over the nvidia-smi process below: on some boxes each poll of the 9-field nvidia-smi query stalls kernel
    launches for 25-50 ms (conformer step 58.7 -> 66-77 ms with 6-7 polls inside the timed region; the end-to-end loop,
    timed without a sampler, was unaffected)."""

    def __init__(self, dev_index: int, period: float = 0.25):
        import threading
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        try:
            self.h = pynvml.nvmlDeviceGetHandleByUUID("GPU-" + str(torch.cuda.get_device_properties(dev_index).uuid))
        except Exception:
            self.h = pynvml.nvmlDeviceGetHandleByIndex(dev_index)
        self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        self._sample()            # first call of each query outside the timed region
        self.rows = []
        self.on = False
        self.period = period
        self.ev = threading.Event()
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _sample(self):
        return (float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)),
                int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)))

    def _run(self):
        while not self.ev.wait(self.period):
            if self.on:
                try:
                    self.rows.append(self._sample())
                except Exception:
                    pass

    def wait_ready(self):
        pass

    def mark(self):
        self.on = True

    def stop(self):
        self.on = False
        self.ev.set()
        self.t.join(timeout=2)
        out = {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "source": "nvml"}
        rows = self.rows
        if rows:
            sm = sorted(r[0] for r in rows)
            out["sm_mhz"] = sm[len(sm) // 2]
            bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
            for nm, bit in bits.items():
                if any(r[1] & bit for r in rows):
                    out["reasons"].append(nm)
        out["samples"] = len(rows)
        return out


def make_sampler(dev_index: int):
    if os.environ.get("B2T_BENCH_NO_SAMPLER") == "1":
        return None
    try:
        return NvmlSampler(dev_index)
    except Exception:
        return ClockSampler(dev_index, period_ms=1000)


class ClockSampler:
    """One nvidia-smi process polling every 200 ms.  It is started (and its driver initialisation waited for) BEFORE
    the warm-up steps: attaching to the driver takes 0.1-0.4 s on some boxes and stalls kernel launches while it does,
    which used to land inside the first timed steps.  mark() is called when the timed region starts; only the samples
    after it are reported."""
    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, period_ms: int = 200):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.t_mark = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", str(period_ms),
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def wait_ready(self, timeout: float = 10.0):
        t0 = time.time()
        while self.p is not None and time.time() - t0 < timeout:
            if os.path.getsize(self.f.name) > 0 or self.p.poll() is not None:
                return
            time.sleep(0.02)

    def mark(self):
        self.t_mark = datetime.datetime.now()

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = []
        with open(self.f.name) as fh:
            for line in fh:
                parts = [x.strip() for x in line.split(",")]
                if len(parts) >= 10:
                    rows.append(parts)
        os.unlink(self.f.name)
        if self.t_mark is not None:
            timed = []
            for r in rows:
                try:
                    if datetime.datetime.strptime(r[0], "%Y/%m/%d %H:%M:%S.%f") >= self.t_mark:
                        timed.append(r)
                except ValueError:
                    timed.append(r)
            rows = timed or rows[-1:]
        rows = [r[1:] for r in rows]
        if not rows:
            return out
        sm = sorted(float(r[1]) for r in rows if r[1].replace(".", "").isdigit())
        if sm:
            out["sm_mhz"] = sm[len(sm) // 2]
        try:
            out["sm_max_mhz"] = float(rows[0][2])
        except ValueError:
            pass
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for i, nm in enumerate(names):
            if any(r[5 + i].lower().startswith("active") for r in rows):
                out["reasons"].append(nm)
        out["samples"] = len(rows)
        return out


def run_ours(args):
    import torch.distributed as dist
    from wav2vec2forbrain_b200 import dist as bdist, models, ops
    rank, local_rank, world = bdist.env_rank_world()
    multi = world > 1
    if multi:
        bdist.init_process_group("nccl")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    workload, precision = args.workload, args.precision
    B = args.batch or PER_GPU_BATCH[workload]
    model, opt = build_model(workload, precision, args.optimizer)
    dp = None
    if multi:
        dp = bdist.DataParallel(model)
        # no gradient clipping in this loop (the reference's default): the optimizer may wait per parameter group
        dp.attach_optimizer(opt, lazy=os.environ.get("B2T_DP_LAZY", "1") == "1")
    hb = make_host_batch(B, seed=1234 + rank, pin=True)
    batch = to_device_batch(models, hb, dev)
    torch.cuda.synchronize()

    def step(bt):
        opt.zero_grad(set_to_none=True)
        out = model.forward(bt)
        out.loss.backward()
        opt.step()          # DP: pre-hook waits for the bucketed all-reduce
        return out

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    bdist.tune_gc()      # (also what a training script would call once after building model and optimizer)
    sampler = make_sampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.wait_ready()
    for _ in range(max(args.warmup, 3)):
        step(batch)
    # Untimed settling steps: LayerDrop changes the allocation sequence from step to step, so the caching allocator keeps
    # growing its pool (cudaMalloc, which synchronises the device) for a few more steps than any fixed warm-up count;
    # measured: 44.7 -> 49.1 GiB reserved over ~25 conformer steps, 59 ms instead of 52 ms per step while it grows.
    # Warm up until the reserved pool is unchanged over 4 consecutive steps (at most 40 extra steps).
    same, last, settle = 0, torch.cuda.memory_reserved(dev), 0
    while settle < 40:
        more = same < 4
        if multi:      # every rank runs the same number of steps (each step holds collectives)
            flag = torch.tensor([1.0 if more else 0.0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            more = flag.item() > 0
        if not more:
            break
        step(batch)
        settle += 1
        now = torch.cuda.memory_reserved(dev)
        same = same + 1 if now == last else 0
        last = now
    # ---- device-resident timing ----
    barrier()
    if sampler:
        sampler.mark()
    l0 = ops.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    marks = []
    for _ in range(args.steps):
        step(batch)
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append(ev)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # per-step device times of this rank (diagnostic: a few slow steps or all of them?)
    per = sorted(a.elapsed_time(b) for a, b in zip([e0] + marks[:-1], marks))
    step_ms = {"min": per[0], "median": per[len(per) // 2], "max": per[-1]}
    launches = ops.LAUNCHES - l0
    clocks = sampler.stop() if sampler else None
    # ---- end-to-end timing: pinned host batch -> device every step, loss read back every step ----
    barrier()
    t0 = time.perf_counter()
    # the device batch is a persistent staging buffer refilled from pinned host memory every step (a fresh
    # allocation per step while the previous step's outputs are still referenced changes the caching
    # allocator's pattern and costs occasional cudaMalloc/cudaFree stalls)
    keys = ("input", "target", "day_idxs", "input_lens", "target_lens")
    dst = {"input": batch.input, "target": batch.target, "day_idxs": batch.day_idxs, "input_lens": batch.input_lens,
           "target_lens": batch.target_lens}
    out = None
    for _ in range(args.steps):
        for k in keys:
            dst[k].copy_(hb[k], non_blocking=True)
        out = step(batch)
        _ = out.loss.item()
        out = None
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    h2d = sum(hb[k].numel() * hb[k].element_size() for k in keys)
    t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    if multi:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = t.tolist()
    peaks = read_peaks()
    # ---- exposed gradient all-reduce per step (N > 1): the same timed loop with the collective switched off ----
    comm = None
    if multi:
        dp.buckets.enabled = False
        for _ in range(2):
            step(batch)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(args.steps):
            step(batch)
        c1.record()
        barrier()
        dp.buckets.enabled = True
        tt = torch.tensor([c0.elapsed_time(c1)], device=dev, dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        nocomm_ms = tt.item() / args.steps
        for p in model.parameters():        # the un-reduced steps let the replicas drift: re-broadcast rank 0's weights
            dist.broadcast(p.data, src=0)
        comm = {"exposed_ms_per_step": ms / args.steps - nocomm_ms, "ms_per_step_without_allreduce": nocomm_ms,
                "allreduce_bytes_per_step": dp.buckets.bytes_per_step, "buckets": len(dp.buckets.buckets),
                "how": "same timed loop with the gradient all-reduce disabled, max over ranks"}
    # every rank runs the roofline steps (they contain the gradient all-reduce); rank 0 reports
    roof = gemm_roofline(model, opt, batch, step, peaks, workload) if not args.no_roofline else None
    # ---- inference: eval() + no_grad forward + greedy CTC decode on the device (train_loop.py:89-109, evaluator.py:108-120)
    infer = None
    if not args.no_inference and rank == 0:
        model.eval()
        with torch.no_grad():
            for _ in range(3):
                ops.greedy_decode(model.forward(batch).logits)
            i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            i0.record()
            for _ in range(args.steps):
                ops.greedy_decode(model.forward(batch).logits)
            i1.record()
            torch.cuda.synchronize()
        model.train()
        ims = i0.elapsed_time(i1) / args.steps
        infer = {"value": B / (ims / 1e3), "unit": "trials/s", "ms_per_batch": ims, "per_gpu_batch": B, "n_gpus": 1,
                 "what": "eval() + no_grad forward + greedy CTC decode kernel, device-resident batch, CUDA events",
                 "roofline_frac": B / (ims / 1e3) * GF_INFER[workload] / (peaks["bf16_tflops_sustained"] * 1e3)}
    if multi:
        dist.barrier()
        torch.cuda.synchronize()
        dist.destroy_process_group()
    if rank != 0:
        return
    tps = B * world * args.steps / (ms / 1e3)
    e2e_tps = B * world * args.steps / (e2e_ms / 1e3)
    line = {
        "metric": "train trials/sec", "value": tps, "unit": "trials/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "settling_steps": settle, "ms_per_step": ms / args.steps, "step_ms_rank0": step_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16x3 (3-term bf16 split operands, fp32 accumulate)" if precision == "bf16x3" else "bf16",
        "data": "synthetic", "config": workload_config(workload, world, B),
        "e2e": {"value": e2e_tps, "unit": "trials/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
        "gpu_launches": launches, "clocks": clocks,
        "step_roofline_frac": tps / world * GF_TRAIN[workload] / (peaks["bf16_tflops_sustained"] * 1e3),
    }
    if roof is not None:
        line["roofline"] = roof
    if infer is not None:
        line["inference"] = infer
    if comm is not None:
        line["comm"] = comm
    line["optimizer"] = ("wav2vec2forbrain_b200.optim.FusedAdam (multi-tensor CUDA kernel)" if args.optimizer == "fused"
                         else "torch.optim.Adam(fused=True)")
    if world == 1 and not args.no_eager:
        # the same-box bar: the reference's step by PyTorch-eager library kernels on this GPU
        del model, opt
        torch.cuda.empty_cache()
        from baseline import eager_gpu
        dbatch = {k: v.to(dev) for k, v in hb.items()}
        eager = {"what": "reference training step executed by PyTorch-eager library kernels (cuDNN GRU, cuBLAS, ATen, "
                         "transformers encoder classes; baseline/eager_gpu.py) on the same GPU, same workload and batch; "
                         "NOT the product and NOT the --impl reference arm", "per_gpu_batch": B}
        for key, ac in (("fp32", False), ("bf16_autocast", True)):
            try:
                eager[key] = eager_gpu.time_eager(workload, dbatch, dev, ac)
                eager[key]["ours_over_eager"] = tps / eager[key]["value"]
            except Exception as ex:       # never lose the product's line to the baseline
                eager[key] = {"error": f"{type(ex).__name__}: {ex}"[:300]}
        eager["fp32"]["note"] = "torch defaults, as the reference runs: fp32 matmul (TF32 off), cuDNN may use TF32"
        line["gpu_eager_baseline"] = eager
    if world == 1 and not args.no_cpu_baseline:
        v, _ = time_oracle(workload, steps=3, warmup=1, B=REFERENCE_SAMPLE_BATCH)
        line["cpu_baseline"] = {"value": v, "unit": "trials/s", "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": f"3 steps of {REFERENCE_SAMPLE_BATCH} trials (T=500) of the same workload after 1 "
                                          "warm-up, torch CPU fp32, all host threads"}
    print(json.dumps(line))


def gemm_roofline(model, opt, batch, step, peaks, workload_key="base"):
    """CUDA-event time of every tcgen05 GEMM launch of one training step vs its algorithmic FLOPs."""
    from wav2vec2forbrain_b200 import engine, ops
    real = ops.gemm
    rec = []

    def timed_gemm(a, b, **kw):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        real(a, b, **kw)
        e.record()
        k_eff = kw["k"] * (a.d2 if kw.get("k_spans_batches") else 1)
        rec.append((s, e, 2.0 * kw["m_per_batch"] * kw.get("batches", 1) * kw["n"] * k_eff,
                    (kw["m_per_batch"], kw.get("batches", 1), kw["n"], k_eff, "A" + ("mn" if a.mn_major else "k"),
                     "B" + ("mn" if b.mn_major else "k"), "sk%d" % kw.get("split_k", 1), "bn%d" % kw.get("block_n", 0),
                     "+".join(x for x in ("out_f32", "out_pre", "out_hi", "bias", "residual", "aux", "colsum", "act") if kw.get(x) is not None))))
    ops.gemm = timed_gemm
    # every launch is timed on the stream it runs on, one at a time: the weight-gradient side stream (which overlaps
    # GEMMs of two streams in the product's step) is switched off for this measuring step
    side_was = engine.Engine._side_enabled
    engine.Engine._side_enabled = False
    try:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        step(batch)
        rec.clear()
        e0.record()
        step(batch)
        e1.record()
        torch.cuda.synchronize()
    finally:
        ops.gemm = real
        engine.Engine._side_enabled = side_was
    t_ms = sum(r[0].elapsed_time(r[1]) for r in rec)
    flops = sum(r[2] for r in rec)
    table = os.environ.get("B2T_GEMM_TABLE")
    if table:   # per-shape breakdown of the step's GEMM launches (debug aid, not part of the JSON line)
        agg = {}
        for s, e, f, key in rec:
            c = agg.setdefault(key, [0, 0.0, 0.0])
            c[0] += 1; c[1] += s.elapsed_time(e) * 1e3; c[2] += f
        with open(table, "w") as fh:
            for key, (n, us, f) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                fh.write(f"{n:4d} x {us / n:8.1f} us = {us:9.1f} us  {f / us / 1e6:7.1f} TF/s alg  {key}\n")
    step_ms = e0.elapsed_time(e1)
    achieved = flops / (t_ms * 1e-3) / 1e12
    peak = peaks["bf16_tflops_sustained"]
    traffic = None   # DRAM bytes per GEMM launch from the committed ncu capture (scripts/gpu_traffic.sh -> scripts/traffic_json.py)
    tpath = os.path.join(REPO, "profiles", "gemm_traffic_r02.json")
    if not os.path.exists(tpath):
        tpath = os.path.join(REPO, "profiles", "gemm_traffic_r01.json")
    if os.path.exists(tpath):
        with open(tpath) as fh:
            tj = json.load(fh).get(workload_key)
        if tj:
            traffic = tj["dram_bytes_per_launch"]
    return {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05/TMA, all launches of one step)", "achieved": achieved,
            "achieved_mma_tflops": achieved * (3.0 if getattr(getattr(getattr(model, "runtime", None), "engine", None), "three", True) else 1.0),
            "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
            "peak_source": peaks["source"] + " (sustained cuBLAS bf16)", "launches": len(rec), "gemm_ms_per_step": t_ms,
            "gemm_share_of_step": t_ms / step_ms,
            "note": "algorithmic FLOPs (2MNK); the bf16x3 mode issues 3 MMAs per algorithmic MAC"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=["base", "frozen", "conformer"])
    ap.add_argument("--precision", default="bf16x3", choices=["bf16x3", "bf16"])
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-eager", action="store_true", help="skip the PyTorch-eager same-box baseline")
    ap.add_argument("--no-inference", action="store_true")
    ap.add_argument("--optimizer", default="fused", choices=["fused", "torch"],
                    help="fused = this repo's multi-tensor Adam kernel; torch = torch.optim.Adam(fused=True)")
    args = ap.parse_args()
    if args.workload is None:
        args.workload = "conformer"     # north-star config: configs[3] at N=1, configs[4] at N>1
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
