"""Forward / backward schedule of the b2p2t_gru+w2v hot path over the C-ABI kernels.

This is the host-side replacement of W2VBrainEncoderModel.forward + loss.backward()
(src/model/w2v_custom_feat_extractor.py:65-123, src/train/train_loop.py:67-70): one explicit
kernel schedule with hand-derived backward, instead of ~2000 autograd nodes of library calls.
All activations are fp32 master copies plus bf16 (hi[, lo]) operand copies written by the
producing kernel's epilogue; weights are split once per optimizer step.

`ops` is looked up through the module global so the CPU test-suite can substitute an emulator of
the C ABI (tests/abi_emulator.py) and check this schedule against the oracle without a GPU; the
product never does.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import os

import torch

from . import ops as _real_ops
from .config import BrainEncoderConfig, W2VConfig, NEURAL_DIM

ops = _real_ops
Tensor = torch.Tensor


def _mix(seed: int, site: int) -> int:
    """per-site dropout seed (splitmix-style, host side)."""
    z = (seed * 0x9E3779B97F4A7C15 + site * 0xD1B54A32D192ED03 + 0x632BE59BD9B4E019) & 0xFFFFFFFFFFFFFFFF
    z ^= z >> 31
    return (z * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF


class Weights:
    """fp32 parameters + cached bf16 operand copies, keyed by parameter version."""

    def __init__(self, three: bool):
        self.three = three
        self._cache: Dict[str, tuple] = {}

    def get(self, key: str, versions: tuple, build):
        hit = self._cache.get(key)
        if hit is not None and hit[0] == versions:
            return hit[1]
        val = build()
        self._cache[key] = (versions, val)
        return val


def _ver(*ts) -> tuple:
    return tuple((t.data_ptr(), t._version) for t in ts if t is not None)


def _split_k_for(m: int, n: int, k_blocks: int, bn: int = 128) -> int:
    """split-K factor of a weight-gradient GEMM ([m, n] output, k_blocks 64-row blocks of contraction): enough work
    items for the 74 CTA pairs of the 256 x 256 pair tiles (b2t_gemm cta_group 2), 148 CTAs for narrow outputs."""
    if n % 128 == 0 and n >= 256 and m > 128:
        tiles = ((m + 255) // 256) * ((n + 255) // 256)
        if tiles >= 56:
            return 1
        return max(1, min(74 // tiles, max(1, k_blocks // 4), 16))
    tiles = ((m + 127) // 128) * ((n + bn - 1) // bn)
    if tiles >= 96:
        return 1
    return max(1, min(148 // tiles, max(1, k_blocks // 4), 16))


class Engine:
    def __init__(self, be: BrainEncoderConfig, w2v: W2VConfig, precision: str = "bf16x3"):
        assert precision in ("bf16x3", "bf16")
        self.be, self.w2v = be, w2v
        self.three = precision == "bf16x3"
        self.precision = precision
        self.weights = Weights(self.three)
        if be.nonlinearity != "softsign":
            raise NotImplementedError("input_layer_nonlinearity other than softsign")  # b2p2t_model.py:107
        if be.fc_act not in ("gelu", "silu", "swish"):
            raise NotImplementedError(f"encoder_fc_activation_function={be.fc_act}")

    # ------------------------------------------------------------------------------------------
    # gradient sink: data parallel training hands finished gradients to the all-reduce while the backward
    # schedule of the same autograd node is still being issued (dist.GradientBuckets.sink); None = single GPU
    # ------------------------------------------------------------------------------------------
    grad_sink = None
    grad_sync = None

    def begin_grads(self) -> None:
        self._emitted = set()

    def emit_grads(self, G: Dict[str, Tensor]) -> None:
        """Everything in G that was not handed over yet is final (a layer only writes its own keys)."""
        if self.grad_sink is None:
            return
        done = getattr(self, "_emitted", None)
        if done is None:
            done = self._emitted = set()
        new = [k for k, v in G.items() if k not in done and v is not None]
        if new:
            if not self._arena_on:
                raise RuntimeError("data-parallel gradient buckets reduce the flat gradient arena in place, but this backward "
                                   "pass is not writing into it (parameters still hold a .grad from an earlier backward: "
                                   "gradient accumulation is not supported under DataParallel; call optimizer.zero_grad())")
            self.finalize_grads(G, only=new)       # a gradient is reduced where it lives: in its arena slot
            done.update(new)
            self.grad_sink([(k, G[k]) for k in new])

    def sync_grads(self, G: Dict[str, Tensor], prefix: str = "") -> None:
        """End of an autograd node: the rest is handed over, and every bucket of this node that has not left yet is
        launched (dist.ArenaBuckets.node_end)."""
        if self.grad_sink is None:
            return
        self.emit_grads(G)
        self.grad_sync(prefix)
        self._emitted = set()

    # ------------------------------------------------------------------------------------------
    # helpers
    # ------------------------------------------------------------------------------------------
    def _w(self, P, name: str):
        """bf16 operand copy of a [N, K] weight (serves forward as K-major and dgrad as MN-major)."""
        w = P[name]
        return self.weights.get(name, _ver(w), lambda: ops.split(w, self.three))

    def _w_cat(self, P, key: str, names: List[str], unfold=None):
        """several [N_i, K] weights stacked along N into one operand (QKV, both GRU directions)."""
        ws = [P[n] for n in names]

        def build():
            n_tot = sum(w.shape[0] for w in ws)
            k = ws[0].shape[1]
            hi = torch.empty(n_tot, k, dtype=torch.bfloat16, device=ws[0].device)
            lo = torch.empty_like(hi) if self.three else None
            r = 0
            for w in ws:
                if unfold is None:
                    h, l = ops.split(w, self.three)
                else:
                    h, l = ops.permute_unfold_cols(w, unfold[0], unfold[1], self.three)
                hi[r:r + w.shape[0]].copy_(h)
                if lo is not None:
                    lo[r:r + w.shape[0]].copy_(l)
                r += w.shape[0]
            return hi, lo
        return self.weights.get(key, _ver(*ws), build)

    @staticmethod
    def _cat_vec(P, names: List[str]) -> Optional[Tensor]:
        vs = [P.get(n) for n in names]
        if any(v is None for v in vs):
            return None
        return torch.cat([v.reshape(-1) for v in vs])

    def _linear_fwd(self, a_hi, a_lo, w_hl, bias, M, N, K, **kw):
        """[M,K] x [N,K]^T with epilogue; outputs selected through kw (want_f32 / want_split / want_pre)."""
        dev = a_hi.device
        want_f32 = kw.pop("want_f32", True)
        want_split = kw.pop("want_split", False)
        want_pre = kw.pop("want_pre", False)
        out = torch.empty(M, N, dtype=torch.float32, device=dev) if want_f32 else None
        pre = torch.empty(M, N, dtype=torch.float32, device=dev) if want_pre else None
        hi = torch.empty(M, N, dtype=torch.bfloat16, device=dev) if want_split else None
        lo = torch.empty(M, N, dtype=torch.bfloat16, device=dev) if (want_split and self.three) else None
        ops.gemm(ops.kmajor(a_hi, a_lo), ops.kmajor(w_hl[0], w_hl[1]), m_per_batch=M, n=N, k=K, out_f32=out, out_pre=pre,
                 out_hi=hi, out_lo=lo, bias=bias, **kw)
        return out, pre, hi, lo

    # ------------------------------------------------------------------------------------------
    # gradient outputs: slots of the flat arena (flat.GradArena) when one is attached and the node's parameters hold
    # no .grad yet (models sets _arena_on per autograd node), else fresh tensors exactly as before
    # ------------------------------------------------------------------------------------------
    grad_arena = None
    _arena_on = False
    arena_copies = 0        # gradients that had to be copied into their slot at the end of a node (diagnostic)

    def _gout(self, name: str, shape, device, zero: bool) -> Tensor:
        """Tensor the gradient of parameter `name` is written to.  Arena slots are zero at the start of the node."""
        if self._arena_on:
            v = self.grad_arena.view(name)
            if v is not None:
                return v if tuple(v.shape) == tuple(shape) else v.view(shape)
        if zero:
            return ops.zeros(tuple(shape), device=device)
        return torch.empty(tuple(shape), dtype=torch.float32, device=device)

    def _gout_cat(self, names, shape, device, zero: bool) -> Tensor:
        """One tensor covering the gradients of several parameters produced by one kernel (row-stacked)."""
        if self._arena_on:
            v = self.grad_arena.view_cat(names, shape)
            if v is not None:
                return v
        if zero:
            return ops.zeros(tuple(shape), device=device)
        return torch.empty(tuple(shape), dtype=torch.float32, device=device)

    def finalize_grads(self, G: Dict[str, Tensor], only=None) -> None:
        """End of an autograd node: every gradient that was not produced in its arena slot is copied there, so that
        .grad tensors, all-reduce buckets and optimizer tables always see the slot."""
        if only is None:
            self.join_side()      # end of the node: autograd and the optimizer read the gradients on this stream
        if not self._arena_on:
            return
        for k in (list(G) if only is None else only):
            v = G[k]
            if v is None:
                continue
            slot = self.grad_arena.view(k)
            if slot is None or self.grad_arena.owns(k, v):
                continue
            self.join_side()      # a gradient produced outside its slot may have come from the weight-gradient stream
            slot.copy_(v.view(slot.shape) if v.is_contiguous() else v.reshape(slot.shape))
            G[k] = slot
            self.arena_copies += 1

    @staticmethod
    def grad_groups(be: BrainEncoderConfig, w2v: W2VConfig):
        """Parameter names whose gradients one kernel writes as one row-stacked tensor (arena layout constraint)."""
        groups = []
        e = "brain_encoder.neural_decoder.encoder."
        if be.bidirectional:
            for l in range(be.gru_layers):
                groups.append([e + f"gru.weight_ih_l{l}", e + f"gru.weight_ih_l{l}_reverse"])
                groups.append([e + f"gru.bias_ih_l{l}", e + f"gru.bias_ih_l{l}_reverse"])
        for i in range(w2v.layers):
            if w2v.kind == "base":
                a = f"w2v_encoder.wav2vec2.encoder.layers.{i}.attention."
                groups.append([a + "q_proj.weight", a + "k_proj.weight", a + "v_proj.weight"])
                groups.append([a + "q_proj.bias", a + "k_proj.bias", a + "v_proj.bias"])
            else:
                a = f"w2v_encoder.wav2vec2_conformer.encoder.layers.{i}.self_attn."
                groups.append([a + "linear_q.weight", a + "linear_k.weight"])
                groups.append([a + "linear_q.bias", a + "linear_k.bias", a + "linear_v.bias"])
        return groups

    # frozen-w2v mode (unfreeze_strategy == "brain_encoder", b2t_gru_w2v_experiment.py:129-133): the optimizer never
    # reads the w2v weight gradients, so the encoder backward is data-gradient only; set by
    # models._EncoderFn.backward around the encoder's schedule, never during the brain encoder's
    _skip_wgrad = False

    def _wgrad(self, dy_hi, dy_lo, x_hi, x_lo, N, K, M, out: Optional[Tensor] = None, alpha: float = 1.0,
               name=None) -> Optional[Tensor]:
        """dW[N,K] = dY[M,N]^T . X[M,K]  (both operands row-major, contracted over rows).  `name`: the parameter
        (or the list of row-stacked parameters) whose gradient this is -> written into its arena slot."""
        if self._skip_wgrad:
            return None
        kb = (M + 63) // 64
        sk = _split_k_for(N, K, kb)
        if out is None:
            if isinstance(name, (list, tuple)):
                out = self._gout_cat(name, (N, K), dy_hi.device, zero=sk > 1)
            elif name is not None:
                out = self._gout(name, (N, K), dy_hi.device, zero=sk > 1)
            else:
                out = ops.zeros(N, K, device=dy_hi.device) if sk > 1 else torch.empty(N, K, dtype=torch.float32, device=dy_hi.device)
        side = self._side()
        if side is None:
            ops.gemm(ops.mnmajor(dy_hi, dy_lo), ops.mnmajor(x_hi, x_lo), m_per_batch=N, n=K, k=M, out_f32=out,
                     atomic=sk > 1, split_k=sk, alpha=alpha)
            return out
        # Weight gradients are off the backward pass's critical path (nothing reads them before the node ends): they go
        # to a second stream, so that their CTAs fill the SMs the data-gradient chain leaves idle in the ramp-up and
        # the last-tile tail of every persistent GEMM.  The side stream waits for everything queued so far (the
        # producer of dY, the zeroing of the slot); operands are kept from reuse until the side stream has read them.
        side.wait_stream(self._main_stream())
        ops.gemm(ops.mnmajor(dy_hi, dy_lo), ops.mnmajor(x_hi, x_lo), m_per_batch=N, n=K, k=M, out_f32=out,
                 atomic=sk > 1, split_k=sk, alpha=alpha, on_stream=side.cuda_stream)
        for t in ((dy_hi, dy_lo, x_hi, x_lo) if self._arena_on else (dy_hi, dy_lo, x_hi, x_lo, out)):   # arena slots persist
            if t is not None:
                t.record_stream(side)
        self._side_dirty = True
        return out

    # ------------------------------------------------------------------------------------------
    # weight-gradient stream (B2T_WGRAD_STREAM=0 switches it off)
    # ------------------------------------------------------------------------------------------
    side_stream = None
    _side_dirty = False
    _side_enabled = os.environ.get("B2T_WGRAD_STREAM", "1") == "1"

    def _side(self):
        if not self._side_enabled or ops is not _real_ops or not torch.cuda.is_available():
            return None     # the CPU test-suite runs the schedule on an emulator of the C ABI: one stream
        if self.side_stream is None:
            self.side_stream = torch.cuda.Stream()
            _real_ops.SIDE_STREAMS.append(self.side_stream)
        return self.side_stream

    _main = None

    def _main_stream(self):
        """torch.cuda.current_stream() costs ~18 us of Python per call; the stream object is cached per raw handle"""
        raw = ops._lib.raw_stream()
        m = self._main
        if m is None or m[0] != raw:
            m = self._main = (raw, torch.cuda.current_stream())
        return m[1]

    def join_side(self) -> None:
        """the current stream waits for every weight gradient issued so far"""
        if self._side_dirty:
            self._main_stream().wait_stream(self.side_stream)
            self._side_dirty = False

    def _dgrad(self, dy_hi, dy_lo, w_hl, M, N, K, **kw):
        """dX[M,K] = dY[M,N] . W[N,K]"""
        dev = dy_hi.device
        want_f32 = kw.pop("want_f32", True)
        want_split = kw.pop("want_split", False)
        out = torch.empty(M, K, dtype=torch.float32, device=dev) if want_f32 else None
        hi = torch.empty(M, K, dtype=torch.bfloat16, device=dev) if want_split else None
        lo = torch.empty(M, K, dtype=torch.bfloat16, device=dev) if (want_split and self.three) else None
        ops.gemm(ops.kmajor(dy_hi, dy_lo), ops.mnmajor(w_hl[0], w_hl[1]), m_per_batch=M, n=K, k=N, out_f32=out, out_hi=hi,
                 out_lo=lo, **kw)
        return out, hi, lo

    # ------------------------------------------------------------------------------------------
    # brain encoder   (b2p2t_model.py:138-183, brain_feature_extractor.py:56-68)
    # ------------------------------------------------------------------------------------------
    def brain_forward(self, P, x: Tensor, day: Tensor, training: bool, seed: int, aug=(0.0, 0.0)):
        be = self.be
        three = self.three
        B, T, C = x.shape
        assert C == NEURAL_DIM
        S: dict = {"B": B, "T": T}
        pre = "brain_encoder."
        taps = P[pre + "gaussian_smoother.weight"]
        ntaps = taps.shape[-1]
        _, y_hi, y_lo = ops.smooth_fwd(x, taps, pad_left=(ntaps - 1) // 2, three=three, white_sd=aug[0], offset_sd=aug[1],
                                       seed=_mix(seed, 1))
        # day affine + softsign: z = y . W[day] + b[day]  (x.W, not x.W^T: the B operand is MN-major W[day])
        dw_hl = self._w(P, pre + "day_weights")
        dev = x.device
        z = torch.empty(B, T, C, dtype=torch.float32, device=dev) if training else None
        s_hi = torch.empty(B, T, C, dtype=torch.bfloat16, device=dev)
        s_lo = torch.empty(B, T, C, dtype=torch.bfloat16, device=dev) if three else None
        ops.gemm(ops.kmajor(y_hi, y_lo), ops.mnmajor(dw_hl[0], dw_hl[1]), m_per_batch=T, batches=B, n=C, k=C, out_pre=z,
                 out_hi=s_hi, out_lo=s_lo, a_batched=True, b_batch_idx=day, bias=P[pre + "day_bias"], bias_batch_stride=C,
                 act="softsign")
        S.update(y_hi=y_hi, y_lo=y_lo, z=z, s_hi=s_hi, s_lo=s_lo, day=day)

        KL, ST = be.unfold_len, be.unfold_stride
        Tp = (T - KL) // ST + 1
        M = B * Tp
        H, ndir, L = be.gru_hidden, be.ndir, be.gru_layers
        e = pre + "neural_decoder.encoder."
        sfx = [""] + (["_reverse"] if ndir == 2 else [])
        S["Tp"] = Tp
        S["gru"] = []
        in_hi, in_lo = None, None
        for l in range(L):
            names = [e + f"gru.weight_ih_l{l}{s}" for s in sfx]
            b_ih = self._cat_vec(P, [e + f"gru.bias_ih_l{l}{s}" for s in sfx]) if be.bias else None
            gi = torch.empty(M, ndir * 3 * H, dtype=torch.float32, device=dev)
            if l == 0:
                w_hl = self._w_cat(P, f"gru_ih_l0_cat", names, unfold=(C, KL))
                ops.gemm(ops.unfold_view(s_hi, s_lo, KL, ST), ops.kmajor(w_hl[0], w_hl[1]), m_per_batch=Tp, batches=B,
                         n=ndir * 3 * H, k=C * KL, out_f32=gi, a_batched=True, bias=b_ih)
            else:
                w_hl = self._w_cat(P, f"gru_ih_l{l}_cat", names)
                ops.gemm(ops.kmajor(in_hi, in_lo), ops.kmajor(w_hl[0], w_hl[1]), m_per_batch=M, n=ndir * 3 * H, k=ndir * H,
                         out_f32=gi, bias=b_ih)
            w_hh = torch.stack([P[e + f"gru.weight_hh_l{l}{s}"] for s in sfx])
            if be.bias:
                b_hh = torch.stack([P[e + f"gru.bias_hh_l{l}{s}"] for s in sfx])
            else:
                b_hh = torch.zeros(ndir, 3 * H, dtype=torch.float32, device=dev)
            h0 = P[e + "hidden_start"][l * ndir:(l + 1) * ndir].contiguous() if be.learnable_h0 else None
            p_drop = be.dropout if (training and l < L - 1) else 0.0
            fw = ops.gru_fwd(gi, w_hh, b_hh, h0, B, Tp, H, ndir, three=three, training=training, drop_p=p_drop,
                             seed=_mix(seed, 100 + l))
            fw.update(w_hh=w_hh, h0=h0, drop_p=p_drop, seed=_mix(seed, 100 + l), in_hi=in_hi, in_lo=in_lo)
            S["gru"].append(fw)
            in_hi, in_lo = fw["out_hi"], fw["out_lo"]
        # FC stack (nn_helper.py:31-49): Linear, act, ..., Linear
        sizes = [ndir * H] + list(be.fc_hidden) + [be.out_size]
        S["fc"] = []
        a_hi, a_lo = in_hi, in_lo
        feats = None
        for i in range(len(sizes) - 1):
            last = i == len(sizes) - 2
            w_hl = self._w(P, e + f"fc.{2 * i}.weight")
            bias = P[e + f"fc.{2 * i}.bias"]
            if last:
                feats, _, _, _ = self._linear_fwd(a_hi, a_lo, w_hl, bias, M, sizes[i + 1], sizes[i])
                S["fc"].append(dict(a_hi=a_hi, a_lo=a_lo, pre=None))
            else:
                _, prea, hi, lo = self._linear_fwd(a_hi, a_lo, w_hl, bias, M, sizes[i + 1], sizes[i], want_f32=False,
                                                   want_split=True, want_pre=training, act=be.fc_act)
                S["fc"].append(dict(a_hi=a_hi, a_lo=a_lo, pre=prea))
                a_hi, a_lo = hi, lo
        return feats, S

    def brain_backward(self, P, S: dict, dfeats: Tensor, G: Dict[str, Tensor]) -> None:
        """dfeats [M, out_size] -> parameter gradients of the brain encoder written into G."""
        be = self.be
        three = self.three
        B, T, Tp = S["B"], S["T"], S["Tp"]
        M = B * Tp
        H, ndir, L = be.gru_hidden, be.ndir, be.gru_layers
        C = NEURAL_DIM
        pre = "brain_encoder."
        e = pre + "neural_decoder.encoder."
        sfx = [""] + (["_reverse"] if ndir == 2 else [])
        dev = dfeats.device
        sizes = [ndir * H] + list(be.fc_hidden) + [be.out_size]
        # ---- FC stack ----
        _, d_hi, d_lo = ops.dropout(dfeats, 0.0, 0, three=three, want_f32=False, want_split=True)
        d_f32 = dfeats
        for i in reversed(range(len(sizes) - 1)):
            st = S["fc"][i]
            wname = e + f"fc.{2 * i}.weight"
            G[wname] = self._wgrad(d_hi, d_lo, st["a_hi"], st["a_lo"], sizes[i + 1], sizes[i], M, name=wname)
            if d_f32 is not None:
                db = self._gout(e + f"fc.{2 * i}.bias", (sizes[i + 1],), dev, zero=True)
                ops.colsum(d_f32, db)
                G[e + f"fc.{2 * i}.bias"] = db
            w_hl = self._w(P, wname)
            if i > 0:
                # gradient w.r.t. the previous Linear's pre-activation, its column sums = that bias gradient
                db_prev = self._gout(e + f"fc.{2 * (i - 1)}.bias", (sizes[i],), dev, zero=True)
                _, d_hi, d_lo = self._dgrad(d_hi, d_lo, w_hl, M, sizes[i + 1], sizes[i], want_f32=False, want_split=True,
                                            aux=S["fc"][i - 1]["pre"], aux_mode=be.fc_act, colsum=db_prev)
                G[e + f"fc.{2 * (i - 1)}.bias"] = db_prev
                d_f32 = None
            else:
                dout, _, _ = self._dgrad(d_hi, d_lo, w_hl, M, sizes[i + 1], sizes[i])
        # ---- GRU stack ----
        for l in reversed(range(L)):
            fw = S["gru"][l]
            dh0 = torch.zeros(ndir, H, dtype=torch.float32, device=dev) if be.learnable_h0 else None
            bw = ops.gru_bwd(dout, fw, fw["w_hh"], fw["h0"], dh0, B, Tp, H, ndir, three=three, drop_p=fw["drop_p"], seed=fw["seed"])
            if dh0 is not None:
                G.setdefault(e + "hidden_start", torch.zeros_like(P[e + "hidden_start"]))[l * ndir:(l + 1) * ndir] = dh0
            # recurrent weights / biases
            for d, s in enumerate(sfx):
                G[e + f"gru.weight_hh_l{l}{s}"] = self._wgrad(bw["dgh_hi"][d], None if bw["dgh_lo"] is None else bw["dgh_lo"][d],
                                                              fw["hprev_hi"][d],
                                                              None if fw["hprev_lo"] is None else fw["hprev_lo"][d], 3 * H, H, M,
                                                              name=e + f"gru.weight_hh_l{l}{s}")
            db_ih = self._gout_cat([e + f"gru.bias_ih_l{l}{s}" for s in sfx], (ndir * 3 * H,), dev, zero=True)
            ops.colsum(bw["dgi"], db_ih)
            dbn = ops.zeros(ndir * H, device=dev)
            ops.colsum(bw["dghn"].reshape(ndir * M, H), dbn[:H], rows=M)
            if ndir == 2:
                ops.colsum(bw["dghn"][1], dbn[H:], rows=M)
            if be.bias:
                for d, s in enumerate(sfx):
                    seg = db_ih[d * 3 * H:(d + 1) * 3 * H]
                    G[e + f"gru.bias_ih_l{l}{s}"] = seg
                    G[e + f"gru.bias_hh_l{l}{s}"] = torch.cat([seg[:2 * H], dbn[d * H:(d + 1) * H]])
            names = [e + f"gru.weight_ih_l{l}{s}" for s in sfx]
            if l > 0:
                w_hl = self._w_cat(P, f"gru_ih_l{l}_cat", names)
                dw = self._wgrad(bw["dgi_hi"], bw["dgi_lo"], fw["in_hi"], fw["in_lo"], ndir * 3 * H, ndir * H, M, name=names)
                for d, n_ in enumerate(names):
                    G[n_] = dw[d * 3 * H:(d + 1) * 3 * H]
                dout, _, _ = self._dgrad(bw["dgi_hi"], bw["dgi_lo"], w_hl, M, ndir * 3 * H, ndir * H)
            else:
                KL, ST = be.unfold_len, be.unfold_stride
                w_hl = self._w_cat(P, "gru_ih_l0_cat", names, unfold=(C, KL))
                dgi_hi3 = bw["dgi_hi"].view(B, Tp, ndir * 3 * H)
                dgi_lo3 = bw["dgi_lo"].view(B, Tp, ndir * 3 * H) if bw["dgi_lo"] is not None else None
                dwp = torch.zeros(ndir * 3 * H, C * KL, dtype=torch.float32, device=dev)
                kb = B * ((Tp + 63) // 64)
                sk = _split_k_for(ndir * 3 * H, C * KL, kb)
                uv = ops.unfold_view(S["s_hi"], S["s_lo"], KL, ST, mn_major=True)
                ops.gemm(ops.mnmajor(dgi_hi3, dgi_lo3), uv, m_per_batch=ndir * 3 * H, n=C * KL, k=Tp, out_f32=dwp,
                         k_spans_batches=True, atomic=True, split_k=sk)
                dw = ops.unpermute_unfold_cols(dwp, C, KL, out=self._gout_cat(names, (ndir * 3 * H, C * KL), dev, zero=False))
                for d, n_ in enumerate(names):
                    G[n_] = dw[d * 3 * H:(d + 1) * 3 * H]
                # ds: dgrad folded through the overlapping unfold view (red.add)
                ds = torch.zeros(B, T, C, dtype=torch.float32, device=dev)
                ops.gemm(ops.kmajor(dgi_hi3, dgi_lo3), ops.mnmajor(w_hl[0], w_hl[1]), m_per_batch=Tp, batches=B, n=C * KL,
                         k=ndir * 3 * H, out_f32=ds, a_batched=True, atomic=True, ldc=ST * C, c_batch_stride=T * C)
        # ---- softsign + day affine ----
        dbias = self._gout(pre + "day_bias", P[pre + "day_bias"].shape, dev, zero=True)
        dz_hi, dz_lo = ops.softsign_bwd(ds, S["z"], S["day"], dbias, three=three)
        G[pre + "day_bias"] = dbias
        dW = self._gout(pre + "day_weights", P[pre + "day_weights"].shape, dev, zero=True)
        ops.gemm(ops.mnmajor(S["y_hi"], S["y_lo"]), ops.mnmajor(dz_hi, dz_lo), m_per_batch=C, batches=B, n=C, k=T, out_f32=dW,
                 a_batched=True, b_batched=True, c_batch_idx=S["day"], c_batch_stride=C * C, atomic=True)
        G[pre + "day_weights"] = dW

    # ------------------------------------------------------------------------------------------
    # wav2vec2 encoder (post-LN and stable-LN)   transformers modeling_wav2vec2.py:668-803
    # ------------------------------------------------------------------------------------------
    def _posconv_forward(self, P, q: str, x: Tensor, B: int, T: int, training: bool):
        cfg = self.w2v
        three = self.three
        D, K, G = cfg.hidden, cfg.pos_conv_k, cfg.pos_conv_groups
        Cg = D // G
        CP = ops.POSCONV_CP
        pad = K // 2
        Tpad = T + K
        gk, v = self._posconv_params(P, q)
        wg_hi, wg_lo, norm = self.weights.get(q + "pos_conv", _ver(gk, v),
                                              lambda: ops.posconv_pack_weight(gk, v, G, three))
        xg_hi, xg_lo = ops.posconv_pack_input(x, B, T, G, pad, Tpad, three)
        M = B * T
        out = torch.empty(M, D, dtype=torch.float32, device=x.device)
        pre = torch.empty(M, D, dtype=torch.float32, device=x.device) if training else None
        bias = P[q + "pos_conv_embed.conv.bias"]
        # all groups in one launch: batch = (group, trial); the weight / bias slot is the group, the output slot
        # (unit = Cg columns) is trial*T*G + group, so the G*B tiles fill the SMs (one group alone is B tiles)
        gi, ci = self._posconv_batch_idx(B, T, G, x.device)
        a = ops.unfold_view(xg_hi.view(G * B, Tpad, CP), None if xg_lo is None else xg_lo.view(G * B, Tpad, CP), K, 1)
        a.d1 = T   # conv output has T+1 frames, the last one is dropped (Wav2Vec2SamePadLayer)
        ops.gemm(a, ops.kmajor(wg_hi.view(G, Cg, K * CP), None if wg_lo is None else wg_lo.view(G, Cg, K * CP)),
                 m_per_batch=T, batches=G * B, n=Cg, k=K * CP, out_f32=out, out_pre=pre, ldc=D, c_batch_stride=Cg,
                 a_batched=True, b_batch_idx=gi, c_batch_idx=ci, bias=bias, bias_batch_stride=Cg, act="gelu", residual=x,
                 block_n=64)
        return out, dict(xg_hi=xg_hi, xg_lo=xg_lo, pre=pre, norm=norm, Tpad=Tpad, pad=pad)

    @staticmethod
    def _posconv_params(P, q: str):
        k0 = q + "pos_conv_embed.conv.parametrizations.weight.original0"
        if k0 in P:
            return P[k0], P[q + "pos_conv_embed.conv.parametrizations.weight.original1"]
        return P[q + "pos_conv_embed.conv.weight_g"], P[q + "pos_conv_embed.conv.weight_v"]

    def _posconv_backward(self, P, q: str, st: dict, dsum: Tensor, B: int, T: int, G_: Dict[str, Tensor]) -> Tensor:
        """dsum = gradient w.r.t. (x + gelu(conv(x))).  Returns the gradient w.r.t. x."""
        cfg = self.w2v
        three = self.three
        D, K, G = cfg.hidden, cfg.pos_conv_k, cfg.pos_conv_groups
        Cg = D // G
        CP = ops.POSCONV_CP
        M = B * T
        dev = dsum.device
        gk, v = self._posconv_params(P, q)
        wg_hi, wg_lo, norm = self.weights.get(q + "pos_conv", _ver(gk, v), lambda: ops.posconv_pack_weight(gk, v, G, three))
        dpre, dp_hi, dp_lo = ops.act_bwd(dsum, st["pre"], "gelu", three=three, want_f32=True, want_split=True)
        db = self._gout(q + "pos_conv_embed.conv.bias", (D,), dev, zero=True)
        ops.colsum(dpre, db)
        G_[q + "pos_conv_embed.conv.bias"] = db
        Tpad = st["Tpad"]
        dwg = ops.zeros(D, K * CP, device=dev)
        dp_hi3 = dp_hi.view(B, T, D)
        dp_lo3 = dp_lo.view(B, T, D) if dp_lo is not None else None
        # dgrad as a convolution of its own (no [T, K*CP] unfolded gradient, no red.add fold):
        #   dx[b,s,i] = sum_{k',o} dpp[b, s+k', o] * Wd[i, k'*CP+o],  dpp = dpre shifted right by K-1-pad frames,
        #   Wd[i, k'*CP+o] = W[o, (K-1-k')*CP+i]  (taps flipped, in/out channels swapped inside the group)
        wd_hi, wd_lo = self.weights.get(q + "pos_conv_dgrad", _ver(gk, v), lambda: self._posconv_dgrad_weight(wg_hi, wg_lo, G, Cg, K, CP))
        dpp_hi, dpp_lo = ops.posconv_pack_input(dpre, B, T, G, K - 1 - st["pad"], Tpad, three)
        dx = torch.empty_like(dsum)
        gi, ci = self._posconv_batch_idx(B, T, G, dev)
        ad = ops.unfold_view(dpp_hi.view(G * B, Tpad, CP), None if dpp_lo is None else dpp_lo.view(G * B, Tpad, CP), K, 1)
        ad.d1 = T
        ops.gemm(ad, ops.kmajor(wd_hi.view(G, Cg, K * CP), None if wd_lo is None else wd_lo.view(G, Cg, K * CP)),
                 m_per_batch=T, batches=G * B, n=Cg, k=K * CP, out_f32=dx, ldc=D, c_batch_stride=Cg, a_batched=True,
                 b_batch_idx=gi, c_batch_idx=ci, residual=dsum, block_n=64)
        if self._skip_wgrad:
            return dx
        for g in range(G):
            wsl = slice(g * Cg, (g + 1) * Cg)
            # A: [B, T, Cg] slice of dpre (row stride D); TMA needs 16-byte aligned base and strides
            if (g * Cg) % 8 == 0 and D % 8 == 0:
                a_k = ops.Operand(dp_hi3[:, :, g * Cg:], None if dp_lo3 is None else dp_lo3[:, :, g * Cg:], Cg, T, B, D, T * D)
            else:
                cpad = (Cg + 7) // 8 * 8
                th = torch.zeros(B, T, cpad, dtype=torch.bfloat16, device=dev)
                th[:, :, :Cg].copy_(dp_hi3[:, :, wsl])
                tl = None
                if dp_lo3 is not None:
                    tl = torch.zeros(B, T, cpad, dtype=torch.bfloat16, device=dev)
                    tl[:, :, :Cg].copy_(dp_lo3[:, :, wsl])
                a_k = ops.Operand(th, tl, Cg, T, B, cpad, T * cpad)
            # wgrad: dwg[o, k*CP+i] = sum_{b,t} dpre[b,t,o] * xg[b, t+k, i]
            uv = ops.unfold_view(st["xg_hi"][g], None if st["xg_lo"] is None else st["xg_lo"][g], K, 1, mn_major=True)
            uv.d1 = T
            ops.gemm(a_k.as_mn(), uv, m_per_batch=Cg, n=K * CP, k=T, out_f32=dwg[wsl], k_spans_batches=True, atomic=True,
                     split_k=max(1, min(8, B)))
        k0 = q + "pos_conv_embed.conv.parametrizations.weight.original0"
        kg, kv = (k0, q + "pos_conv_embed.conv.parametrizations.weight.original1") if k0 in P else \
            (q + "pos_conv_embed.conv.weight_g", q + "pos_conv_embed.conv.weight_v")
        dg, dv = ops.posconv_weight_bwd(dwg, gk, v, norm, dg_out=self._gout(kg, gk.shape, dev, zero=False),
                                        dv_out=self._gout(kv, v.shape, dev, zero=False))
        G_[kg], G_[kv] = dg, dv
        return dx

    def _posconv_batch_idx(self, B: int, T: int, G: int, dev):
        key = ("posconv_idx", B, T, G, str(dev))
        hit = self.weights._cache.get(key)
        if hit is None:
            g = torch.arange(G, dtype=torch.int32).repeat_interleave(B)
            b = torch.arange(B, dtype=torch.int32).repeat(G)
            hit = (g.to(dev), (b * (T * G) + g).to(dev))
            self.weights._cache[key] = hit
        return hit

    @staticmethod
    def _posconv_dgrad_weight(wg_hi: Tensor, wg_lo: Optional[Tensor], G: int, Cg: int, K: int, CP: int):
        """[D, K*CP] packed conv weight (row o, column k*CP+i) -> the data-gradient operand (row i, column
        k'*CP+o with k' = K-1-k): a pure permutation of the bf16 hi/lo pairs, so it is exact."""
        def one(w):
            if w is None:
                return None
            w4 = w.view(G, Cg, K, CP)[:, :, :, :Cg].flip(2).permute(0, 3, 2, 1)    # [G, i, k', o]
            out = torch.zeros(G, Cg, K, CP, dtype=w.dtype, device=w.device)
            out[:, :, :, :Cg] = w4
            return out.view(G * Cg, K * CP)
        return one(wg_hi), one(wg_lo)

    def w2v_base_forward(self, P, feats: Tensor, B: int, T: int, training: bool, seed: int, skip_layers=()):
        cfg = self.w2v
        three = self.three
        D, F, NH = cfg.hidden, cfg.ffn, cfg.heads
        hd = D // NH
        M = B * T
        q = "w2v_encoder.wav2vec2.encoder."
        tr = training
        p_h = cfg.hidden_dropout if tr else 0.0
        p_att = cfg.attention_dropout if tr else 0.0
        p_act = cfg.activation_dropout if tr else 0.0
        S: dict = {"B": B, "T": T, "layers": [], "skip": set(skip_layers)}
        xs, S["posconv"] = self._posconv_forward(P, q, feats, B, T, training)
        S["xs"] = xs
        if not cfg.stable_ln:
            h, h_hi, h_lo, mean, rstd = ops.layernorm_fwd(xs, P[q + "layer_norm.weight"], P[q + "layer_norm.bias"], cfg.ln_eps,
                                                          three=three, drop_p=p_h, seed=_mix(seed, 200))
            S["ln0"] = (mean, rstd)
        else:
            h, h_hi, h_lo = ops.dropout(xs, p_h, _mix(seed, 200), three=three, want_f32=True, want_split=False)
        for i in range(cfg.layers):
            if i in S["skip"]:
                S["layers"].append(None)
                continue
            Lq = q + f"layers.{i}."
            st: dict = {"seed": seed}
            wqkv = self._w_cat(P, Lq + "qkv", [Lq + "attention.q_proj.weight", Lq + "attention.k_proj.weight",
                                               Lq + "attention.v_proj.weight"])
            bqkv = self._cat_vec(P, [Lq + "attention.q_proj.bias", Lq + "attention.k_proj.bias", Lq + "attention.v_proj.bias"])
            if cfg.stable_ln:
                _, a_hi, a_lo, m1, r1 = ops.layernorm_fwd(h, P[Lq + "layer_norm.weight"], P[Lq + "layer_norm.bias"], cfg.ln_eps,
                                                          three=three, want_f32=False)
                st.update(ln1=(m1, r1), ln1_in=h)
            else:
                a_hi, a_lo = h_hi, h_lo
            st.update(in_hi=a_hi, in_lo=a_lo)
            qkv, _, _, _ = self._linear_fwd(a_hi, a_lo, wqkv, bqkv, M, 3 * D, D)
            _, c_hi, c_lo, probs = ops.attention_fwd(qkv, B, T, NH, hd, three=three, save_probs=training, drop_p=p_att,
                                                     seed=_mix(seed, 1000 + 10 * i + 1))
            st.update(qkv=qkv, probs=probs, c_hi=c_hi, c_lo=c_lo)
            u, _, _, _ = self._linear_fwd(c_hi, c_lo, self._w(P, Lq + "attention.out_proj.weight"),
                                          P[Lq + "attention.out_proj.bias"], M, D, D, residual=h, drop_p=p_h,
                                          drop_seed=_mix(seed, 1000 + 10 * i + 2))
            if cfg.stable_ln:
                _, h1_hi, h1_lo, m2, r2 = ops.layernorm_fwd(u, P[Lq + "final_layer_norm.weight"], P[Lq + "final_layer_norm.bias"],
                                                            cfg.ln_eps, three=three, want_f32=False)
                res2 = u
            else:
                h1, h1_hi, h1_lo, m2, r2 = ops.layernorm_fwd(u, P[Lq + "layer_norm.weight"], P[Lq + "layer_norm.bias"],
                                                             cfg.ln_eps, three=three)
                res2 = h1
            st.update(u=u, ln2=(m2, r2), h1_hi=h1_hi, h1_lo=h1_lo)
            _, pre1, f_hi, f_lo = self._linear_fwd(h1_hi, h1_lo, self._w(P, Lq + "feed_forward.intermediate_dense.weight"),
                                                   P[Lq + "feed_forward.intermediate_dense.bias"], M, F, D, want_f32=False,
                                                   want_split=True, want_pre=training, act=cfg.hidden_act, drop_p=p_act,
                                                   drop_seed=_mix(seed, 1000 + 10 * i + 3))
            # (the conformer's SiLU feed-forward stores the backward factor instead, b2t_gemm pre_mode 1; for GELU the
            # extra erf-derivative in the forward epilogue measured +29 us against -26 us in backward: not taken)
            st.update(pre1=pre1, f_hi=f_hi, f_lo=f_lo)
            u2, _, _, _ = self._linear_fwd(f_hi, f_lo, self._w(P, Lq + "feed_forward.output_dense.weight"),
                                           P[Lq + "feed_forward.output_dense.bias"], M, D, F, residual=res2, drop_p=p_h,
                                           drop_seed=_mix(seed, 1000 + 10 * i + 4))
            if cfg.stable_ln:
                h = u2
            else:
                h, h_hi, h_lo, m3, r3 = ops.layernorm_fwd(u2, P[Lq + "final_layer_norm.weight"],
                                                          P[Lq + "final_layer_norm.bias"], cfg.ln_eps, three=three)
                st.update(u2=u2, ln3=(m3, r3))
            S["layers"].append(st)
        if cfg.stable_ln:
            hin = h
            h, h_hi, h_lo, mf, rf = ops.layernorm_fwd(hin, P[q + "layer_norm.weight"], P[q + "layer_norm.bias"], cfg.ln_eps,
                                                      three=three)
            S["lnf"] = (mf, rf, hin)
        S["p"] = (p_h, p_att, p_act)
        S["seed"] = seed
        return h, S

    def w2v_base_backward(self, P, S: dict, dh: Tensor, G: Dict[str, Tensor]) -> Tensor:
        """dh = gradient w.r.t. the encoder output [M, D]; returns the gradient w.r.t. feats."""
        cfg = self.w2v
        three = self.three
        D, F, NH = cfg.hidden, cfg.ffn, cfg.heads
        hd = D // NH
        B, T = S["B"], S["T"]
        M = B * T
        dev = dh.device
        q = "w2v_encoder.wav2vec2.encoder."
        p_h, p_att, p_act = S["p"]
        seed = S["seed"]

        def gz(name, n):
            return self._gout(name, (n,), dev, zero=True)

        ds_ws = torch.empty(B, NH, T, T, dtype=torch.float32, device=dev)
        if cfg.stable_ln:
            mf, rf, hin = S["lnf"]
            G[q + "layer_norm.weight"], G[q + "layer_norm.bias"] = gz(q + "layer_norm.weight", D), gz(q + "layer_norm.bias", D)
            dh, _, _ = ops.layernorm_bwd(dh, hin, P[q + "layer_norm.weight"], mf, rf, G[q + "layer_norm.weight"],
                                         G[q + "layer_norm.bias"], three=three, want_split=False)
        for i in reversed(range(cfg.layers)):
            self.emit_grads(G)   # gradients of the layers above are complete: start their all-reduce (data parallel)
            st = S["layers"][i]
            if st is None:
                continue
            Lq = q + f"layers.{i}."
            s2, s3, s4, s1 = (_mix(seed, 1000 + 10 * i + k) for k in (2, 3, 4, 1))
            # ---- FFN block ----
            db2 = gz(Lq + "feed_forward.output_dense.bias", D)
            if cfg.stable_ln:
                # u2 = u + drop(FFN(LN(u)));  dh is d(u2)
                du2 = dh
                _, g_hi, g_lo = ops.dropout(du2, p_h, s4, three=three, want_f32=False, want_split=True)
                ops.colsum(ops.dropout(du2, p_h, s4, three=three)[0] if p_h > 0 else du2, db2)
            else:
                m3, r3 = st["ln3"]
                G[Lq + "final_layer_norm.weight"], G[Lq + "final_layer_norm.bias"] = (gz(Lq + "final_layer_norm.weight", D),
                                                                                      gz(Lq + "final_layer_norm.bias", D))
                du2, g_hi, g_lo = ops.layernorm_bwd(dh, st["u2"], P[Lq + "final_layer_norm.weight"], m3, r3,
                                                    G[Lq + "final_layer_norm.weight"], G[Lq + "final_layer_norm.bias"], db2,
                                                    three=three, br_drop_p=p_h, br_seed=s4)
            G[Lq + "feed_forward.output_dense.bias"] = db2
            G[Lq + "feed_forward.output_dense.weight"] = self._wgrad(g_hi, g_lo, st["f_hi"], st["f_lo"], D, F, M,
                                                                     name=Lq + "feed_forward.output_dense.weight")
            db1 = gz(Lq + "feed_forward.intermediate_dense.bias", F)
            _, dp_hi, dp_lo = self._dgrad(g_hi, g_lo, self._w(P, Lq + "feed_forward.output_dense.weight"), M, D, F,
                                          want_f32=False, want_split=True, aux=st["pre1"], aux_mode=cfg.hidden_act,
                                          drop_p=p_act, drop_seed=s3, colsum=db1)
            G[Lq + "feed_forward.intermediate_dense.bias"] = db1
            G[Lq + "feed_forward.intermediate_dense.weight"] = self._wgrad(dp_hi, dp_lo, st["h1_hi"], st["h1_lo"], F, D, M,
                                                                           name=Lq + "feed_forward.intermediate_dense.weight")
            w1 = self._w(P, Lq + "feed_forward.intermediate_dense.weight")
            m2, r2 = st["ln2"]
            dbo = gz(Lq + "attention.out_proj.bias", D)
            if cfg.stable_ln:
                # d(LN2 out) -> LN2 backward on u, plus the residual du2
                dl, _, _ = self._dgrad(dp_hi, dp_lo, w1, M, F, D)
                G[Lq + "final_layer_norm.weight"], G[Lq + "final_layer_norm.bias"] = (gz(Lq + "final_layer_norm.weight", D),
                                                                                      gz(Lq + "final_layer_norm.bias", D))
                dxl, _, _ = ops.layernorm_bwd(dl, st["u"], P[Lq + "final_layer_norm.weight"], m2, r2,
                                              G[Lq + "final_layer_norm.weight"], G[Lq + "final_layer_norm.bias"], three=three,
                                              want_split=False)
                du = ops.axpby(dxl, du2)
                _, b_hi, b_lo = ops.dropout(du, p_h, s2, three=three, want_f32=False, want_split=True)
                ops.colsum(ops.dropout(du, p_h, s2, three=three)[0] if p_h > 0 else du, dbo)
            else:
                dh1, _, _ = self._dgrad(dp_hi, dp_lo, w1, M, F, D, residual=du2)
                G[Lq + "layer_norm.weight"], G[Lq + "layer_norm.bias"] = gz(Lq + "layer_norm.weight", D), gz(Lq + "layer_norm.bias", D)
                du, b_hi, b_lo = ops.layernorm_bwd(dh1, st["u"], P[Lq + "layer_norm.weight"], m2, r2, G[Lq + "layer_norm.weight"],
                                                   G[Lq + "layer_norm.bias"], dbo, three=three, br_drop_p=p_h, br_seed=s2)
            # ---- attention block ----
            G[Lq + "attention.out_proj.bias"] = dbo
            G[Lq + "attention.out_proj.weight"] = self._wgrad(b_hi, b_lo, st["c_hi"], st["c_lo"], D, D, M,
                                                              name=Lq + "attention.out_proj.weight")
            dctx, _, _ = self._dgrad(b_hi, b_lo, self._w(P, Lq + "attention.out_proj.weight"), M, D, D)
            dqkv, dq_hi, dq_lo = ops.attention_bwd(st["qkv"], st["probs"], dctx, ds_ws, B, T, NH, hd, three=three,
                                                   drop_p=p_att, seed=s1)
            qkv_names = [Lq + f"attention.{nm}" for nm in ("q_proj", "k_proj", "v_proj")]
            dbq = self._gout_cat([n_ + ".bias" for n_ in qkv_names], (3 * D,), dev, zero=True)
            ops.colsum(dqkv, dbq)
            dwq = self._wgrad(dq_hi, dq_lo, st["in_hi"], st["in_lo"], 3 * D, D, M, name=[n_ + ".weight" for n_ in qkv_names])
            for j, nm in enumerate(("q_proj", "k_proj", "v_proj")):
                G[Lq + f"attention.{nm}.weight"] = None if dwq is None else dwq[j * D:(j + 1) * D]
                G[Lq + f"attention.{nm}.bias"] = dbq[j * D:(j + 1) * D]
            wqkv = self._w_cat(P, Lq + "qkv", [Lq + "attention.q_proj.weight", Lq + "attention.k_proj.weight",
                                               Lq + "attention.v_proj.weight"])
            if cfg.stable_ln:
                dl, _, _ = self._dgrad(dq_hi, dq_lo, wqkv, M, 3 * D, D)
                m1, r1 = st["ln1"]
                G[Lq + "layer_norm.weight"], G[Lq + "layer_norm.bias"] = gz(Lq + "layer_norm.weight", D), gz(Lq + "layer_norm.bias", D)
                dxl, _, _ = ops.layernorm_bwd(dl, st["ln1_in"], P[Lq + "layer_norm.weight"], m1, r1, G[Lq + "layer_norm.weight"],
                                              G[Lq + "layer_norm.bias"], three=three, want_split=False)
                dh = ops.axpby(dxl, du)
            else:
                dh, _, _ = self._dgrad(dq_hi, dq_lo, wqkv, M, 3 * D, D, residual=du)
        # ---- encoder input: [LN ->] dropout, positional conv ----
        if not cfg.stable_ln:
            mean, rstd = S["ln0"]
            G[q + "layer_norm.weight"], G[q + "layer_norm.bias"] = gz(q + "layer_norm.weight", D), gz(q + "layer_norm.bias", D)
            dxs, _, _ = ops.layernorm_bwd(dh, S["xs"], P[q + "layer_norm.weight"], mean, rstd, G[q + "layer_norm.weight"],
                                          G[q + "layer_norm.bias"], three=three, want_split=False, out_drop_p=p_h,
                                          out_seed=_mix(seed, 200))
        else:
            dxs = ops.dropout(dh, p_h, _mix(seed, 200), three=three)[0] if p_h > 0 else dh
        return self._posconv_backward(P, q, S["posconv"], dxs, B, T, G)

    # ------------------------------------------------------------------------------------------
    # lm_head + CTC     (w2v_custom_feat_extractor.py:149-152, 59, 81-90)
    # ------------------------------------------------------------------------------------------
    def head_forward(self, P, h: Tensor, B: int, T: int, training: bool, seed: int):
        cfg = self.w2v
        p_f = cfg.final_dropout if training else 0.0
        hd, hd_hi, hd_lo = ops.dropout(h, p_f, _mix(seed, 300), three=self.three, want_f32=True, want_split=True)
        logits, _, _, _ = self._linear_fwd(hd_hi, hd_lo, self._w(P, "w2v_encoder.lm_head.weight"),
                                           P["w2v_encoder.lm_head.bias"], B * T, cfg.vocab, cfg.hidden, block_n=64)
        return logits.view(B, T, cfg.vocab), hd, dict(hd_hi=hd_hi, hd_lo=hd_lo, p_f=p_f, seed=_mix(seed, 300), B=B, T=T)

    def head_backward(self, P, S: dict, dlogits: Tensor, G: Dict[str, Tensor]) -> Tensor:
        cfg = self.w2v
        B, T = S["B"], S["T"]
        M = B * T
        d2 = dlogits.reshape(M, cfg.vocab)
        _, d_hi, d_lo = ops.dropout(d2, 0.0, 0, three=self.three, want_f32=False, want_split=True)
        db = self._gout("w2v_encoder.lm_head.bias", (cfg.vocab,), d2.device, zero=True)
        ops.colsum(d2, db)
        G["w2v_encoder.lm_head.bias"] = db
        G["w2v_encoder.lm_head.weight"] = self._wgrad(d_hi, d_lo, S["hd_hi"], S["hd_lo"], cfg.vocab, cfg.hidden, M,
                                                      name="w2v_encoder.lm_head.weight")
        dh, _, _ = self._dgrad(d_hi, d_lo, self._w(P, "w2v_encoder.lm_head.weight"), M, cfg.vocab, cfg.hidden,
                               drop_p=S["p_f"], drop_seed=S["seed"])
        return dh

    # ------------------------------------------------------------------------------------------
    # whole model
    # ------------------------------------------------------------------------------------------
    def processed_lens(self, input_lens: Tensor) -> Tensor:
        """b2p2t_model.py:170-173: trunc((len - kernel) / stride) as int32."""
        be = self.be
        return ((input_lens - be.unfold_len) / be.unfold_stride).to(torch.int32)

    def forward(self, P, x: Tensor, day: Tensor, training: bool, seed: int = 0, skip_layers=(), aug=(0.0, 0.0)):
        """-> (logits [B,T',V], hidden [B,T',D], feats [B*T', D], saved)"""
        feats, Sb = self.brain_forward(P, x, day, training, seed, aug)
        B, Tp = Sb["B"], Sb["Tp"]
        if self.w2v.kind == "base":
            h, Se = self.w2v_base_forward(P, feats, B, Tp, training, seed, skip_layers)
        else:
            from . import conformer
            h, Se = conformer.forward(self, P, feats, B, Tp, training, seed, skip_layers)
        logits, hd, Sh = self.head_forward(P, h, B, Tp, training, seed)
        return logits, hd.view(B, Tp, -1), feats, dict(brain=Sb, enc=Se, head=Sh)

    def backward(self, P, saved: dict, dlogits: Tensor, dfeats_extra: Optional[Tensor] = None) -> Dict[str, Tensor]:
        G: Dict[str, Tensor] = {}
        dh = self.head_backward(P, saved["head"], dlogits, G)
        if self.w2v.kind == "base":
            dfeats = self.w2v_base_backward(P, saved["enc"], dh, G)
        else:
            from . import conformer
            dfeats = conformer.backward(self, P, saved["enc"], dh, G)
        if dfeats_extra is not None:
            dfeats = ops.axpby(dfeats, dfeats_extra)
        self.brain_backward(P, saved["brain"], dfeats, G)
        return G
