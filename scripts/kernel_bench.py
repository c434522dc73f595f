"""CUDA-event timings of the non-GEMM kernels at the shapes of the base (B=64) and conformer (B=32) steps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from wav2vec2forbrain_b200 import ops

def timeit(name, fn, iters=10, bytes_=None):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(4_000_000)   # ~2 ms of GPU backlog so that the host's launch overhead is not what is timed
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    extra = f"  {bytes_ / us / 1e3:7.1f} GB/s algorithmic" if bytes_ else ""
    print(f"{name:48s} {us:9.1f} us{extra}", flush=True)

def rnd(*s): return torch.randn(*s, device="cuda")

def gru(B, T=118, H=512, ndir=2):
    M = B * T
    gi = rnd(M, ndir * 3 * H); w = rnd(ndir, 3 * H, H) * H ** -0.5; b = rnd(ndir, 3 * H) * 0.1; dout = rnd(M, ndir * H)
    fw = ops.gru_fwd(gi, w, b, None, B, T, H, ndir)
    timeit(f"gru_fwd B={B}", lambda: ops.gru_fwd(gi, w, b, None, B, T, H, ndir))
    timeit(f"gru_bwd B={B}", lambda: ops.gru_bwd(dout, fw, w, None, None, B, T, H, ndir))

def attn(B, H, d=64, T=118):
    qkv = rnd(B * T, 3 * H * d); dctx = rnd(B * T, H * d)
    _, _, _, probs = ops.attention_fwd(qkv, B, T, H, d)
    ws = torch.empty(1, device="cuda")
    timeit(f"attention_fwd B={B} H={H}", lambda: ops.attention_fwd(qkv, B, T, H, d, drop_p=0.1, seed=5))
    timeit(f"attention_bwd B={B} H={H}", lambda: ops.attention_bwd(qkv, probs, dctx, ws, B, T, H, d, drop_p=0.1, seed=5))

def ln(M, n):
    x = rnd(M, n); dy = rnd(M, n); dres = rnd(M, n); g = rnd(n); b = rnd(n)
    _, _, _, mean, rstd = ops.layernorm_fwd(x, g, b, 1e-5)
    dg, db, dc = (torch.zeros(n, device="cuda") for _ in range(3))
    timeit(f"layernorm_fwd [{M},{n}] f32+hi/lo", lambda: ops.layernorm_fwd(x, g, b, 1e-5), bytes_=M * n * 12)
    timeit(f"layernorm_bwd [{M},{n}] dres+dcol+hi/lo", lambda: ops.layernorm_bwd(dy, x, g, mean, rstd, dg, db, dc, dres=dres, br_drop_p=0.1, br_seed=3),
           bytes_=M * n * 20)

def conv(B, T=118, D=1024, K=31):
    M = B * T
    g = rnd(M, 2 * D); w = rnd(D, 1, K) * 0.2; ga = rnd(D); be = rnd(D); rm = torch.zeros(D, device="cuda"); rv = torch.ones(D, device="cuda")
    ds = rnd(M, D)
    dwo, stats = ops.glu_dwconv_fwd(g, w, B, T)
    _, _, mean, rstd = ops.bn_silu_fwd(dwo, stats, ga, be, 1e-5, 0.1, rm, rv, None, True)
    dgam, dbet = ops.bn_silu_bwd_stats(ds, dwo, mean, rstd, ga, be)
    timeit("glu_dwconv_fwd", lambda: ops.glu_dwconv_fwd(g, w, B, T), bytes_=M * D * 12)
    timeit("bn_silu_fwd", lambda: ops.bn_silu_fwd(dwo, stats, ga, be, 1e-5, 0.1, rm, rv, None, True), bytes_=M * D * 8)
    timeit("bn_silu_bwd_stats", lambda: ops.bn_silu_bwd_stats(ds, dwo, mean, rstd, ga, be), bytes_=M * D * 8)
    timeit("glu_dwconv_bwd", lambda: ops.glu_dwconv_bwd(ds, dwo, g, w, mean, rstd, ga, be, dgam, dbet, B, T, True), bytes_=M * D * 24)

def rope(B, T=118, D=1024, hd=64):
    M = B * T
    x = rnd(M, D); g = rnd(D); b = rnd(D); inv = torch.exp(-torch.arange(0, hd, 2, device="cuda").float() / hd * 9.21)
    cos, sin = ops.rope_table(inv, T, hd)
    drot = rnd(M, D); dplain = rnd(M, D)
    timeit(f"ln_rope_fwd [{M},{D}] plain+rotated hi/lo", lambda: ops.layernorm_rope_fwd(x, g, b, 1e-5, cos, sin, T, hd), bytes_=M * D * 12)
    timeit(f"rope_bwd_add [{M},{D}]", lambda: ops.rope_bwd_add(drot, dplain, cos, sin, T, hd), bytes_=M * D * 12)


def pre(B, T=500, C=256):
    x = rnd(B, T, C); taps = torch.softmax(rnd(20), 0).repeat(C, 1).contiguous()
    timeit(f"smooth_fwd B={B} (hi/lo out)", lambda: ops.smooth_fwd(x, taps, 9), bytes_=B * T * C * 8)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "pre":
        pre(64); pre(32); sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "ln":
        ln(3776, 1024); sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "rope":
        rope(32); sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "conv":
        conv(32); sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "gru":
        gru(32); sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "attn":
        attn(32, 16); attn(64, 12); sys.exit(0)
    gru(64); gru(32)
    attn(64, 12); attn(32, 16)
    ln(7552, 768); ln(3776, 1024)
    conv(32)
    rope(32)
    pre(64)
