"""Times b2t_gemm on the GEMM shapes of the conformer-large / base steps (CUDA events, L2 flushed between
iterations by cycling through operand copies larger than the 126 MB L2)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from wav2vec2forbrain_b200 import ops

def bench(name, m, n, k, three, a_mn=False, b_mn=False, split_k=1, bn=0, iters=20, out="f32", cg=0, stream_k=0):
    dev = "cuda"
    ncopy = max(2, int(200e6 // ((m * k + n * k) * 2 * (2 if three else 1))) + 1)
    A = [torch.randn((k, m) if a_mn else (m, k), device=dev) for _ in range(ncopy)]
    B = [torch.randn((k, n) if b_mn else (n, k), device=dev) for _ in range(ncopy)]
    As = [ops.split(a, three) for a in A]
    Bs = [ops.split(b, three) for b in B]
    del A, B
    o = torch.zeros(m, n, device=dev)
    hi = torch.empty(m, n, device=dev, dtype=torch.bfloat16) if out == "split" else None
    lo = torch.empty(m, n, device=dev, dtype=torch.bfloat16) if (out == "split" and three) else None
    def run(i):
        a, b = As[i % ncopy], Bs[i % ncopy]
        oa = ops.mnmajor(*a) if a_mn else ops.kmajor(*a)
        ob = ops.mnmajor(*b) if b_mn else ops.kmajor(*b)
        ops.gemm(oa, ob, m_per_batch=m, n=n, k=k, out_f32=o if out == "f32" or split_k > 1 else None, out_hi=hi, out_lo=lo,
                 atomic=split_k > 1, split_k=split_k, block_n=bn, cta_group=cg, stream_k=stream_k)
    for i in range(3): run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(6_000_000)   # ~3 ms of GPU backlog: the host's ~25 us per op call must not be what is timed
    e0.record()
    for i in range(iters): run(i)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    tf = 2.0 * m * n * k / us / 1e6
    print(f"{name:34s} m={m:5d} n={n:5d} k={k:5d} terms={'3' if three else '1'} cg={cg or 'auto':>4} bn={bn or 'auto':>4} splitk={split_k:2d} sk={stream_k:2d}  {us:8.1f} us  "
          f"{tf:7.1f} TF/s alg  {tf * (3 if three else 1):7.1f} TF/s mma", flush=True)

def bench_epi(name, m, n, k, bn=128, b_mn=False, iters=20, cg=0, **flags):
    """ffn-shaped GEMM with a chosen set of fused epilogue stages (bf16x3)."""
    dev = "cuda"
    ncopy = 3
    As = [ops.split(torch.randn(m, k, device=dev), True) for _ in range(ncopy)]
    Bs = [ops.split(torch.randn((k, n) if b_mn else (n, k), device=dev), True) for _ in range(ncopy)]
    kw = {}
    if flags.get("f32"): kw["out_f32"] = torch.empty(m, n, device=dev)
    if flags.get("pre"): kw["out_pre"] = torch.empty(m, n, device=dev)
    if flags.get("hi"):
        kw["out_hi"] = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        kw["out_lo"] = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
    if flags.get("bias"): kw["bias"] = torch.randn(n, device=dev)
    if flags.get("res"): kw["residual"] = torch.randn(m, n, device=dev)
    if flags.get("aux"): kw["aux"] = torch.randn(m, n, device=dev); kw["aux_mode"] = flags["aux"]
    if flags.get("colsum"): kw["colsum"] = torch.zeros(n, device=dev)
    if flags.get("act"): kw["act"] = flags["act"]
    if flags.get("drop"): kw["drop_p"] = 0.1; kw["drop_seed"] = 7
    if flags.get("prefac"): kw["pre_mode"] = 1
    def run(i):
        a, b = As[i % ncopy], Bs[i % ncopy]
        ops.gemm(ops.kmajor(*a), ops.mnmajor(*b) if b_mn else ops.kmajor(*b), m_per_batch=m, n=n, k=k, block_n=bn, cta_group=cg, **kw)
    for i in range(3): run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(6_000_000)
    e0.record()
    for i in range(iters): run(i)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    print(f"{name:44s} cg={cg} bn={bn} {us:8.1f} us  {2.0 * m * n * k / us / 1e6:7.1f} TF/s alg", flush=True)


def shapes_cg():
    """single-CTA tiles (auto block_n) against CTA-pair tiles on the GEMM shapes of the conformer-large / base steps"""
    M = 3776
    for three in (True, False):
        for cg, sk in ((1, -1), (2, -1), (2, 1)):
            ops.STREAM_K = sk > 0
            kw = dict(cg=cg, stream_k=sk)
            bench("ffn1 fwd (N=4096,K=1024)", M, 4096, 1024, three, out="split", **kw)
            bench("ffn2 fwd (N=1024,K=4096)", M, 1024, 4096, three, **kw)
            bench("qk proj (N=2048,K=1024)", M, 2048, 1024, three, out="split", **kw)
            bench("out proj (N=1024,K=1024)", M, 1024, 1024, three, **kw)
            bench("dgrad ffn2 (B mn-major)", M, 4096, 1024, three, b_mn=True, out="split", **kw)
            bench("dgrad ffn1 (B mn-major)", M, 1024, 4096, three, b_mn=True, **kw)
            bench("wgrad ffn1 (both mn-major)", 4096, 1024, M, three, a_mn=True, b_mn=True, **kw)
            if sk < 0:
                for spl in (2, 4):
                    bench("wgrad 1024x1024 splitk", 1024, 1024, M, three, a_mn=True, b_mn=True, split_k=spl, **kw)
            else:
                bench("wgrad 1024x1024 stream-k", 1024, 1024, M, three, a_mn=True, b_mn=True, **kw)
            bench("base ffn1 (M=7552,N=3072,K=768)", 7552, 3072, 768, three, out="split", **kw)
            bench("base ffn2 (M=7552,N=768,K=3072)", 7552, 768, 3072, three, **kw)
            bench("base qkv (M=7552,N=2304,K=768)", 7552, 2304, 768, three, **kw)
            bench("big square 8192^3", 8192, 8192, 8192, three, iters=5, **kw)


def shapes_sk():
    """pair tiles: whole-tile waves against the hybrid stream-K schedule (B2T_GEMM_DEBUG=8 drops the partial traffic)"""
    M = 3776
    for sk in (-1, 1):
        ops.STREAM_K = sk > 0
        kw = dict(cg=2, stream_k=sk)
        bench("ffn1 fwd (N=4096,K=1024)", M, 4096, 1024, True, out="split", **kw)
        bench("ffn2 fwd (N=1024,K=4096)", M, 1024, 4096, True, **kw)
        bench("qk proj (N=2048,K=1024)", M, 2048, 1024, True, out="split", **kw)
        bench("out proj (N=1024,K=1024)", M, 1024, 1024, True, **kw)
        bench("dgrad ffn1 (B mn-major)", M, 1024, 4096, True, b_mn=True, **kw)
        bench("wgrad ffn1 (both mn-major)", 4096, 1024, M, True, a_mn=True, b_mn=True, **kw)
        bench("wgrad 1024x1024", 1024, 1024, M, True, a_mn=True, b_mn=True, split_k=4 if sk < 0 else 1, **kw)
        bench("base ffn1 (M=7552,N=3072,K=768)", 7552, 3072, 768, True, out="split", **kw)
        bench("base ffn2 (M=7552,N=768,K=3072)", 7552, 768, 3072, True, **kw)
        bench("base qkv (M=7552,N=2304,K=768)", 7552, 2304, 768, True, **kw)
        bench("even: 74 tiles (M=9472,N=512,K=1024)", 9472, 512, 1024, True, **kw)
        for kk in (64, 256, 1024):
            bench(f"fixed cost: 60 tiles, K={kk}", M, 1024, kk, True, **kw)
            bench(f"fixed cost: 60 tiles, K={kk}, 1 term", M, 1024, kk, False, **kw)
        bench("111 tiles (M=9472,N=768,K=1024)", 9472, 768, 1024, True, **kw)


def shapes_fixed():
    """launch + prologue + first load + epilogue + teardown: one k-block GEMMs (B2T_GEMM_DEBUG strips pieces, CG=1 only)"""
    M = 3776
    x = torch.zeros(1 << 20, device="cuda")
    def empty_launches():
        torch.cuda._sleep(6_000_000)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): ops.axpby(x, None, 1.0, 0.0, out=x)
        e1.record(); torch.cuda.synchronize()
        print(f"20 back-to-back 4 MB elementwise launches: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us each", flush=True)
    empty_launches()
    for cg in (1, 2):
        for n in (1024, 4096):
            for kk in (64, 128, 1024):
                bench(f"fixed cost K={kk}", M, n, kk, True, cg=cg, stream_k=-1)
                bench(f"fixed cost K={kk} hi/lo out", M, n, kk, True, cg=cg, stream_k=-1, out="split")


EPI_CASES = [("f32", dict(f32=1)), ("hi/lo", dict(hi=1)), ("f32+hi/lo", dict(f32=1, hi=1)), ("hi/lo+aux(silu')", dict(hi=1, aux="silu")),
             ("hi/lo+colsum", dict(hi=1, colsum=1)), ("hi/lo+aux(silu')+colsum+drop", dict(hi=1, aux="silu", colsum=1, drop=1)),
             ("pre+hi/lo+bias+silu+drop", dict(pre=1, hi=1, bias=1, act="silu", drop=1)), ("pre+hi/lo+bias+gelu", dict(pre=1, hi=1, bias=1, act="gelu")),
             ("prefac+hi/lo+bias+silu+drop", dict(pre=1, hi=1, bias=1, act="silu", drop=1, prefac=1)),
             ("prefac+hi/lo+bias+gelu+drop", dict(pre=1, hi=1, bias=1, act="gelu", drop=1, prefac=1)),
             ("hi/lo+aux(mul)+colsum", dict(hi=1, aux="mul", colsum=1)),
             ("f32+bias", dict(f32=1, bias=1)), ("f32+bias+res", dict(f32=1, bias=1, res=1)), ("f32+bias+drop+res", dict(f32=1, bias=1, res=1, drop=1))]


if __name__ == "__main__":
    M = 3776
    if len(sys.argv) > 1 and sys.argv[1] == "epi2":   # fused epilogues on pair tiles (env B2T_GEMM_NO_TMA_EPILOGUE=1: transposing epilogue)
        ops.STREAM_K = False
        for nm, fl in EPI_CASES:
            bench_epi("ffn1 (N=4096,K=1024) " + nm, M, 4096, 1024, bn=0, cg=2, **fl)
        for nm, fl in EPI_CASES:
            bench_epi("outproj (N=1024,K=1024) " + nm, M, 1024, 1024, bn=0, cg=2, **fl)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "epi3":   # short run for ncu: single-wave GEMM with / without the residual input
        ops.STREAM_K = False
        bench_epi("outproj f32+bias", M, 1024, 1024, bn=0, iters=2, f32=1, bias=1)
        bench_epi("outproj f32+bias+res", M, 1024, 1024, bn=0, iters=2, f32=1, bias=1, res=1)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "prof2":   # short run for ncu --set full: the pair kernel, epilogue-bound and MMA-bound
        ops.STREAM_K = False
        bench("ffn1 K=64 (epilogue-bound)", M, 4096, 64, True, cg=2, stream_k=-1, out="split", iters=1)
        bench("ffn1 K=1024", M, 4096, 1024, True, cg=2, stream_k=-1, out="split", iters=1)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "fixed":
        shapes_fixed()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "sk":
        shapes_sk()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "cg":
        shapes_cg()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "epi1":   # the two heavy in-step epilogues, short (for ncu --set full)
        ops.STREAM_K = False
        bench_epi("ffn1 dgrad hi/lo+aux(silu')+colsum+drop", M, 4096, 1024, bn=0, b_mn=True, iters=2, hi=1, aux="silu", colsum=1, drop=1)
        bench_epi("ffn1 fwd pre+hi/lo+bias+silu+drop", M, 4096, 1024, bn=0, iters=2, pre=1, hi=1, bias=1, act="silu", drop=1)
        bench_epi("ffn2 fwd f32+bias+drop+res", M, 1024, 4096, bn=0, iters=2, f32=1, bias=1, res=1, drop=1)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "epi":
        ops.STREAM_K = False
        for bn in (128, 256):
            for nm, fl in [("f32", dict(f32=1)), ("hi/lo", dict(hi=1)), ("f32+hi/lo", dict(f32=1, hi=1)), ("hi/lo+aux(silu')", dict(hi=1, aux="silu")),
                           ("hi/lo+aux(gelu')", dict(hi=1, aux="gelu")), ("hi/lo+colsum", dict(hi=1, colsum=1)),
                           ("hi/lo+aux(silu')+colsum", dict(hi=1, aux="silu", colsum=1)),
                           ("pre+hi/lo+bias+silu", dict(pre=1, hi=1, bias=1, act="silu")),
                           ("pre+hi/lo+bias+gelu", dict(pre=1, hi=1, bias=1, act="gelu")),
                           ("hi/lo+bias+silu", dict(hi=1, bias=1, act="silu")), ("pre+hi/lo", dict(pre=1, hi=1)),
                           ("f32+bias", dict(f32=1, bias=1)), ("f32+bias+res", dict(f32=1, bias=1, res=1)),
                           ("f32+bias+drop+res", dict(f32=1, bias=1, res=1, drop=1))]:
                bench_epi("ffn1 " + nm, M, 4096, 1024, bn=bn, **fl)
            bench_epi("outproj f32+bias+res", M, 1024, 1024, bn=bn, f32=1, bias=1, res=1)
            bench_epi("outproj f32", M, 1024, 1024, bn=bn, f32=1)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "prof":     # short run for ncu --set full: 4 launches per shape
        ops.STREAM_K = False
        if len(sys.argv) > 2:     # component timing with B2T_GEMM_DEBUG set by the caller
            for bn in (128, 256):
                bench("ffn1 fwd (N=4096,K=1024)", M, 4096, 1024, True, bn=bn)
                bench("ffn2 fwd (N=1024,K=4096)", M, 1024, 4096, True, bn=bn)
                bench("out proj (N=1024,K=1024)", M, 1024, 1024, True, bn=bn)
                bench("big square 8192^3", 8192, 8192, 8192, True, bn=bn, iters=3)
            sys.exit(0)
        for three in (True, False):
            bench("ffn1 fwd (N=4096,K=1024)", M, 4096, 1024, three, bn=128, iters=1)
            bench("ffn2 fwd (N=1024,K=4096)", M, 1024, 4096, three, bn=256, iters=1)
        sys.exit(0)
    for sk in (False, True):
        ops.STREAM_K = sk
        print(f"--- stream-K {'on' if sk else 'off'}", flush=True)
        for three in ((True, False) if not sk or len(sys.argv) > 1 else (True,)):
            for bn in (128, 256):
                bench("ffn1 fwd (N=4096,K=1024)", M, 4096, 1024, three, bn=bn)
                bench("ffn2 fwd (N=1024,K=4096)", M, 1024, 4096, three, bn=bn)
                bench("qk proj (N=2048,K=1024)", M, 2048, 1024, three, bn=bn)
                bench("out proj (N=1024,K=1024)", M, 1024, 1024, three, bn=bn)
                bench("ffn1 fwd split out", M, 4096, 1024, three, bn=bn, out="split")
                bench("dgrad ffn2 (B mn-major)", M, 4096, 1024, three, b_mn=True, bn=bn)
                bench("wgrad ffn1 (both mn-major)", 4096, 1024, M, three, a_mn=True, b_mn=True, bn=bn)
                bench("wgrad 1024x1024 splitk", 1024, 1024, M, three, a_mn=True, b_mn=True, split_k=2, bn=bn)
                bench("base ffn1 (M=7552,N=3072,K=768)", 7552, 3072, 768, three, bn=bn)
                bench("base ffn2 (M=7552,N=768,K=3072)", 7552, 768, 3072, three, bn=bn)
                bench("base qkv (M=7552,N=2304,K=768)", 7552, 2304, 768, three, bn=bn)
            bench("big square 8192^3", 8192, 8192, 8192, three, bn=128, iters=5)
            bench("big square 8192^3", 8192, 8192, 8192, three, bn=256, iters=5)
