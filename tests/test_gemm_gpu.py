"""tcgen05 GEMM (b2t_gemm) against fp64 torch matmul of the same operands."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from wav2vec2forbrain_b200 import ops
    return ops


def rel(a, b):
    a = a.double()
    b = b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, generator=g, device="cuda") * scale


@pytest.mark.parametrize("three", [False, True])
@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (300, 200, 192), (1000, 768, 520)])
def test_kmajor(three, m, n, k):
    ops = _ops()
    a, b = rnd(m, k, seed=1), rnd(n, k, seed=2)
    ah, al = ops.split(a, three)
    bh, bl = ops.split(b, three)
    out = torch.full((m, n), float("nan"), device="cuda")
    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out)
    if three:
        ref = a.double() @ b.double().t()
        tol = 2e-5
    else:
        ref = ah.double() @ bh.double().t()
        tol = 2e-6
    assert rel(out, ref) < tol


@pytest.mark.parametrize("a_mn,b_mn", [(True, False), (False, True), (True, True)])
@pytest.mark.parametrize("three", [False, True])
def test_mn_major(a_mn, b_mn, three):
    ops = _ops()
    m, n, k = 328, 264, 200
    a, b = rnd(m, k, seed=3), rnd(n, k, seed=4)
    ah, al = ops.split(a, three, transpose=a_mn)   # [k, m] when a_mn
    bh, bl = ops.split(b, three, transpose=b_mn)
    A = ops.mnmajor(ah, al) if a_mn else ops.kmajor(ah, al)
    Bm = ops.mnmajor(bh, bl) if b_mn else ops.kmajor(bh, bl)
    out = torch.full((m, n), float("nan"), device="cuda")
    ops.gemm(A, Bm, m_per_batch=m, n=n, k=k, out_f32=out)
    ref = a.double() @ b.double().t() if three else \
        (ah.double().t() if a_mn else ah.double()) @ (bh.double() if b_mn else bh.double().t())
    assert rel(out, ref) < (2e-5 if three else 2e-6)


@pytest.mark.parametrize("bn", [64, 128, 256])
def test_block_n_and_epilogue(bn):
    ops = _ops()
    m, n, k = 260, 328, 256
    a, b = rnd(m, k, seed=5, scale=0.3), rnd(n, k, seed=6, scale=0.3)
    bias, res = rnd(n, seed=7), rnd(m, n, seed=8)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    out = torch.empty(m, n, device="cuda")
    pre = torch.empty(m, n, device="cuda")
    hi = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    lo = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    colsum = torch.zeros(n, device="cuda")
    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, out_pre=pre, out_hi=hi,
             out_lo=lo, bias=bias, residual=res, colsum=colsum, act="gelu", alpha=0.5, block_n=bn)
    z = 0.5 * (a.double() @ b.double().t()) + bias.double()
    ref = torch.nn.functional.gelu(z) + res.double()
    assert rel(pre, z) < 2e-5
    assert rel(out, ref) < 2e-5
    assert rel(hi.double() + lo.double(), ref) < 3e-5
    assert rel(colsum, ref.sum(0)) < 1e-4


@pytest.mark.parametrize("mode", ["gelu", "silu", "softsign"])
def test_aux_derivative(mode):
    ops = _ops()
    m, n, k = 130, 96, 128
    a, b, aux = rnd(m, k, seed=9), rnd(n, k, seed=10), rnd(m, n, seed=11)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    out = torch.empty(m, n, device="cuda")
    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, aux=aux, aux_mode=mode)
    x = aux.double().requires_grad_(True)
    f = {"gelu": torch.nn.functional.gelu, "silu": torch.nn.functional.silu, "softsign": torch.nn.functional.softsign}[mode]
    f(x).sum().backward()
    ref = (a.double() @ b.double().t()) * x.grad
    assert rel(out, ref) < 2e-5


def test_split_k_atomic():
    ops = _ops()
    m, n, k = 256, 192, 2048
    a, b = rnd(m, k, seed=12), rnd(n, k, seed=13)
    ah, al = ops.split(a, transpose=True)
    bh, bl = ops.split(b, transpose=True)
    out = torch.ones(m, n, device="cuda")
    ops.gemm(ops.mnmajor(ah, al), ops.mnmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, atomic=True, split_k=5)
    ref = a.double() @ b.double().t() + 1.0
    assert rel(out, ref) < 2e-5


def test_batched_day_affine_pattern():
    """A batched [B, T, 256]; B operand picked per batch through b_batch_idx (MN-major W[day])."""
    ops = _ops()
    Bn, T, D, days = 5, 200, 256, 7
    y = rnd(Bn, T, D, seed=14)
    w = rnd(days, D, D, seed=15, scale=0.1)      # z = y @ w[day]
    bias = None
    day = torch.tensor([3, 0, 6, 3, 1], device="cuda", dtype=torch.int32)
    yh, yl = ops.split(y)
    wh, wl = ops.split(w)
    out = torch.empty(Bn, T, D, device="cuda")
    # B operand: n = output feature j contiguous, k = input feature c rows, d2 = day  -> MN-major
    ops.gemm(ops.kmajor(yh, yl), ops.mnmajor(wh, wl), m_per_batch=T, batches=Bn, n=D, k=D, out_f32=out,
             a_batched=True, b_batch_idx=day, act="softsign")
    ref = torch.nn.functional.softsign(torch.einsum("btd,bdk->btk", y.double(), w.double()[day.long()]))
    assert rel(out, ref) < 2e-5


def test_unfold_view_forward():
    """GRU layer-0 input projection as an implicit GEMM over the overlapping unfold view."""
    ops = _ops()
    Bn, T, C, KL, S, N = 3, 100, 256, 32, 4, 192
    s = rnd(Bn, T, C, seed=16)
    w = rnd(N, C * KL, seed=17, scale=0.05)      # reference column order c*KL + k
    sh, sl = ops.split(s)
    wh, wl = ops.permute_unfold_cols(w, C, KL)
    Tp = (T - KL) // S + 1
    out = torch.empty(Bn, Tp, N, device="cuda")
    ops.gemm(ops.unfold_view(sh, sl, KL, S), ops.kmajor(wh, wl), m_per_batch=Tp, batches=Bn, n=N, k=C * KL, out_f32=out,
             a_batched=True)
    u = s.double().unfold(1, KL, S)              # [B, Tp, C, KL] -> feature c*KL + k
    ref = u.reshape(Bn, Tp, C * KL) @ w.double().t()
    # K = 8192: the tensor core's fp32 accumulate truncates, error grows with the number of k-steps
    assert rel(out, ref) < 6e-5


def test_unfold_view_wgrad_and_fold():
    ops = _ops()
    Bn, T, C, KL, S, N = 3, 100, 256, 32, 4, 192
    s = rnd(Bn, T, C, seed=18)
    dg = rnd(Bn, (T - KL) // S + 1, N, seed=19)
    w = rnd(N, C * KL, seed=20, scale=0.05)
    Tp = dg.shape[1]
    sh, sl = ops.split(s)
    dh, dl = ops.split(dg)
    # wgrad: dW[n, k*C+c] = sum_{b,j} dg[b,j,n] * s[b, S*j + k, c]
    dw = torch.zeros(N, C * KL, device="cuda")
    ops.gemm(ops.mnmajor(dh, dl), ops.unfold_view(sh, sl, KL, S, mn_major=True), m_per_batch=N, n=C * KL, k=Tp,
             out_f32=dw, k_spans_batches=True, atomic=True, split_k=2)
    dw_ref_order = ops.unpermute_unfold_cols(dw, C, KL)
    u = s.double().unfold(1, KL, S).reshape(Bn * Tp, C * KL)
    ref = dg.double().reshape(Bn * Tp, N).t() @ u
    assert rel(dw_ref_order, ref) < 2e-5
    # dgrad folded straight into ds through the overlapping output view (red.add)
    wh, wl = ops.permute_unfold_cols(w, C, KL)
    wth, wtl = ops.split(w, transpose=False)     # unused; keeps the allocator honest
    ds = torch.zeros(Bn, T, C, device="cuda")
    ops.gemm(ops.kmajor(dh, dl), ops.mnmajor(wh, wl), m_per_batch=Tp, batches=Bn, n=C * KL, k=N, out_f32=ds,
             a_batched=True, atomic=True, ldc=S * C, c_batch_stride=T * C)
    sd = s.double().requires_grad_(True)
    (sd.unfold(1, KL, S).reshape(Bn, Tp, C * KL) @ w.double().t() * dg.double()).sum().backward()
    assert rel(ds, sd.grad) < 2e-5


def test_large_shapes_both_modes():
    ops = _ops()
    m, n, k = 7552, 3072, 768
    a, b = rnd(m, k, seed=21), rnd(n, k, seed=22, scale=0.03)
    for three in (False, True):
        ah, al = ops.split(a, three)
        bh, bl = ops.split(b, three)
        out = torch.empty(m, n, device="cuda")
        ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out)
        ref = (a.double() @ b.double().t()) if three else (ah.double() @ bh.double().t())
        assert rel(out, ref) < (2e-5 if three else 2e-6)


def test_simt_debug_gemm():
    ops = _ops()
    a, b, bias = rnd(70, 50, seed=23), rnd(40, 50, seed=24), rnd(40, seed=25)
    out = ops.gemm_simt(a, b, bias)
    assert rel(out, a.double() @ b.double().t() + bias.double()) < 1e-5


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("m,n,k", [(260, 328, 256), (1000, 768, 520), (3776, 1024, 1024), (128, 128, 4096)])
def test_stream_k_full_epilogue(bn, m, n, k):
    """stream-K schedule (k-block ranges per CTA, partial tiles summed through the workspace by the owner CTA)
    with every fused epilogue stage on; twice in a row (the flags must be left clean)."""
    ops = _ops()
    a, b = rnd(m, k, seed=5, scale=0.3), rnd(n, k, seed=6, scale=0.3)
    bias, res = rnd(n, seed=7), rnd(m, n, seed=8)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    z = 0.5 * (a.double() @ b.double().t()) + bias.double()
    ref = torch.nn.functional.gelu(z) + res.double()
    for _ in range(2):
        out = torch.full((m, n), float("nan"), device="cuda")
        pre = torch.empty(m, n, device="cuda")
        hi = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        lo = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        colsum = torch.zeros(n, device="cuda")
        ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, out_pre=pre, out_hi=hi,
                 out_lo=lo, bias=bias, residual=res, colsum=colsum, act="gelu", alpha=0.5, block_n=bn, stream_k=1)
        assert rel(pre, z) < 2e-5
        assert rel(out, ref) < 2e-5
        assert rel(hi.double() + lo.double(), ref) < 3e-5
        assert rel(colsum, ref.sum(0)) < 1e-4


@pytest.mark.parametrize("stream_k", [-1, 1])
def test_stream_k_wgrad_shape(stream_k):
    """weight-gradient shape (both operands MN-major, long K, few tiles): stream-K replaces split-K + red.add."""
    ops = _ops()
    m, n, k = 512, 384, 3776
    a, b = rnd(m, k, seed=9), rnd(n, k, seed=10)
    ah, al = ops.split(a, True, transpose=True)
    bh, bl = ops.split(b, True, transpose=True)
    out = torch.full((m, n), float("nan"), device="cuda")
    ops.gemm(ops.mnmajor(ah, al), ops.mnmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, stream_k=stream_k)
    assert rel(out, a.double() @ b.double().t()) < 2e-5


# ------------------------------------------------------------------------------------------------------
# cta_group::2 (CTA pair, 256 x 256 tiles): same contracts as the single-CTA tiles
# ------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("three", [False, True])
@pytest.mark.parametrize("m,n,k", [(256, 256, 64), (392, 520, 200), (1000, 768, 520), (3776, 1024, 1024)])
def test_pair_all_major_combinations(a_mn, b_mn, three, m, n, k):
    """odd row-block counts (the pair's second CTA has no rows), ragged N (zero-filled B rows), ragged K"""
    ops = _ops()
    a, b = rnd(m, k, seed=3), rnd(n, k, seed=4)
    ah, al = ops.split(a, three, transpose=a_mn)
    bh, bl = ops.split(b, three, transpose=b_mn)
    A = ops.mnmajor(ah, al) if a_mn else ops.kmajor(ah, al)
    Bm = ops.mnmajor(bh, bl) if b_mn else ops.kmajor(bh, bl)
    out = torch.full((m, n), float("nan"), device="cuda")
    ops.gemm(A, Bm, m_per_batch=m, n=n, k=k, out_f32=out, cta_group=2)
    ref = a.double() @ b.double().t() if three else \
        (ah.double().t() if a_mn else ah.double()) @ (bh.double() if b_mn else bh.double().t())
    assert rel(out, ref) < (2e-5 if three else 2e-6)
    # and the single-CTA kernel on the same operands gives the same sums up to accumulation order
    out1 = torch.empty_like(out)
    ops.gemm(A, Bm, m_per_batch=m, n=n, k=k, out_f32=out1, cta_group=1)
    assert rel(out, out1) < 1e-6


@pytest.mark.parametrize("stream_k", [-1, 1])
@pytest.mark.parametrize("m,n,k", [(260, 328, 256), (1000, 768, 520), (3776, 1024, 1024), (3776, 4096, 256)])
def test_pair_full_epilogue(stream_k, m, n, k):
    """every fused epilogue stage on a CTA pair, whole-tile and stream-K schedules; twice (flags left clean)"""
    ops = _ops()
    a, b = rnd(m, k, seed=5, scale=0.3), rnd(n, k, seed=6, scale=0.3)
    bias, res = rnd(n, seed=7), rnd(m, n, seed=8)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    z = 0.5 * (a.double() @ b.double().t()) + bias.double()
    ref = torch.nn.functional.gelu(z) + res.double()
    for _ in range(2):
        out = torch.full((m, n), float("nan"), device="cuda")
        pre = torch.empty(m, n, device="cuda")
        hi = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        lo = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        colsum = torch.zeros(n, device="cuda")
        ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, out_pre=pre, out_hi=hi,
                 out_lo=lo, bias=bias, residual=res, colsum=colsum, act="gelu", alpha=0.5, stream_k=stream_k, cta_group=2)
        assert rel(pre, z) < 2e-5
        assert rel(out, ref) < 2e-5
        assert rel(hi.double() + lo.double(), ref) < 3e-5
        assert rel(colsum, ref.sum(0)) < 1e-4


def test_pair_split_k_atomic_and_dropout_mask():
    ops = _ops()
    m, n, k = 1024, 512, 3776
    a, b = rnd(m, k, seed=12), rnd(n, k, seed=13)
    ah, al = ops.split(a, transpose=True)
    bh, bl = ops.split(b, transpose=True)
    out = torch.ones(m, n, device="cuda")
    ops.gemm(ops.mnmajor(ah, al), ops.mnmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, atomic=True, split_k=5, cta_group=2)
    assert rel(out, a.double() @ b.double().t() + 1.0) < 2e-5
    # the dropout mask is a function of (seed, output offset) only: pair and single-CTA tiles agree element by element
    m, n, k = 520, 768, 128
    a, b = rnd(m, k, seed=14), rnd(n, k, seed=15)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    o1, o2 = torch.empty(m, n, device="cuda"), torch.empty(m, n, device="cuda")
    for o, cg in ((o1, 1), (o2, 2)):
        ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=o, drop_p=0.3, drop_seed=99, cta_group=cg)
    assert torch.equal(o1 == 0, o2 == 0)
    assert 0.25 < (o2 == 0).float().mean().item() < 0.35
    assert rel(o2, o1) < 1e-6


def test_pair_batched_operands():
    """batches > 1: tile rows never cross a batch, the second CTA of a batch's last pair is empty (T = 300 -> 3 row blocks)"""
    ops = _ops()
    Bn, T, D, days = 5, 300, 256, 7
    y = rnd(Bn, T, D, seed=14)
    w = rnd(days, D, D, seed=15, scale=0.1)
    day = torch.tensor([3, 0, 6, 3, 1], device="cuda", dtype=torch.int32)
    yh, yl = ops.split(y)
    wh, wl = ops.split(w)
    out = torch.empty(Bn, T, D, device="cuda")
    ops.gemm(ops.kmajor(yh, yl), ops.mnmajor(wh, wl), m_per_batch=T, batches=Bn, n=D, k=D, out_f32=out,
             a_batched=True, b_batch_idx=day, act="softsign", cta_group=2)
    ref = torch.nn.functional.softsign(torch.einsum("btd,bdk->btk", y.double(), w.double()[day.long()]))
    assert rel(out, ref) < 2e-5


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("act", ["silu", "gelu", "softsign"])
@pytest.mark.parametrize("n", [768, 330])      # 330: unaligned for bf16 rows of 8 -> fallback epilogue
def test_pre_mode_stores_the_backward_factor(cg, act, n):
    """pre_mode 1: out_pre = act'(z) * dropout factor * out_scale, from the same keep / drop decisions as the value;
    aux_mode 'mul' multiplies by it in the backward GEMM."""
    ops = _ops()
    m, k = 520, 256
    a, b, bias = rnd(m, k, seed=21, scale=0.3), rnd(n, k, seed=22, scale=0.3), rnd(n, seed=23)
    ah, al = ops.split(a)
    bh, bl = ops.split(b)
    out = torch.empty(m, n, device="cuda")
    pre = torch.empty(m, n, device="cuda")
    p_drop, scale = 0.3, 0.5
    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, out_pre=pre, bias=bias, act=act,
             drop_p=p_drop, drop_seed=1234, out_scale=scale, pre_mode=1, cta_group=cg)
    z = (a.double() @ b.double().t() + bias.double()).requires_grad_(True)
    fz = {"silu": torch.nn.functional.silu, "gelu": torch.nn.functional.gelu, "softsign": torch.nn.functional.softsign}[act](z)
    dz, = torch.autograd.grad(fz.sum(), z)
    # the mask of the value: the same seed / offsets without pre_mode
    ref_out = torch.empty(m, n, device="cuda")
    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=ref_out, bias=bias, act=act,
             drop_p=p_drop, drop_seed=1234, out_scale=scale, cta_group=1)
    assert torch.equal(out == 0, ref_out == 0) and rel(out, ref_out) < 1e-6
    keep = (ref_out != 0).double()
    assert 0.25 < 1.0 - keep.mean().item() < 0.35
    assert rel(out, fz.detach() * keep / (1 - p_drop) * scale) < 2e-5
    assert rel(pre, dz * keep / (1 - p_drop) * scale) < 2e-5
    if n % 8:
        return      # (an MN-major operand needs 16-byte aligned rows)
    # backward: dX = (dY . W) * factor, bias gradient = its column sums
    dy, w = rnd(m, 192, seed=24, scale=0.3), rnd(192, n, seed=25, scale=0.3)     # dY [m, 192], W [192, n]
    dyh, dyl = ops.split(dy)
    wh, wl = ops.split(w)
    dx = torch.empty(m, n, device="cuda")
    cs = torch.zeros(n, device="cuda")
    ops.gemm(ops.kmajor(dyh, dyl), ops.mnmajor(wh, wl), m_per_batch=m, n=n, k=192, out_f32=dx, aux=pre, aux_mode="mul", colsum=cs,
             cta_group=cg)
    ref_dx = (dy.double() @ w.double()) * pre.double()
    assert rel(dx, ref_dx) < 2e-5
    assert rel(cs, ref_dx.sum(0)) < 1e-4


def test_dynamic_schedule_cluster_launch_control():
    """opt-in dynamic schedule (one cluster per work item, cluster launch control): same results as the fixed grid, on
    multi-wave pair tiles with every epilogue stage, single-CTA tiles, split-K, and a grid smaller than the device"""
    ops = _ops()
    from wav2vec2forbrain_b200 import _lib
    lib = _lib.lib()
    cases = [(3776, 4096, 256, 2, 1), (3776, 1024, 1024, 2, 1), (2048, 512, 3776, 2, 4), (1000, 328, 520, 1, 1), (260, 256, 64, 2, 1)]
    try:
        for m, n, k, cg, sk in cases:
            a, b = rnd(m, k, seed=31, scale=0.3), rnd(n, k, seed=32, scale=0.3)
            bias, res = rnd(n, seed=33), rnd(m, n, seed=34)
            ah, al = ops.split(a)
            bh, bl = ops.split(b)
            outs = []
            for dyn in (0, 1):
                lib.b2t_gemm_set_dynamic_schedule(dyn)
                if sk > 1:
                    out = torch.ones(m, n, device="cuda")
                    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, atomic=True, split_k=sk,
                             cta_group=cg)
                    outs.append((out,))
                else:
                    out = torch.full((m, n), float("nan"), device="cuda")
                    hi = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
                    cs = torch.zeros(n, device="cuda")
                    ops.gemm(ops.kmajor(ah, al), ops.kmajor(bh, bl), m_per_batch=m, n=n, k=k, out_f32=out, out_hi=hi, bias=bias,
                             residual=res, colsum=cs, act="gelu", drop_p=0.1, drop_seed=5, cta_group=cg)
                    outs.append((out, hi.float(), cs))
                torch.cuda.synchronize()
            for x0, x1 in zip(*outs):
                assert torch.isfinite(x1).all()
                assert rel(x1, x0) < 1e-5
            ref = a.double() @ b.double().t()
            if sk > 1:
                assert rel(outs[1][0], ref + 1.0) < 2e-5
    finally:
        lib.b2t_gemm_set_dynamic_schedule(0)
