// Fused multi-head self-attention forward / backward, fp32 SIMT, one CTA per (trial, head).
// Replaces matmul*scale -> softmax -> dropout -> matmul of transformers
// modeling_wav2vec2.py:466-549 (eager attention) and modeling_wav2vec2_conformer.py:420-527.
// T' is ~118 on this path so a whole head (K, V tiles and one probability row per warp) lives in
// shared memory; scores never touch HBM except the saved probabilities needed by backward.
// Optional key-length mask (kv_lens); the reference path passes none.
#include "common.cuh"
#include "../../include/b2t_sm100.h"
#include <math_constants.h>

namespace {

constexpr int ATT_THREADS = 256;
constexpr int ATT_WARPS = ATT_THREADS / 32;

// qkv [B*T, ld] with q at col q_off + h*d, k at k_off + h*d, v at v_off + h*d
__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_kernel(const float* __restrict__ qkv, long ld, int q_off, int k_off, int v_off, const int* __restrict__ kv_lens,
                float* __restrict__ probs, float* __restrict__ ctx, bf16* __restrict__ ctx_hi, bf16* __restrict__ ctx_lo,
                long ld_ctx, int T, int H, int d, float scale, float drop_p, unsigned long long seed) {
    extern __shared__ float sm[];
    const int dp = d + 1;
    float* Ks = sm;                      // [T][d+1]
    float* Vs = Ks + T * dp;             // [T][d+1]
    float* qrow = Vs + T * dp;           // [warps][64]
    float* prow = qrow + ATT_WARPS * 64; // [warps][T]
    const int b = blockIdx.x / H, h = blockIdx.x % H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* base = qkv + (long)b * T * ld;
    for (int i = threadIdx.x; i < T * d; i += ATT_THREADS) {
        const int t = i / d, c = i - t * d;
        Ks[t * dp + c] = base[(long)t * ld + k_off + h * d + c];
        Vs[t * dp + c] = base[(long)t * ld + v_off + h * d + c];
    }
    __syncthreads();
    const int kv_len = kv_lens ? min(kv_lens[b], T) : T;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    float* q = qrow + warp * 64;
    float* p = prow + warp * T;
    for (int i = warp; i < T; i += ATT_WARPS) {
        for (int c = lane; c < d; c += 32) q[c] = base[(long)i * ld + q_off + h * d + c] * scale;
        __syncwarp();
        float mx = -CUDART_INF_F;
        for (int j = lane; j < T; j += 32) {
            float s = -CUDART_INF_F;
            if (j < kv_len) {
                s = 0.0f;
                const float* kr = Ks + j * dp;
#pragma unroll 8
                for (int c = 0; c < d; ++c) s = fmaf(q[c], kr[c], s);
            }
            p[j] = s;
            mx = fmaxf(mx, s);
        }
        mx = warp_max(mx);
        float sum = 0.0f;
        for (int j = lane; j < T; j += 32) {
            const float e = (p[j] == -CUDART_INF_F) ? 0.0f : expf(p[j] - mx);
            p[j] = e;
            sum += e;
        }
        sum = warp_sum(sum);
        const float inv = sum > 0.0f ? 1.0f / sum : 0.0f;
        const long prow_off = (((long)b * H + h) * T + i) * T;
        for (int j = lane; j < T; j += 32) {
            float pv = p[j] * inv;
            if (probs) probs[prow_off + j] = pv;
            if (drop_p > 0.0f)
                pv = hash_uniform(seed, (unsigned long long)(prow_off + j)) < drop_p ? 0.0f : pv * drop_scale;
            p[j] = pv;
        }
        __syncwarp();
        for (int c = lane; c < d; c += 32) {
            float acc = 0.0f;
#pragma unroll 4
            for (int j = 0; j < T; ++j) acc = fmaf(p[j], Vs[j * dp + c], acc);
            const long off = ((long)b * T + i) * ld_ctx + h * d + c;
            if (ctx) ctx[off] = acc;
            if (ctx_hi) {
                bf16 hh, ll;
                split_bf16(acc, hh, ll);
                ctx_hi[off] = hh;
                if (ctx_lo) ctx_lo[off] = ll;
            }
        }
        __syncwarp();
    }
}

// dctx [B*T, ld_ctx]; probs saved by forward; ds_ws scratch [B,H,T,T]; dqkv packed like qkv
__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_kernel(const float* __restrict__ qkv, long ld, int q_off, int k_off, int v_off, const float* __restrict__ probs,
                const float* __restrict__ dctx, long ld_ctx, float* __restrict__ ds_ws, float* __restrict__ dqkv,
                bf16* __restrict__ dqkv_hi, bf16* __restrict__ dqkv_lo, int T, int H, int d, float scale, float drop_p,
                unsigned long long seed) {
    extern __shared__ float sm[];
    const int dp = d + 1;
    float* As = sm;                       // phase A: K, phase B: Q
    float* Bs = As + T * dp;              // phase A: V, phase B: dO
    float* rowbuf = Bs + T * dp;          // [warps][64] dO row
    float* prow = rowbuf + ATT_WARPS * 64;   // [warps][T] dS row
    const int b = blockIdx.x / H, h = blockIdx.x % H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* base = qkv + (long)b * T * ld;
    const float* dob = dctx + (long)b * T * ld_ctx + h * d;
    float* dbase = dqkv + (long)b * T * ld;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    auto store = [&](int t, int col, float v) {
        const long off = (long)t * ld + col;
        dbase[off] = v;
        if (dqkv_hi) {
            const long goff = (long)b * T * ld + off;
            bf16 hh, ll;
            split_bf16(v, hh, ll);
            dqkv_hi[goff] = hh;
            if (dqkv_lo) dqkv_lo[goff] = ll;
        }
    };

    for (int i = threadIdx.x; i < T * d; i += ATT_THREADS) {
        const int t = i / d, c = i - t * d;
        As[t * dp + c] = base[(long)t * ld + k_off + h * d + c];
        Bs[t * dp + c] = base[(long)t * ld + v_off + h * d + c];
    }
    __syncthreads();
    // ---- phase A: rows.  dP = dO V^T, dS = P * (dP - sum(dP*P)), dQ = scale * dS K
    float* dor = rowbuf + warp * 64;
    float* dsr = prow + warp * T;
    for (int i = warp; i < T; i += ATT_WARPS) {
        for (int c = lane; c < d; c += 32) dor[c] = dob[(long)i * ld_ctx + c];
        __syncwarp();
        const long prow_off = (((long)b * H + h) * T + i) * T;
        float delta = 0.0f;
        for (int j = lane; j < T; j += 32) {
            float dpv = 0.0f;
            const float* vr = Bs + j * dp;
#pragma unroll 8
            for (int c = 0; c < d; ++c) dpv = fmaf(dor[c], vr[c], dpv);
            if (drop_p > 0.0f)
                dpv = hash_uniform(seed, (unsigned long long)(prow_off + j)) < drop_p ? 0.0f : dpv * drop_scale;
            const float pj = probs[prow_off + j];
            dsr[j] = dpv;
            delta += dpv * pj;
        }
        delta = warp_sum(delta);
        for (int j = lane; j < T; j += 32) {
            const float v = probs[prow_off + j] * (dsr[j] - delta) * scale;
            dsr[j] = v;
            ds_ws[prow_off + j] = v;
        }
        __syncwarp();
        for (int c = lane; c < d; c += 32) {
            float acc = 0.0f;
#pragma unroll 4
            for (int j = 0; j < T; ++j) acc = fmaf(dsr[j], As[j * dp + c], acc);
            store(i, q_off + h * d + c, acc);
        }
        __syncwarp();
    }
    __threadfence_block();
    __syncthreads();
    // ---- phase B: columns.  dK = dS^T Q, dV = Pdrop^T dO
    for (int i = threadIdx.x; i < T * d; i += ATT_THREADS) {
        const int t = i / d, c = i - t * d;
        As[t * dp + c] = base[(long)t * ld + q_off + h * d + c];
        Bs[t * dp + c] = dob[(long)t * ld_ctx + c];
    }
    __syncthreads();
    const long head_off = ((long)b * H + h) * T * T;
    for (int j = warp; j < T; j += ATT_WARPS) {
        float dk0 = 0.0f, dk1 = 0.0f, dv0 = 0.0f, dv1 = 0.0f;
        const int c0 = lane, c1 = lane + 32;
        for (int i = 0; i < T; ++i) {
            const long o = head_off + (long)i * T + j;
            const float dsv = ds_ws[o];
            float pv = probs[o];
            if (drop_p > 0.0f) pv = hash_uniform(seed, (unsigned long long)o) < drop_p ? 0.0f : pv * drop_scale;
            if (c0 < d) {
                dk0 = fmaf(dsv, As[i * dp + c0], dk0);
                dv0 = fmaf(pv, Bs[i * dp + c0], dv0);
            }
            if (c1 < d) {
                dk1 = fmaf(dsv, As[i * dp + c1], dk1);
                dv1 = fmaf(pv, Bs[i * dp + c1], dv1);
            }
        }
        if (c0 < d) {
            store(j, k_off + h * d + c0, dk0);
            store(j, v_off + h * d + c0, dv0);
        }
        if (c1 < d) {
            store(j, k_off + h * d + c1, dk1);
            store(j, v_off + h * d + c1, dv1);
        }
    }
}

size_t attn_smem(int T, int d) { return ((size_t)2 * T * (d + 1) + ATT_WARPS * 64 + (size_t)ATT_WARPS * T) * sizeof(float); }

}  // namespace

// single-tile tensor-core kernels (attention_mma.cu)
int b2t_attn_mma_supported(int T, int d, long ld, long ld_ctx, int q_off, int k_off, int v_off);
int b2t_attn_mma_fwd(const float* qkv, long ld, int q_off, int k_off, int v_off, const int* kv_lens, float* probs, float* ctx,
                     void* ctx_hi, void* ctx_lo, long ld_ctx, int B, int T, int H, int d, float scale, float drop_p,
                     unsigned long long seed, cudaStream_t stream);
int b2t_attn_mma_bwd(const float* qkv, long ld, int q_off, int k_off, int v_off, const float* probs, const float* dctx,
                     long ld_ctx, float* dqkv, void* dqkv_hi, void* dqkv_lo, int B, int T, int H, int d, float scale,
                     float drop_p, unsigned long long seed, cudaStream_t stream);

extern "C" int b2t_attention_fwd(const float* qkv, int64_t ld, int q_off, int k_off, int v_off, const int32_t* kv_lens,
                                 float* probs, float* ctx, void* ctx_hi, void* ctx_lo, int64_t ld_ctx, int B, int T, int H,
                                 int d, float scale, float drop_p, uint64_t seed, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(qkv && (ctx || ctx_hi) && B > 0 && T > 0 && H > 0 && d > 0 && d <= 64, "b2t_attention_fwd: bad arguments (d=%d)", d);
    B2T_CHECK_ARG(drop_p >= 0.0f && drop_p < 1.0f, "b2t_attention_fwd: drop_p");
    if (b2t_attn_mma_supported(T, d, ld, ld_ctx, q_off, k_off, v_off))
        return b2t_attn_mma_fwd(qkv, ld, q_off, k_off, v_off, kv_lens, probs, ctx, ctx_hi, ctx_lo, ld_ctx, B, T, H, d, scale,
                                drop_p, (unsigned long long)seed, stream);
    const size_t smem = attn_smem(T, d);
    B2T_CHECK_ARG(smem <= 220 * 1024, "b2t_attention_fwd: T=%d does not fit the single-tile kernel", T);
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        attr = true;
    }
    attn_fwd_kernel<<<B * H, ATT_THREADS, smem, stream>>>(qkv, ld, q_off, k_off, v_off, kv_lens, probs, ctx, (bf16*)ctx_hi,
                                                          (bf16*)ctx_lo, ld_ctx, T, H, d, scale, drop_p,
                                                          (unsigned long long)seed);
    B2T_CHECK_LAUNCH("b2t_attention_fwd");
    return 0;
}

extern "C" int b2t_attention_bwd(const float* qkv, int64_t ld, int q_off, int k_off, int v_off, const float* probs,
                                 const float* dctx, int64_t ld_ctx, float* ds_ws, float* dqkv, void* dqkv_hi, void* dqkv_lo,
                                 int B, int T, int H, int d, float scale, float drop_p, uint64_t seed, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(qkv && probs && dctx && ds_ws && dqkv && B > 0 && T > 0 && H > 0 && d > 0 && d <= 64,
                  "b2t_attention_bwd: bad arguments");
    if (b2t_attn_mma_supported(T, d, ld, ld_ctx, q_off, k_off, v_off))
        return b2t_attn_mma_bwd(qkv, ld, q_off, k_off, v_off, probs, dctx, ld_ctx, dqkv, dqkv_hi, dqkv_lo, B, T, H, d, scale,
                                drop_p, (unsigned long long)seed, stream);
    const size_t smem = attn_smem(T, d);
    B2T_CHECK_ARG(smem <= 220 * 1024, "b2t_attention_bwd: T=%d does not fit the single-tile kernel", T);
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        attr = true;
    }
    attn_bwd_kernel<<<B * H, ATT_THREADS, smem, stream>>>(qkv, ld, q_off, k_off, v_off, probs, dctx, ld_ctx, ds_ws, dqkv,
                                                          (bf16*)dqkv_hi, (bf16*)dqkv_lo, T, H, d, scale, drop_p,
                                                          (unsigned long long)seed);
    B2T_CHECK_LAUNCH("b2t_attention_bwd");
    return 0;
}
