// CTC loss forward+backward and greedy decode, one warp per trial.
// Replaces log_softmax + torch.nn.CTCLoss(blank=0, reduction="mean", zero_infinity=True)
// (src/model/w2v_custom_feat_extractor.py:59,81-90) and logits.argmax(-1) + tokenizer collapse
// (src/train/evaluator.py:108-120).
//
// vocab (32) == warp width: lane v owns vocabulary entry v for the log-softmax and for the
// per-class occupancy; the extended label lattice (2S+1 states) is spread KS states per lane and
// the alpha/beta recursions exchange their one/two-state halos with warp shuffles, so the T'
// dependent steps run without any block barrier.  Emits loss and d(loss)/d(logits) directly.
#include "common.cuh"
#include "../../include/b2t_sm100.h"
#include <math_constants.h>

namespace {

__device__ __forceinline__ float lse3(float a, float b, float c) {
    const float m = fmaxf(a, fmaxf(b, c));
    if (m == -CUDART_INF_F) return -CUDART_INF_F;
    return m + logf(expf(a - m) + expf(b - m) + expf(c - m));
}

template <int KS>
__global__ void __launch_bounds__(32) ctc_kernel(const float* __restrict__ logits, long ld_t, long ld_b,
                                                 const long long* __restrict__ targets, int s_max,
                                                 const int* __restrict__ input_lens, const long long* __restrict__ target_lens,
                                                 float* __restrict__ alpha_ws, float* __restrict__ nll_out,
                                                 float* __restrict__ loss_out, float* __restrict__ grad, int B, int T, int V,
                                                 int zero_infinity) {
    extern __shared__ float smem[];
    float* lp = smem;              // [T][32]
    float* occ = smem + T * 32;    // [32]
    const int b = blockIdx.x;
    const int lane = threadIdx.x;
    const float NEG = -CUDART_INF_F;
    int Tlen = input_lens[b];
    Tlen = Tlen < 0 ? 0 : (Tlen > T ? T : Tlen);
    long long S_ll = target_lens[b];
    const int S = (int)(S_ll < 0 ? 0 : (S_ll > s_max ? s_max : S_ll));
    const int L = 2 * S + 1;
    const float* lg = logits + (long)b * ld_b;

    // log-softmax of every frame (lane = vocabulary entry)
    for (int t = 0; t < T; ++t) {
        const float x = lane < V ? lg[(long)t * ld_t + lane] : NEG;
        const float m = warp_max(x);
        const float e = lane < V ? expf(x - m) : 0.0f;
        const float lse = m + logf(warp_sum(e));
        lp[t * 32 + lane] = x - lse;
    }
    __syncwarp();

    // lattice bookkeeping for this lane's KS states
    int lab[KS];
    bool skip_a[KS], skip_b[KS], valid[KS];
    bool bad_label = false;
#pragma unroll
    for (int i = 0; i < KS; ++i) {
        const int s = lane * KS + i;
        valid[i] = s < L;
        const int tgt_i = (s - 1) >> 1;
        lab[i] = (valid[i] && (s & 1)) ? (int)targets[(long)b * s_max + tgt_i] : 0;
        // a label outside the vocabulary (wrong tokenizer, a -100 padding value inside target_len) would index the
        // shared-memory log-prob / occupancy tables out of bounds: clamp it for memory safety and treat the trial as
        // infeasible below (ATen's CTCLoss raises here; a kernel cannot, so the trial reports an infinite loss — 0 under
        // zero_infinity — and gets no gradient)
        if (lab[i] < 0 || lab[i] >= V) {
            bad_label = true;
            lab[i] = 0;
        }
        // transitions s-2 -> s and s -> s+2 are allowed between different non-blank labels
        skip_a[i] = valid[i] && (s & 1) && s >= 3 && targets[(long)b * s_max + tgt_i] != targets[(long)b * s_max + tgt_i - 1];
        skip_b[i] = valid[i] && (s & 1) && (s + 2 < L) &&
                    targets[(long)b * s_max + tgt_i] != targets[(long)b * s_max + tgt_i + 1];
    }

    float* aw = alpha_ws + (long)b * T * (32 * KS);
    float a[KS];
    float nll = CUDART_INF_F;
    if (Tlen > 0) {
#pragma unroll
        for (int i = 0; i < KS; ++i) {
            const int s = lane * KS + i;
            a[i] = (valid[i] && s < 2) ? lp[lab[i]] : NEG;
            aw[lane * KS + i] = a[i];
        }
        for (int t = 1; t < Tlen; ++t) {
            const float up1 = __shfl_up_sync(0xffffffffu, a[KS - 1], 1);
            const float up2 = __shfl_up_sync(0xffffffffu, KS >= 2 ? a[KS >= 2 ? KS - 2 : 0] : 0.0f, 1);
            float na[KS];
#pragma unroll
            for (int i = 0; i < KS; ++i) {
                float p1, p2;
                if (i >= 1) p1 = a[i - 1];
                else p1 = lane > 0 ? up1 : NEG;
                if (i >= 2) p2 = a[i - 2];
                else if (i == 1) p2 = lane > 0 ? up1 : NEG;
                else p2 = lane > 0 ? up2 : NEG;
                if (!skip_a[i]) p2 = NEG;
                na[i] = valid[i] ? lse3(a[i], p1, p2) + lp[t * 32 + lab[i]] : NEG;
            }
#pragma unroll
            for (int i = 0; i < KS; ++i) {
                a[i] = na[i];
                aw[(long)t * (32 * KS) + lane * KS + i] = a[i];
            }
        }
        // log-likelihood = lse(alpha_{T-1}(L-1), alpha_{T-1}(L-2))
        float mine = NEG;
#pragma unroll
        for (int i = 0; i < KS; ++i) {
            const int s = lane * KS + i;
            if (valid[i] && (s == L - 1 || s == L - 2)) mine = lse3(mine, a[i], NEG);
        }
        const float m = warp_max(mine);
        float ll = NEG;
        if (m != NEG) ll = m + logf(warp_sum(mine == NEG ? 0.0f : expf(mine - m)));
        nll = -ll;
    }
    if (__any_sync(0xffffffffu, bad_label)) nll = CUDART_INF_F;
    const bool infinite = !(nll < CUDART_INF_F);   // inf or nan
    const float denom = (float)(S > 0 ? S : 1) * (float)B;
    if (lane == 0) {
        const float reported = (infinite && zero_infinity) ? 0.0f : nll;
        if (nll_out) nll_out[b] = reported;
        if (loss_out) atomicAdd(loss_out, reported / denom);
    }
    if (!grad) return;
    float* gb = grad + (long)b * T * V;
    // frames past the input length (and infinite losses) get zero gradient
    for (int t = (infinite && zero_infinity) ? 0 : Tlen; t < T; ++t)
        if (lane < V) gb[(long)t * V + lane] = 0.0f;
    if (Tlen == 0 || (infinite && zero_infinity)) return;

    const float gscale = 1.0f / denom;
    float bt[KS];
    for (int t = Tlen - 1; t >= 0; --t) {
        if (t == Tlen - 1) {
#pragma unroll
            for (int i = 0; i < KS; ++i) {
                const int s = lane * KS + i;
                bt[i] = (valid[i] && (s == L - 1 || s == L - 2)) ? lp[t * 32 + lab[i]] : NEG;
            }
        } else {
            const float dn1 = __shfl_down_sync(0xffffffffu, bt[0], 1);
            const float dn2 = __shfl_down_sync(0xffffffffu, KS >= 2 ? bt[KS >= 2 ? 1 : 0] : 0.0f, 1);
            float nb[KS];
#pragma unroll
            for (int i = 0; i < KS; ++i) {
                float n1, n2;
                if (i + 1 < KS) n1 = bt[i + 1];
                else n1 = lane < 31 ? dn1 : NEG;
                if (i + 2 < KS) n2 = bt[i + 2];
                else if (i + 2 == KS) n2 = lane < 31 ? dn1 : NEG;
                else n2 = lane < 31 ? dn2 : NEG;
                if (!skip_b[i]) n2 = NEG;
                nb[i] = valid[i] ? lse3(bt[i], n1, n2) + lp[t * 32 + lab[i]] : NEG;
            }
#pragma unroll
            for (int i = 0; i < KS; ++i) bt[i] = nb[i];
        }
        // occupancy per class: sum over states with that label of exp(alpha + beta)
        float w[KS];
        float wm = NEG;
#pragma unroll
        for (int i = 0; i < KS; ++i) {
            w[i] = valid[i] ? aw[(long)t * (32 * KS) + lane * KS + i] + bt[i] : NEG;
            wm = fmaxf(wm, w[i]);
        }
        wm = warp_max(wm);
        occ[lane] = 0.0f;
        __syncwarp();
        if (wm != NEG) {
#pragma unroll
            for (int i = 0; i < KS; ++i)
                if (w[i] != NEG) atomicAdd(&occ[lab[i]], expf(w[i] - wm));
        }
        __syncwarp();
        if (lane < V) {
            const float l = lp[t * 32 + lane];
            float g = expf(l);
            if (wm != NEG && occ[lane] > 0.0f) g -= occ[lane] * expf(wm + nll - l);
            gb[(long)t * V + lane] = g * gscale;
        }
        __syncwarp();
    }
}

// argmax over the vocabulary for every frame, then collapse repeats and drop blanks (id 0)
__global__ void __launch_bounds__(32) greedy_kernel(const float* __restrict__ logits, long ld_t, long ld_b,
                                                    int* __restrict__ argmax_out, int* __restrict__ ids_out,
                                                    int* __restrict__ lens_out, int T, int V) {
    const int b = blockIdx.x, lane = threadIdx.x;
    const float* lg = logits + (long)b * ld_b;
    int prev = -1, count = 0;
    for (int t = 0; t < T; ++t) {
        float v = lane < V ? lg[(long)t * ld_t + lane] : -CUDART_INF_F;
        int idx = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, v, o);
            const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
            // NaN-free inputs: larger value wins, ties go to the lower index (torch.argmax)
            if (ov > v || (ov == v && oi < idx)) {
                v = ov;
                idx = oi;
            }
        }
        if (lane == 0) {
            if (argmax_out) argmax_out[(long)b * T + t] = idx;
            if (idx != prev && idx != 0 && ids_out) ids_out[(long)b * T + count++] = idx;
            prev = idx;
        }
    }
    if (lane == 0) {
        if (ids_out)
            for (int i = count; i < T; ++i) ids_out[(long)b * T + i] = 0;
        if (lens_out) lens_out[b] = count;
    }
}

}  // namespace

extern "C" int64_t b2t_ctc_workspace_floats(int B, int T, int s_max) {
    const int L = 2 * s_max + 1;
    int ks = 4;
    while (32 * ks < L) ks *= 2;
    return (int64_t)B * T * 32 * ks;
}

extern "C" int b2t_ctc_fwd_bwd(const float* logits, int64_t ld_t, int64_t ld_b, const int64_t* targets, int s_max,
                               const int32_t* input_lens, const int64_t* target_lens, float* alpha_ws, float* nll,
                               float* loss, float* grad, int B, int T, int V, int zero_infinity, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(logits && targets && input_lens && target_lens && alpha_ws, "b2t_ctc_fwd_bwd: null pointer");
    B2T_CHECK_ARG(B > 0 && T > 0 && V > 0 && V <= 32 && s_max >= 0, "b2t_ctc_fwd_bwd: bad sizes B=%d T=%d V=%d S=%d", B, T, V,
                  s_max);
    const int L = 2 * s_max + 1;
    B2T_CHECK_ARG(L <= 32 * 32, "b2t_ctc_fwd_bwd: target length %d too long (max 511)", s_max);
    const size_t smem = ((size_t)T * 32 + 32) * sizeof(float);
    B2T_CHECK_ARG(smem <= 200 * 1024, "b2t_ctc_fwd_bwd: T=%d too long for the shared-memory log-prob table", T);
#define LAUNCH(KS)                                                                                                       \
    do {                                                                                                                 \
        static bool attr = false;                                                                                        \
        if (!attr) {                                                                                                     \
            B2T_CUDA(cudaFuncSetAttribute(ctc_kernel<KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));      \
            attr = true;                                                                                                 \
        }                                                                                                                \
        ctc_kernel<KS><<<B, 32, smem, stream>>>(logits, ld_t, ld_b, (const long long*)targets, s_max, input_lens,        \
                                                (const long long*)target_lens, alpha_ws, nll, loss, grad, B, T, V,       \
                                                zero_infinity);                                                         \
    } while (0)
    if (L <= 128) LAUNCH(4);
    else if (L <= 256) LAUNCH(8);
    else if (L <= 512) LAUNCH(16);
    else LAUNCH(32);
#undef LAUNCH
    B2T_CHECK_LAUNCH("b2t_ctc_fwd_bwd");
    return 0;
}

// ------------------------------------------------------------------------------------------
// Allocation-free variant of b2t_ctc_fwd_bwd (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The alpha lattice lives in a file-scope device array instead of a caller-provided buffer.  Every GPU that loads the
// library holds these 8 MiB until exit.  All calls share the array: calls that overlap on the GPU (two streams, two
// host threads) race.  Nothing in it has to start at zero - the kernel writes alpha[t] of a trial before the beta pass
// reads it - so stale contents of the previous call are harmless.
// ------------------------------------------------------------------------------------------
#define B2T_CTC_STATIC_MAX_FLOATS (2 * 1024 * 1024)   /* = b2t_ctc_static_max_floats(); e.g. B=128, T=118, S<=63 */
static __device__ __align__(256) float g_ctc_static_ws[B2T_CTC_STATIC_MAX_FLOATS];

extern "C" int64_t b2t_ctc_static_max_floats(void) { return (int64_t)B2T_CTC_STATIC_MAX_FLOATS; }

extern "C" int b2t_ctc_fwd_bwd_static(const float* logits, int64_t ld_t, int64_t ld_b, const int64_t* targets, int s_max,
                                      const int32_t* input_lens, const int64_t* target_lens, float* nll, float* loss,
                                      float* grad, int B, int T, int V, int zero_infinity, b2t_stream stream_) {
    B2T_CHECK_ARG(B > 0 && T > 0 && s_max >= 0, "b2t_ctc_fwd_bwd_static: bad sizes B=%d T=%d S=%d", B, T, s_max);
    const int64_t need = b2t_ctc_workspace_floats(B, T, s_max);
    B2T_CHECK_ARG(need <= (int64_t)B2T_CTC_STATIC_MAX_FLOATS,
                  "b2t_ctc_fwd_bwd_static: B=%d T=%d S=%d needs %lld workspace floats, the static scratch holds %d "
                  "(use b2t_ctc_fwd_bwd with a caller-owned workspace)", B, T, s_max, (long long)need, B2T_CTC_STATIC_MAX_FLOATS);
    // The array's name is a host-side stand-in in host code, not a device address: resolve it once per device, with
    // that device current.
    static float* ws_of_device[64] = {};
    int dev = -1;
    B2T_CUDA(cudaGetDevice(&dev));
    B2T_CHECK_ARG(dev >= 0 && dev < 64, "b2t_ctc_fwd_bwd_static: device index %d out of range", dev);
    float* ws = ws_of_device[dev];
    if (!ws) {
        B2T_CUDA(cudaGetSymbolAddress((void**)&ws, g_ctc_static_ws));
        B2T_CHECK_ARG(ws != nullptr, "b2t_ctc_fwd_bwd_static: cudaGetSymbolAddress returned NULL");
        ws_of_device[dev] = ws;
    }
    return b2t_ctc_fwd_bwd(logits, ld_t, ld_b, targets, s_max, input_lens, target_lens, ws, nll, loss, grad, B, T, V,
                           zero_infinity, stream_);
}

extern "C" int b2t_greedy_decode(const float* logits, int64_t ld_t, int64_t ld_b, int32_t* argmax, int32_t* ids,
                                 int32_t* lens, int B, int T, int V, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(logits && (argmax || ids) && B > 0 && T > 0 && V > 0 && V <= 32, "b2t_greedy_decode: bad arguments");
    greedy_kernel<<<B, 32, 0, stream>>>(logits, ld_t, ld_b, argmax, ids, lens, T, V);
    B2T_CHECK_LAUNCH("b2t_greedy_decode");
    return 0;
}
