// Kernels either side of the model step (SURVEY.md §8f "next" rows):
//   n1  multi-tensor Adam (L2 weight-decay form) + global-norm gradient clipping
//         src/experiments/experiment.py:25-28, src/experiments/b2t_gru_w2v_experiment.py:138-145,
//         src/train/train_loop.py:72-78
//   n3  on-device WER / CER counts of the greedy CTC decode
//         src/train/evaluator.py:108-120,163-229 (+ transformers tokenization_wav2vec2.py:296-345 decode rules)
//   n4  input pipeline: block-wise z-scoring and batch collation of device-resident trials
//         src/datasets/preprocessing.py:60-72,100-113, src/datasets/brain2text.py:178-211
// All HBM-bound streaming work: float4 accesses, grids of one CTA per 64K-element chunk (hundreds to thousands of
// CTAs), no host synchronisation anywhere (norms, clip coefficients and error counts stay on the device).
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

constexpr int MT_CHUNK = B2T_MT_CHUNK;   // elements per CTA
constexpr int MT_THREADS = 256;

struct MtTensor {          // mirror of b2t_mt_tensor
    float* p;
    float* g;
    float* m;
    float* v;
    long n;
    int group;
    int pad;
};
static_assert(sizeof(MtTensor) == sizeof(b2t_mt_tensor), "b2t_mt_tensor layout");

struct AdamGroups {
    b2t_adam_group g[B2T_ADAM_MAX_GROUPS];
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const b2t_adam_group& h, float gs) {
    // torch.optim.Adam (single-tensor path, amsgrad=False, maximize=False), same operation order
    g *= gs;
    g = fmaf(h.weight_decay, p, g);                 // grad.add(param, alpha=weight_decay)
    m = m + (g - m) * h.one_minus_beta1;            // exp_avg.lerp_(grad, 1 - beta1)
    v = v * h.beta2;                                // exp_avg_sq.mul_(beta2)
    v = fmaf(h.one_minus_beta2 * g, g, v);          //           .addcmul_(grad, grad, value=1 - beta2)
    const float denom = sqrtf(v) / h.bias_c2_sqrt + h.eps;
    p = p - h.step_size * (m / denom);              // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(MT_THREADS) adam_multi_kernel(const MtTensor* __restrict__ tensors,
                                                                const int2* __restrict__ chunks, AdamGroups groups,
                                                                const float* __restrict__ grad_scale) {
    const int2 c = chunks[blockIdx.x];
    const MtTensor t = tensors[c.x];
    const b2t_adam_group h = groups.g[t.group];
    const float gs = grad_scale ? *grad_scale : 1.0f;
    const long off = (long)c.y * MT_CHUNK;
    const int len = (int)min((long)MT_CHUNK, t.n - off);
    float* p = t.p + off;
    const float* g = t.g + off;
    float* m = t.m + off;
    float* v = t.v + off;
    const bool vec = ((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0);
    int done = 0;
    if (vec) {
        const int n4 = len >> 2;
        // four independent 16-byte loads per array in flight per thread
        for (int i = threadIdx.x; i < n4; i += MT_THREADS) {
            float4 P = reinterpret_cast<float4*>(p)[i];
            const float4 G = reinterpret_cast<const float4*>(g)[i];
            float4 M = reinterpret_cast<float4*>(m)[i];
            float4 V = reinterpret_cast<float4*>(v)[i];
            adam_one(P.x, G.x, M.x, V.x, h, gs);
            adam_one(P.y, G.y, M.y, V.y, h, gs);
            adam_one(P.z, G.z, M.z, V.z, h, gs);
            adam_one(P.w, G.w, M.w, V.w, h, gs);
            reinterpret_cast<float4*>(p)[i] = P;
            reinterpret_cast<float4*>(m)[i] = M;
            reinterpret_cast<float4*>(v)[i] = V;
        }
        done = n4 << 2;
    }
    for (int i = done + threadIdx.x; i < len; i += MT_THREADS) {
        float P = p[i], M = m[i], V = v[i];
        adam_one(P, g[i], M, V, h, gs);
        p[i] = P;
        m[i] = M;
        v[i] = V;
    }
}

__global__ void __launch_bounds__(MT_THREADS) sumsq_multi_kernel(const MtTensor* __restrict__ tensors,
                                                                 const int2* __restrict__ chunks, double* __restrict__ out) {
    const int2 c = chunks[blockIdx.x];
    const MtTensor t = tensors[c.x];
    const long off = (long)c.y * MT_CHUNK;
    const int len = (int)min((long)MT_CHUNK, t.n - off);
    const float* g = t.g + off;
    float acc = 0.0f;
    int done = 0;
    if ((((uintptr_t)g) & 15) == 0) {
        const int n4 = len >> 2;
        for (int i = threadIdx.x; i < n4; i += MT_THREADS) {
            const float4 G = reinterpret_cast<const float4*>(g)[i];
            acc += G.x * G.x + G.y * G.y + G.z * G.z + G.w * G.w;
        }
        done = n4 << 2;
    }
    for (int i = done + threadIdx.x; i < len; i += MT_THREADS) acc += g[i] * g[i];
    acc = warp_sum(acc);
    __shared__ float part[MT_THREADS / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < MT_THREADS / 32; ++w) s += (double)part[w];
        atomicAdd(out, s);
    }
}

// torch.nn.utils.clip_grad_norm_: coef = max_norm / (total_norm + 1e-6), clamped to 1, grads *= coef
__global__ void __launch_bounds__(MT_THREADS) clip_scale_multi_kernel(const MtTensor* __restrict__ tensors,
                                                                      const int2* __restrict__ chunks,
                                                                      const double* __restrict__ sumsq, float max_norm,
                                                                      float* __restrict__ norm_out) {
    const float total = (float)sqrt(*sumsq);
    if (blockIdx.x == 0 && threadIdx.x == 0 && norm_out) *norm_out = total;
    const float coef = max_norm / (total + 1e-6f);
    if (!(coef < 1.0f)) return;      // clamped coefficient is 1: leave the gradients untouched (no write traffic)
    const int2 c = chunks[blockIdx.x];
    const MtTensor t = tensors[c.x];
    const long off = (long)c.y * MT_CHUNK;
    const int len = (int)min((long)MT_CHUNK, t.n - off);
    float* g = t.g + off;
    int done = 0;
    if ((((uintptr_t)g) & 15) == 0) {
        const int n4 = len >> 2;
        for (int i = threadIdx.x; i < n4; i += MT_THREADS) {
            float4 G = reinterpret_cast<float4*>(g)[i];
            G.x *= coef; G.y *= coef; G.z *= coef; G.w *= coef;
            reinterpret_cast<float4*>(g)[i] = G;
        }
        done = n4 << 2;
    }
    for (int i = done + threadIdx.x; i < len; i += MT_THREADS) g[i] *= coef;
}

// =============================================================================================================
// WER / CER counts of the greedy decode.  One warp per trial.
//   prediction string: tokenizer.batch_decode(argmax ids, group_tokens=True)   evaluator.py:111-114
//   label string:      tokenizer.batch_decode(target ids, group_tokens=False)  evaluator.py:115-119
// decode rules (tokenization_wav2vec2.py:296-345 + clean_up_tokenization): group repeats, drop <pad>, word
// delimiter -> ' ', join, strip, then every " ' " -> "'" (the only clean-up rule an upper-case vocabulary can hit);
// optionally cut the prediction after the first "</s>" (evaluator.py:169-181).  CER = Levenshtein over characters /
// target characters (edit_distance.SequenceMatcher.distance), WER = Levenshtein over whitespace-split words / target
// words (torcheval WordErrorRate).  counts[4] += {char_dist, target_chars, word_dist, target_words}.
// =============================================================================================================
struct Seq {
    unsigned char* ch;    // characters
    int n;
};

__device__ int build_string(unsigned char* out, const int* ids, int n_ids, bool group, const unsigned char* vocab, int V,
                            int pad_id, int delim_id, int eos_id, int cleanup) {
    // serial (lane 0): <= a few hundred tokens
    int n = 0, prev = -1;
    for (int i = 0; i < n_ids; ++i) {
        const int id = ids[i];
        if (group && id == prev) continue;
        prev = id;
        if (id == pad_id || id < 0 || id >= V) continue;
        if (id == delim_id) {
            out[n++] = ' ';
            continue;
        }
        const unsigned char* e = vocab + id * 8;
        const int len = e[0];
        for (int k = 0; k < len; ++k) out[n++] = e[1 + k];
    }
    // strip
    int a = 0, b = n;
    while (a < b && out[a] == ' ') ++a;
    while (b > a && out[b - 1] == ' ') --b;
    int m = 0;
    if (cleanup) {
        // str.replace(" ' ", "'"): left to right, non-overlapping
        int i = a;
        while (i < b) {
            if (i + 2 < b && out[i] == ' ' && out[i + 1] == '\'' && out[i + 2] == ' ') {
                out[m++] = '\'';
                i += 3;
            } else {
                out[m++] = out[i++];
            }
        }
    } else {
        for (int i = a; i < b; ++i) out[m++] = out[i];
    }
    if (eos_id >= 0 && eos_id < V) {
        // evaluator.py:169-176 cut_after_eos_token: keep up to and including the first occurrence of the eos string
        const unsigned char* e = vocab + eos_id * 8;
        const int len = e[0];
        for (int k = 0; len > 0 && k + len <= m; ++k) {
            bool hit = true;
            for (int q = 0; q < len; ++q) hit = hit && (out[k + q] == e[1 + q]);
            if (hit) {
                m = k + len;
                break;
            }
        }
    }
    return m;
}

__device__ int split_words(const unsigned char* s, int n, unsigned long long* words) {
    // str.split(): maximal runs of non-blank characters; a word is represented by a 64-bit FNV-1a hash + length
    int nw = 0, i = 0;
    while (i < n) {
        while (i < n && s[i] == ' ') ++i;
        if (i >= n) break;
        unsigned long long h = 1469598103934665603ull;
        int len = 0;
        while (i < n && s[i] != ' ') {
            h = (h ^ s[i]) * 1099511628211ull;
            ++i;
            ++len;
        }
        words[nw++] = h ^ ((unsigned long long)len << 56);
    }
    return nw;
}

// Levenshtein distance, one warp: rows = a (length na), columns = b (length nb); prev/cur have nb+1 ints.
template <typename Tel>
__device__ int levenshtein_warp(const Tel* a, int na, const Tel* b, int nb, int* prev, int* cur) {
    const int lane = threadIdx.x & 31;
    for (int j = lane; j <= nb; j += 32) prev[j] = j;
    __syncwarp();
    for (int i = 1; i <= na; ++i) {
        const Tel ai = a[i - 1];
        // cur[j] = min(t[j], cur[j-1] + 1) with t[j] = min(prev[j] + 1, prev[j-1] + (a_i != b_j)), cur[0] = i
        //        = j + min_{k <= j} (t[k] - k)                                   (t[0] = i)
        int carry = i;          // min over the columns of the previous chunks of (t[k] - k)
        if (lane == 0) cur[0] = i;
        for (int j0 = 1; j0 <= nb; j0 += 32) {
            const int j = j0 + lane;
            int t = 0x3fffffff;
            if (j <= nb) t = min(prev[j] + 1, prev[j - 1] + (ai != b[j - 1] ? 1 : 0)) - j;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t = min(t, u);
            }
            t = min(t, carry);
            if (j <= nb) cur[j] = t + j;
            carry = __shfl_sync(0xffffffffu, t, 31);
        }
        __syncwarp();
        int* tmp = prev;
        prev = cur;
        cur = tmp;
    }
    return prev[nb];
}

__global__ void __launch_bounds__(32) ctc_error_counts_kernel(const int* __restrict__ pred_ids, int T, long ld_pred,
                                                              const long long* __restrict__ targets, int S,
                                                              const unsigned char* __restrict__ vocab, int V, int pad_id,
                                                              int delim_id, int eos_id, int cleanup,
                                                              unsigned long long* __restrict__ counts,
                                                              int* __restrict__ per_trial, int max_pred_chars,
                                                              int max_tgt_chars) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int b = blockIdx.x;
    const int lane = threadIdx.x;
    // layout: words_p [T] u64 | words_t [S] u64 | dp rows 2 x (max(max_tgt_chars, S) + 1) int | ids [max(T,S)] int |
    //         pred chars | target chars
    unsigned long long* wp = reinterpret_cast<unsigned long long*>(smem);
    unsigned long long* wt = wp + (max_pred_chars / 2 + 1);
    const int dpw = max_tgt_chars + 1;
    int* dp0 = reinterpret_cast<int*>(wt + (max_tgt_chars / 2 + 1));
    int* dp1 = dp0 + dpw;
    int* ids = dp1 + dpw;
    unsigned char* sp = reinterpret_cast<unsigned char*>(ids + max(T, S));
    unsigned char* st = sp + max_pred_chars;
    __shared__ int n_sh[4];
    for (int i = lane; i < T; i += 32) ids[i] = pred_ids[(long)b * ld_pred + i];
    __syncwarp();
    if (lane == 0) n_sh[0] = build_string(sp, ids, T, true, vocab, V, pad_id, delim_id, eos_id, cleanup);
    __syncwarp();
    for (int i = lane; i < S; i += 32) ids[i] = (int)targets[(long)b * S + i];
    __syncwarp();
    if (lane == 0) {
        n_sh[1] = build_string(st, ids, S, false, vocab, V, pad_id, delim_id, -1, cleanup);
        n_sh[2] = split_words(sp, n_sh[0], wp);
        n_sh[3] = split_words(st, n_sh[1], wt);
    }
    __syncwarp();
    const int np = n_sh[0], nt = n_sh[1], nwp = n_sh[2], nwt = n_sh[3];
    const int cd = levenshtein_warp<unsigned char>(sp, np, st, nt, dp0, dp1);
    __syncwarp();
    const int wd = levenshtein_warp<unsigned long long>(wp, nwp, wt, nwt, dp0, dp1);
    if (lane == 0) {
        atomicAdd(counts + 0, (unsigned long long)cd);
        atomicAdd(counts + 1, (unsigned long long)nt);
        atomicAdd(counts + 2, (unsigned long long)wd);
        atomicAdd(counts + 3, (unsigned long long)nwt);
        if (per_trial) {
            per_trial[b * 4 + 0] = cd;
            per_trial[b * 4 + 1] = nt;
            per_trial[b * 4 + 2] = wd;
            per_trial[b * 4 + 3] = nwt;
        }
    }
}

// =============================================================================================================
// Input pipeline: trials live device-resident as one flat [sum_T, C] fp32 array with offsets.
// =============================================================================================================
// pass 1: ws[block][c][0] += sum x, ws[block][c][2] += frames        pass 2: ws[block][c][1] += sum (x - mean)^2
template <int PASS>
__global__ void __launch_bounds__(256) block_stats_kernel(const float* __restrict__ x, const long* __restrict__ offs,
                                                          const int* __restrict__ trial_block, int C, double* __restrict__ ws) {
    const int trial = blockIdx.x;
    const int blk = trial_block[trial];
    if (blk < 0) return;
    const long f0 = offs[trial], f1 = offs[trial + 1];
    // thread = channel (strided), rows split over blockIdx.y
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        double* w = ws + ((long)blk * C + c) * 3;
        double mean = 0.0;
        if (PASS == 2) mean = w[0] / w[2];
        double acc = 0.0;
        for (long f = f0 + blockIdx.y; f < f1; f += gridDim.y) {
            const double v = (double)x[f * C + c];
            if (PASS == 1) acc += v;
            else acc += (v - mean) * (v - mean);
        }
        if (PASS == 1) {
            atomicAdd(w + 0, acc);
            if (blockIdx.y == 0) atomicAdd(w + 2, (double)(f1 - f0));
        } else {
            atomicAdd(w + 1, acc);
        }
    }
}

// (x - mean) / (std + eps), std = population standard deviation (np.std), computed as numpy does in float64 and
// rounded to float32 at the end (torch.tensor(..., dtype=float32), brain2text.py:131)
__global__ void __launch_bounds__(256) block_zscore_apply_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                 const long* __restrict__ offs,
                                                                 const int* __restrict__ trial_block, int C,
                                                                 const double* __restrict__ ws, double eps) {
    const int trial = blockIdx.x;
    const int blk = trial_block[trial];
    const long f0 = offs[trial], f1 = offs[trial + 1];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        double mean = 0.0, inv = 1.0;
        if (blk >= 0) {
            const double* w = ws + ((long)blk * C + c) * 3;
            mean = w[0] / w[2];
            inv = 1.0 / (sqrt(w[1] / w[2]) + eps);
        }
        for (long f = f0 + blockIdx.y; f < f1; f += gridDim.y) y[f * C + c] = (float)(((double)x[f * C + c] - mean) * inv);
    }
}

// out[i, t, :] = frames[offs[idx[i]] + t, :] for t < len, 0 up to t_max; small per-trial fields gathered by the
// first CTA row of each trial (brain2text.py:178-211)
__global__ void __launch_bounds__(256) collate_kernel(const float4* __restrict__ frames, const long* __restrict__ offs,
                                                      const int* __restrict__ idx, float4* __restrict__ out, int t_max,
                                                      int c4, const long long* __restrict__ labels, int s_all,
                                                      const long long* __restrict__ label_lens,
                                                      const long long* __restrict__ days, long long* __restrict__ out_tgt,
                                                      int s_out, long long* __restrict__ out_in_lens,
                                                      long long* __restrict__ out_tgt_lens, long long* __restrict__ out_days) {
    const int i = blockIdx.y;
    const int trial = idx[i];
    const long f0 = offs[trial];
    const int len = (int)min((long)t_max, offs[trial + 1] - f0);
    const long n4 = (long)t_max * c4;
    const long lim = (long)len * c4;
    const float4* src = frames + f0 * c4;
    float4* dst = out + (long)i * n4;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < n4; e += (long)gridDim.x * blockDim.x)
        dst[e] = e < lim ? src[e] : z;
    if (blockIdx.x == 0) {
        for (int s = threadIdx.x; s < s_out; s += blockDim.x)
            out_tgt[(long)i * s_out + s] = s < s_all ? labels[(long)trial * s_all + s] : 0;
        if (threadIdx.x == 0) {
            out_in_lens[i] = len;
            out_tgt_lens[i] = label_lens[trial];
            out_days[i] = days[trial];
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------
extern "C" int b2t_adam_multi(const b2t_mt_tensor* tensors, const int32_t* chunks, int n_chunks, const b2t_adam_group* groups,
                              int n_groups, const float* grad_scale, b2t_stream stream) {
    B2T_CHECK_ARG(n_groups >= 1 && n_groups <= B2T_ADAM_MAX_GROUPS, "b2t_adam_multi: 1..%d parameter groups, got %d",
                  B2T_ADAM_MAX_GROUPS, n_groups);
    B2T_CHECK_ARG(n_chunks >= 0, "b2t_adam_multi: n_chunks < 0");
    if (n_chunks == 0) return B2T_OK;
    B2T_CHECK_ARG(tensors && chunks && groups, "b2t_adam_multi: null table");
    AdamGroups g;
    memset(&g, 0, sizeof(g));
    for (int i = 0; i < n_groups; ++i) {
        g.g[i] = groups[i];
        B2T_CHECK_ARG(groups[i].bias_c2_sqrt > 0.f, "b2t_adam_multi: group %d: step must be >= 1", i);
    }
    adam_multi_kernel<<<n_chunks, MT_THREADS, 0, (cudaStream_t)stream>>>(reinterpret_cast<const MtTensor*>(tensors),
                                                                         reinterpret_cast<const int2*>(chunks), g, grad_scale);
    B2T_CHECK_LAUNCH("b2t_adam_multi");
    return B2T_OK;
}

extern "C" int b2t_grad_sumsq_multi(const b2t_mt_tensor* tensors, const int32_t* chunks, int n_chunks, double* sumsq,
                                    b2t_stream stream) {
    B2T_CHECK_ARG(sumsq, "b2t_grad_sumsq_multi: null output");
    if (n_chunks <= 0) return B2T_OK;
    sumsq_multi_kernel<<<n_chunks, MT_THREADS, 0, (cudaStream_t)stream>>>(reinterpret_cast<const MtTensor*>(tensors),
                                                                          reinterpret_cast<const int2*>(chunks), sumsq);
    B2T_CHECK_LAUNCH("b2t_grad_sumsq_multi");
    return B2T_OK;
}

extern "C" int b2t_grad_clip_multi(const b2t_mt_tensor* tensors, const int32_t* chunks, int n_chunks, const double* sumsq,
                                   float max_norm, float* total_norm_out, b2t_stream stream) {
    B2T_CHECK_ARG(sumsq && max_norm > 0.f, "b2t_grad_clip_multi: bad arguments");
    if (n_chunks <= 0) return B2T_OK;
    clip_scale_multi_kernel<<<n_chunks, MT_THREADS, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const MtTensor*>(tensors), reinterpret_cast<const int2*>(chunks), sumsq, max_norm, total_norm_out);
    B2T_CHECK_LAUNCH("b2t_grad_clip_multi");
    return B2T_OK;
}

extern "C" int b2t_ctc_error_counts(const int32_t* pred_ids, int T, int64_t ld_pred, const int64_t* targets, int S,
                                    const uint8_t* vocab_chars, int V, int pad_id, int delim_id, int eos_id, int cleanup,
                                    int64_t* counts, int32_t* per_trial, int B, b2t_stream stream) {
    B2T_CHECK_ARG(B >= 0 && T >= 0 && S >= 0 && V > 0 && V <= 256, "b2t_ctc_error_counts: bad sizes B=%d T=%d S=%d V=%d", B, T, S, V);
    if (B == 0) return B2T_OK;
    B2T_CHECK_ARG(pred_ids && targets && vocab_chars && counts, "b2t_ctc_error_counts: null pointer");
    const int max_pred = T * 7 + 8, max_tgt = S * 7 + 8;
    const int tmax = T > S ? T : S;
    size_t smem = (size_t)(max_pred / 2 + 1) * 8 + (size_t)(max_tgt / 2 + 1) * 8 + (size_t)2 * (max_tgt + 1) * 4 + (size_t)tmax * 4 +
                  max_pred + max_tgt + 16;
    B2T_CHECK_ARG(smem <= 200 * 1024, "b2t_ctc_error_counts: T=%d / S=%d too long for one CTA's shared memory", T, S);
    static bool attr = false;
    if (!attr || smem > 48 * 1024) {
        B2T_CUDA(cudaFuncSetAttribute(ctc_error_counts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr = true;
    }
    ctc_error_counts_kernel<<<B, 32, smem, (cudaStream_t)stream>>>(
        pred_ids, T, ld_pred, reinterpret_cast<const long long*>(targets), S, vocab_chars, V, pad_id, delim_id, eos_id, cleanup,
        reinterpret_cast<unsigned long long*>(counts), per_trial, max_pred, max_tgt);
    B2T_CHECK_LAUNCH("b2t_ctc_error_counts");
    return B2T_OK;
}

extern "C" int b2t_block_zscore(const float* x, float* y, const int64_t* trial_offsets, const int32_t* trial_block,
                                int n_trials, int n_blocks, int C, double* ws, double eps, b2t_stream stream) {
    B2T_CHECK_ARG(n_trials >= 0 && n_blocks >= 0 && C > 0, "b2t_block_zscore: bad sizes");
    if (n_trials == 0) return B2T_OK;
    B2T_CHECK_ARG(x && y && trial_offsets && trial_block && ws, "b2t_block_zscore: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    B2T_CUDA(cudaMemsetAsync(ws, 0, sizeof(double) * 3 * (size_t)n_blocks * C, s));
    dim3 grid(n_trials, 8);
    block_stats_kernel<1><<<grid, 256, 0, s>>>(x, reinterpret_cast<const long*>(trial_offsets), trial_block, C, ws);
    block_stats_kernel<2><<<grid, 256, 0, s>>>(x, reinterpret_cast<const long*>(trial_offsets), trial_block, C, ws);
    block_zscore_apply_kernel<<<grid, 256, 0, s>>>(x, y, reinterpret_cast<const long*>(trial_offsets), trial_block, C, ws, eps);
    B2T_CHECK_LAUNCH("b2t_block_zscore");
    return B2T_OK;
}

extern "C" int b2t_collate(const float* frames, const int64_t* trial_offsets, const int32_t* idx, int B, int C, int t_max,
                           float* out, const int64_t* labels, int s_all, const int64_t* label_lens, const int64_t* days,
                           int64_t* out_targets, int s_out, int64_t* out_input_lens, int64_t* out_target_lens,
                           int64_t* out_days, b2t_stream stream) {
    B2T_CHECK_ARG(B >= 0 && C > 0 && C % 4 == 0 && t_max >= 0, "b2t_collate: C must be a multiple of 4 (C=%d)", C);
    if (B == 0) return B2T_OK;
    B2T_CHECK_ARG(frames && trial_offsets && idx && out && labels && label_lens && days && out_targets && out_input_lens &&
                      out_target_lens && out_days, "b2t_collate: null pointer");
    const long n4 = (long)t_max * (C / 4);
    int gx = (int)((n4 + 256 * 4 - 1) / (256 * 4));
    if (gx < 1) gx = 1;
    if (gx > 64) gx = 64;
    dim3 grid(gx, B);
    collate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(frames), reinterpret_cast<const long*>(trial_offsets), idx,
        reinterpret_cast<float4*>(out), t_max, C / 4, reinterpret_cast<const long long*>(labels), s_all,
        reinterpret_cast<const long long*>(label_lens), reinterpret_cast<const long long*>(days),
        reinterpret_cast<long long*>(out_targets), s_out, reinterpret_cast<long long*>(out_input_lens),
        reinterpret_cast<long long*>(out_target_lens), reinterpret_cast<long long*>(out_days));
    B2T_CHECK_LAUNCH("b2t_collate");
    return B2T_OK;
}
