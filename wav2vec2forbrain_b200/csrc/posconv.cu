// Layout kernels around the positional convolution of wav2vec2
// (transformers modeling_wav2vec2.py:326-368: weight-normed grouped conv1d, k=128, groups=16,
// padding 64, last frame dropped, GELU).  The convolution itself runs on the tcgen05 GEMM as a
// per-group implicit GEMM: the input is re-laid out group-major with the group's channels padded
// to 64 (= one 128-byte swizzle row), so that frame t's receptive field [t, t+K) x 64 is ONE
// contiguous run of K*64 elements and the im2col is just an overlapping-stride TMA view.
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

constexpr int CP = 64;   // padded channels per group

// x [B,T,D] -> xg [G][B][Tp][CP] (bf16 hi/lo), xg[g][b][pad_left + t][i] = x[b][t][g*Cg + i], zeros elsewhere
__global__ void pack_input_kernel(const float* __restrict__ x, bf16* __restrict__ hi, bf16* __restrict__ lo, int B, int T,
                                  int D, int G, int Cg, int pad_left, int Tp) {
    const long n = (long)G * B * Tp * CP;
    for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long)gridDim.x * blockDim.x) {
        const int i = (int)(idx % CP);
        long r = idx / CP;
        const int tp = (int)(r % Tp);
        r /= Tp;
        const int b = (int)(r % B);
        const int g = (int)(r / B);
        const int t = tp - pad_left;
        float v = 0.0f;
        if (i < Cg && t >= 0 && t < T) v = x[((long)b * T + t) * D + g * Cg + i];
        bf16 h, l;
        split_bf16(v, h, l);
        hi[idx] = h;
        if (lo) lo[idx] = l;
    }
}

// norm[k] = || v[:, :, k] ||_2  (weight_norm dim=2: one norm per tap over (out, in))
__global__ void __launch_bounds__(256) tap_norm_kernel(const float* __restrict__ v, float* __restrict__ norm, long per_tap,
                                                       int K) {
    __shared__ float red[8];
    const int k = blockIdx.x;
    float s = 0.0f;
    for (long e = threadIdx.x; e < per_tap; e += blockDim.x) {
        const float w = v[e * K + k];
        s += w * w;
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int i = 0; i < 8; ++i) t += red[i];
        norm[k] = sqrtf(t);
    }
}

// wn[o][i][k] = v[o][i][k] * g[k] / norm[k]  ->  wg[g][o_local][k*CP + i] (zero for i >= Cg)
__global__ void pack_weight_kernel(const float* __restrict__ gk, const float* __restrict__ v, const float* __restrict__ norm,
                                   bf16* __restrict__ hi, bf16* __restrict__ lo, int D, int Cg, int K, int G) {
    const int Og = D / G;
    const long n = (long)D * K * CP;
    for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long)gridDim.x * blockDim.x) {
        const int i = (int)(idx % CP);
        long r = idx / CP;
        const int k = (int)(r % K);
        const int o = (int)(r / K);   // global output channel == g*Og + o_local (rows are group-contiguous)
        float w = 0.0f;
        if (i < Cg) w = v[((long)o * Cg + i) * K + k] * gk[k] / norm[k];
        bf16 h, l;
        split_bf16(w, h, l);
        hi[idx] = h;
        if (lo) lo[idx] = l;
        (void)Og;
    }
}

// dx[b][t][g*Cg + i] += dxg[g][b][pad_left + t][i]
__global__ void unpack_dx_kernel(const float* __restrict__ dxg, float* __restrict__ dx, int B, int T, int D, int G, int Cg,
                                 int pad_left, int Tp) {
    const long n = (long)B * T * D;
    for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long)gridDim.x * blockDim.x) {
        const int c = (int)(idx % D);
        const long bt = idx / D;
        const int t = (int)(bt % T), b = (int)(bt / T);
        const int g = c / Cg, i = c - g * Cg;
        dx[idx] += dxg[(((long)g * B + b) * Tp + pad_left + t) * CP + i];
    }
}

// weight-norm backward, one CTA per tap k:
//   s = sum_{o,i} dwn*v ; dg[k] = s/norm ; dv = dwn*g/norm - v*g*s/norm^3,   dwn[o][i][k] = dwg[o][k*CP + i]
__global__ void __launch_bounds__(256) weight_bwd_kernel(const float* __restrict__ dwg, const float* __restrict__ gk,
                                                         const float* __restrict__ v, const float* __restrict__ norm,
                                                         float* __restrict__ dg, float* __restrict__ dv, int D, int Cg,
                                                         int K) {
    __shared__ float red[8];
    __shared__ float s_total;
    const int k = blockIdx.x;
    const long per_tap = (long)D * Cg;
    float s = 0.0f;
    for (long e = threadIdx.x; e < per_tap; e += blockDim.x) {
        const long o = e / Cg;
        const int i = (int)(e - o * Cg);
        s += dwg[(o * K + k) * CP + i] * v[e * K + k];
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int i = 0; i < 8; ++i) t += red[i];
        s_total = t;
        dg[k] = t / norm[k];
    }
    __syncthreads();
    const float nk = norm[k], g = gk[k];
    const float a = g / nk, bcoef = g * s_total / (nk * nk * nk);
    for (long e = threadIdx.x; e < per_tap; e += blockDim.x) {
        const long o = e / Cg;
        const int i = (int)(e - o * Cg);
        dv[e * K + k] = dwg[(o * K + k) * CP + i] * a - v[e * K + k] * bcoef;
    }
}

// out = dy * act'(pre)  -> fp32 and/or bf16 hi/lo
__global__ void act_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ pre, float* __restrict__ out,
                               bf16* __restrict__ hi, bf16* __restrict__ lo, long n, int mode) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float a = pre[i];
        float d;
        if (mode == 1) d = gelu_erf_grad(a);
        else if (mode == 2) d = silu_grad(a);
        else if (mode == 3) {
            const float q = 1.0f + fabsf(a);
            d = 1.0f / (q * q);
        } else d = 1.0f;
        const float v = dy[i] * d;
        if (out) out[i] = v;
        if (hi) {
            bf16 h, l;
            split_bf16(v, h, l);
            hi[i] = h;
            if (lo) lo[i] = l;
        }
    }
}

inline int grid_for(long n, int block) {
    long g = (n + block - 1) / block;
    const long cap = 148L * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

extern "C" int b2t_posconv_pack_input(const float* x, void* hi, void* lo, int B, int T, int D, int G, int pad_left, int Tp,
                                      b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && hi && B > 0 && T > 0 && D > 0 && G > 0 && D % G == 0 && D / G <= CP && Tp >= T + pad_left,
                  "b2t_posconv_pack_input: bad arguments");
    const long n = (long)G * B * Tp * CP;
    pack_input_kernel<<<grid_for(n, 256), 256, 0, stream>>>(x, (bf16*)hi, (bf16*)lo, B, T, D, G, D / G, pad_left, Tp);
    B2T_CHECK_LAUNCH("b2t_posconv_pack_input");
    return 0;
}

extern "C" int b2t_posconv_pack_weight(const float* g, const float* v, float* norm, void* hi, void* lo, int D, int Cg, int K,
                                       int G, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(g && v && norm && hi && D > 0 && Cg > 0 && Cg <= CP && K > 0 && G > 0 && D % G == 0,
                  "b2t_posconv_pack_weight: bad arguments");
    tap_norm_kernel<<<K, 256, 0, stream>>>(v, norm, (long)D * Cg, K);
    B2T_CHECK_LAUNCH("b2t_posconv_pack_weight(norm)");
    const long n = (long)D * K * CP;
    pack_weight_kernel<<<grid_for(n, 256), 256, 0, stream>>>(g, v, norm, (bf16*)hi, (bf16*)lo, D, Cg, K, G);
    B2T_CHECK_LAUNCH("b2t_posconv_pack_weight");
    return 0;
}

extern "C" int b2t_posconv_unpack_dx(const float* dxg, float* dx, int B, int T, int D, int G, int pad_left, int Tp,
                                     b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(dxg && dx && B > 0 && T > 0 && D > 0 && G > 0 && D % G == 0, "b2t_posconv_unpack_dx: bad arguments");
    unpack_dx_kernel<<<grid_for((long)B * T * D, 256), 256, 0, stream>>>(dxg, dx, B, T, D, G, D / G, pad_left, Tp);
    B2T_CHECK_LAUNCH("b2t_posconv_unpack_dx");
    return 0;
}

extern "C" int b2t_posconv_weight_bwd(const float* dwg, const float* g, const float* v, const float* norm, float* dg,
                                      float* dv, int D, int Cg, int K, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(dwg && g && v && norm && dg && dv && D > 0 && Cg > 0 && K > 0, "b2t_posconv_weight_bwd: bad arguments");
    weight_bwd_kernel<<<K, 256, 0, stream>>>(dwg, g, v, norm, dg, dv, D, Cg, K);
    B2T_CHECK_LAUNCH("b2t_posconv_weight_bwd");
    return 0;
}

extern "C" int b2t_act_bwd(const float* dy, const float* pre, float* out, void* hi, void* lo, int64_t n, int mode,
                           b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(dy && pre && (out || hi) && n > 0 && mode >= 0 && mode <= 3, "b2t_act_bwd: bad arguments");
    act_bwd_kernel<<<grid_for(n, 256), 256, 0, stream>>>(dy, pre, out, (bf16*)hi, (bf16*)lo, n, mode);
    B2T_CHECK_LAUNCH("b2t_act_bwd");
    return 0;
}
