// Stage 1 of the path (b2p2t_model.py:138-167): [optional augmentation] -> Gaussian smoothing.
// The day affine + softsign that follow run as a batched tcgen05 GEMM with a softsign epilogue
// (b2t_gemm, b_batch_idx = day), the unfold is never materialised (unfold view of b2t_gemm).
//
// HBM-bound stencil: one CTA = one trial x TT frames x all channels; the (TT + taps - 1) x C
// input window is staged once in shared memory with coalesced float4 loads, every output
// frame then reads its `taps` neighbours from smem.  Algorithmic traffic: 4 B read + 4 B write
// (bf16 hi+lo) per element.
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

constexpr int TT = 32;   // output frames per CTA

__device__ __forceinline__ float gauss_from_hash(unsigned long long seed, unsigned long long idx) {
    // Box-Muller on two counter-based uniforms
    const float u1 = fmaxf(hash_uniform(seed, 2 * idx), 1e-7f);
    const float u2 = hash_uniform(seed, 2 * idx + 1);
    return sqrtf(-2.0f * __logf(u1)) * __cosf(6.28318530717958647692f * u2);
}

// x [B,T,C] fp32; taps [C, ntaps] (the registered buffer gaussian_smoother.weight [C,1,ntaps]);
// y = sum_k taps[c][k] * xa[b, t + k - pad_left, c], zero outside [0,T)
// xa = x + white_sd*N(0,1) per element + offset_sd*N(0,1) per (b,c)   (train_loop.py:46-62, intended form)
// CC: compile-time channel count (0 = runtime).  ncu (round 2): the kernel was ISSUE-bound, not memory-bound - 24.7 M warp
// instructions for 5.2 M FMAs, most of them the runtime integer divisions of the index arithmetic (window rows, and a
// per-CTA transposition of the tap table that cost as much as the stencil itself).
template <int NT, int CC>
__global__ void __launch_bounds__(256, 3) smooth_fwd_kernel(const float* __restrict__ x, const float* __restrict__ taps,
                                                         float* __restrict__ y, bf16* __restrict__ y_hi,
                                                         bf16* __restrict__ y_lo, int T, int C, int ntaps, int pad_left,
                                                         float white_sd, float offset_sd, unsigned long long seed) {
    if (CC) C = CC;   // index arithmetic below becomes shifts
    extern __shared__ float win[];   // [(TT + ntaps - 1)][C], then the taps transposed to [ntaps][C]
    const int b = blockIdx.y;
    const int t0 = blockIdx.x * TT;
    const int rows = TT + ntaps - 1;
    // taps [C][ntaps] -> smem [ntaps][C]: the stencil loop then reads one conflict-free float4 per tap instead of
    // four strided global loads (those made the kernel run at 170 GB/s, 2.6 % of HBM)
    float* tps = win + rows * C;
    if (NT == 0) {   // generic tap count: table transposed into smem [ntaps][C]
        for (int i = threadIdx.x; i < ntaps * C; i += blockDim.x) tps[i] = __ldg(taps + (i % C) * ntaps + i / C);
    } else {
        // [C][NT] table copied as it is, rows padded to NT + 1 words: coalesced global reads (10 in flight per thread),
        // conflict-free smem writes, and the stencil's per-channel reads (lane stride 2 (NT + 1) words) are 2-way at worst
        for (int i0 = threadIdx.x; i0 < NT * C; i0 += 10 * blockDim.x) {
            float tv[10];
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                const int i = i0 + j * blockDim.x;
                tv[j] = i < NT * C ? __ldg(taps + i) : 0.0f;
            }
#pragma unroll
            for (int j = 0; j < 10; ++j) {
                const int i = i0 + j * blockDim.x;
                if (i < NT * C) tps[i + i / (NT > 0 ? NT : 1)] = tv[j];
            }
        }
    }
    const float* xb = x + (long)b * T * C;
    const int c4n = C >> 2;
    // window rows -> smem, 8 independent 16-byte loads per thread in flight before the first is used: as one
    // load -> (optional noise) -> store loop ptxas kept the order, i.e. 13 serialised L2 / DRAM round trips per CTA with
    // 3 CTAs per SM to hide them - the kernel ran at 1.4 TB/s (0.2 of HBM) on a pure streaming problem
    constexpr int LB = 8;
    const bool noisy = white_sd > 0.0f || offset_sd > 0.0f;
    for (int i0 = threadIdx.x; i0 < rows * c4n; i0 += LB * blockDim.x) {
        float4 v[LB];
#pragma unroll
        for (int j = 0; j < LB; ++j) {
            const int i = i0 + j * blockDim.x;
            const int r = i / c4n, c4 = i - r * c4n;
            const int t = t0 + r - pad_left;
            v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < rows * c4n && t >= 0 && t < T) v[j] = __ldg(reinterpret_cast<const float4*>(xb + (long)t * C + c4 * 4));
        }
#pragma unroll
        for (int j = 0; j < LB; ++j) {
            const int i = i0 + j * blockDim.x;
            if (i >= rows * c4n) break;
            const int r = i / c4n, c4 = i - r * c4n;
            const int t = t0 + r - pad_left;
            if (noisy && t >= 0 && t < T) {
                float* pv = reinterpret_cast<float*>(&v[j]);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int c = c4 * 4 + e;
                    if (white_sd > 0.0f) pv[e] += white_sd * gauss_from_hash(seed, ((unsigned long long)b * T + t) * C + c);
                    if (offset_sd > 0.0f)
                        pv[e] += offset_sd * gauss_from_hash(seed ^ 0x5851F42D4C957F2Dull, (unsigned long long)b * C + c);
                }
            }
            *reinterpret_cast<float4*>(win + r * C + c4 * 4) = v[j];
        }
    }
    __syncthreads();
    auto store_out = [&](int t, int c4, const float (&acc)[4]) {
        const long off = ((long)b * T + t) * C + c4 * 4;
        if (y) *reinterpret_cast<float4*>(y + off) = make_float4(acc[0], acc[1], acc[2], acc[3]);
        if (y_hi) {
            bf16 h[4], l[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) split_bf16(acc[j], h[j], l[j]);
            __nv_bfloat162 p0 = __halves2bfloat162(h[0], h[1]), p1 = __halves2bfloat162(h[2], h[3]);
            uint2 o;
            o.x = *reinterpret_cast<uint32_t*>(&p0);
            o.y = *reinterpret_cast<uint32_t*>(&p1);
            *reinterpret_cast<uint2*>(y_hi + off) = o;
            if (y_lo) {
                p0 = __halves2bfloat162(l[0], l[1]);
                p1 = __halves2bfloat162(l[2], l[3]);
                o.x = *reinterpret_cast<uint32_t*>(&p0);
                o.y = *reinterpret_cast<uint32_t*>(&p1);
                *reinterpret_cast<uint2*>(y_lo + off) = o;
            }
        }
    };
    if (NT > 0) {
        // register-blocked over time: a thread owns 4 channels x FR consecutive frames, keeps the NT taps in
        // registers and reads each window row ONCE (FR + NT - 1 smem loads for FR outputs instead of 2 NT per
        // output); per output the taps are still applied in k = 0..NT-1 order (same rounding as the plain loop)
        // Two channels per thread (not four): the 20 taps of 4 channels are 80 registers, which with the accumulators put
        // the kernel at ~150 registers = ONE 256-thread CTA per SM, 7 waves of load -> sync -> compute -> store with
        // nothing to overlap them (1.4-1.8 TB/s).  40 tap registers leave room for 3 CTAs per SM (the smem limit).
        constexpr int FR = 8;
        constexpr int NTC = NT > 0 ? NT : 1;
        const int c2n = C >> 1;
        for (int i = threadIdx.x; i < (TT / FR) * c2n; i += blockDim.x) {
            const int tg = i / c2n, c2 = i - tg * c2n;
            if (t0 + tg * FR >= T) break;
            float2 w[NTC];
#pragma unroll
            for (int k = 0; k < NTC; ++k) w[k] = make_float2(tps[(c2 * 2) * (NTC + 1) + k], tps[(c2 * 2 + 1) * (NTC + 1) + k]);
            float acc[FR][2];
#pragma unroll
            for (int o = 0; o < FR; ++o) acc[o][0] = acc[o][1] = 0.0f;
#pragma unroll
            for (int r = 0; r < FR + NTC - 1; ++r) {
                const float2 v = *reinterpret_cast<const float2*>(win + (tg * FR + r) * C + c2 * 2);
#pragma unroll
                for (int o = 0; o < FR; ++o) {
                    const int k = r - o;
                    if (k >= 0 && k < NTC) {
                        acc[o][0] = fmaf(w[k].x, v.x, acc[o][0]);
                        acc[o][1] = fmaf(w[k].y, v.y, acc[o][1]);
                    }
                }
            }
#pragma unroll
            for (int o = 0; o < FR; ++o) {
                const int t = t0 + tg * FR + o;
                if (t >= T) break;
                const long off = ((long)b * T + t) * C + c2 * 2;
                if (y) *reinterpret_cast<float2*>(y + off) = make_float2(acc[o][0], acc[o][1]);
                if (y_hi) {
                    __nv_bfloat162 h = __floats2bfloat162_rn(acc[o][0], acc[o][1]);
                    *reinterpret_cast<__nv_bfloat162*>(y_hi + off) = h;
                    if (y_lo) {
                        const float2 hf = __bfloat1622float2(h);
                        *reinterpret_cast<__nv_bfloat162*>(y_lo + off) = __floats2bfloat162_rn(acc[o][0] - hf.x, acc[o][1] - hf.y);
                    }
                }
            }
        }
        return;
    }
    for (int i = threadIdx.x; i < TT * c4n; i += blockDim.x) {
        const int r = i / c4n, c4 = i - r * c4n;
        const int t = t0 + r;
        if (t >= T) break;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int k = 0; k < ntaps; ++k) {
            const float4 v = *reinterpret_cast<const float4*>(win + (r + k) * C + c4 * 4);
            const float4 w = *reinterpret_cast<const float4*>(tps + k * C + c4 * 4);
            acc[0] = fmaf(w.x, v.x, acc[0]);
            acc[1] = fmaf(w.y, v.y, acc[1]);
            acc[2] = fmaf(w.z, v.z, acc[2]);
            acc[3] = fmaf(w.w, v.w, acc[3]);
        }
        store_out(t, c4, acc);
    }
}

// dz = ds * softsign'(z) -> bf16 hi/lo (operand of the day-weight gradient GEMM) and
// dbias[day_b][c] += sum_t dz[b,t,c]
__global__ void __launch_bounds__(256) softsign_bwd_kernel(const float* __restrict__ ds, const float* __restrict__ z,
                                                           const int* __restrict__ day, bf16* __restrict__ dz_hi,
                                                           bf16* __restrict__ dz_lo, float* __restrict__ dbias, int T,
                                                           int C, int rows_per_block) {
    const int b = blockIdx.y;
    const int t0 = blockIdx.x * rows_per_block;
    const int d = day[b];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float acc = 0.0f;
        for (int r = 0; r < rows_per_block; ++r) {
            const int t = t0 + r;
            if (t >= T) break;
            const long off = ((long)b * T + t) * C + c;
            const float den = 1.0f + fabsf(z[off]);
            const float g = ds[off] / (den * den);
            acc += g;
            bf16 h, l;
            split_bf16(g, h, l);
            dz_hi[off] = h;
            if (dz_lo) dz_lo[off] = l;
        }
        atomicAdd(dbias + (long)d * C + c, acc);
    }
}

}  // namespace

extern "C" int b2t_smooth_fwd(const float* x, const float* taps, float* y, void* y_hi, void* y_lo, int B, int T, int C,
                              int ntaps, int pad_left, float white_sd, float offset_sd, uint64_t seed, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && taps && (y || y_hi), "b2t_smooth_fwd: null pointer");
    B2T_CHECK_ARG(B > 0 && T > 0 && C > 0 && C % 4 == 0 && ntaps > 0 && pad_left >= 0 && pad_left < ntaps,
                  "b2t_smooth_fwd: bad sizes B=%d T=%d C=%d taps=%d pad=%d", B, T, C, ntaps, pad_left);
    const size_t smem = (size_t)(TT + 2 * ntaps) * C * sizeof(float);   // window + tap table (rows padded by one word)
    B2T_CHECK_ARG(smem <= 200 * 1024, "b2t_smooth_fwd: window of %zu bytes does not fit in shared memory", smem);
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(smooth_fwd_kernel<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        B2T_CUDA(cudaFuncSetAttribute(smooth_fwd_kernel<20, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        B2T_CUDA(cudaFuncSetAttribute(smooth_fwd_kernel<20, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr = true;
    }
    dim3 grid((T + TT - 1) / TT, B);
#define SCALL(NT_, CC_)                                                                                                  \
    smooth_fwd_kernel<NT_, CC_><<<grid, 256, smem, stream>>>(x, taps, y, (bf16*)y_hi, (bf16*)y_lo, T, C, ntaps, pad_left,  \
                                                             white_sd, offset_sd, (unsigned long long)seed)
    if (ntaps == 20 && C == 256) SCALL(20, 256);   // GaussianSmoothing(256, 20, sigma), b2p2t_model.py:116
    else if (ntaps == 20) SCALL(20, 0);
    else SCALL(0, 0);
#undef SCALL
    B2T_CHECK_LAUNCH("b2t_smooth_fwd");
    return 0;
}

extern "C" int b2t_softsign_bwd(const float* ds, const float* z, const int32_t* day, void* dz_hi, void* dz_lo, float* dbias,
                                int B, int T, int C, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(ds && z && day && dz_hi && dbias && B > 0 && T > 0 && C > 0, "b2t_softsign_bwd: bad arguments");
    const int rows = 16;
    dim3 grid((T + rows - 1) / rows, B);
    softsign_bwd_kernel<<<grid, 256, 0, stream>>>(ds, z, day, (bf16*)dz_hi, (bf16*)dz_lo, dbias, T, C, rows);
    B2T_CHECK_LAUNCH("b2t_softsign_bwd");
    return 0;
}
