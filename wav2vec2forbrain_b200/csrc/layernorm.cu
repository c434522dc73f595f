// LayerNorm forward/backward (one warp per row, row held in registers: one HBM read, one write),
// plus the small HBM-bound helpers that sit between GEMMs: dropout(+split), column sums.
// Replaces nn.LayerNorm / nn.Dropout calls of transformers modeling_wav2vec2.py:690-693,592-609
// and modeling_wav2vec2_conformer.py:577-631.
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

template <int MAXP>
__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, float eps, float* __restrict__ y,
                                                     bf16* __restrict__ y_hi, bf16* __restrict__ y_lo,
                                                     float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
                                                     int n, float drop_p, unsigned long long seed) {
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const float inv_n = 1.0f / (float)n;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    for (int row = warp_global; row < rows; row += nwarps) {
        const float* xr = x + (long)row * n;
        float v[MAXP];
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            v[i] = c < n ? xr[c] : 0.0f;
            s += v[i];
        }
        const float mean = warp_sum(s) * inv_n;
        float q = 0.0f;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            const float d = c < n ? v[i] - mean : 0.0f;
            q += d * d;
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_n + eps);
        if (lane == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            if (c < n) {
                const long off = (long)row * n + c;
                float o = (v[i] - mean) * rstd * gamma[c] + beta[c];
                if (drop_p > 0.0f) o = hash_uniform(seed, (unsigned long long)off) < drop_p ? 0.0f : o * drop_scale;
                if (y) y[off] = o;
                if (y_hi) {
                    bf16 h, l;
                    split_bf16(o, h, l);
                    y_hi[off] = h;
                    if (y_lo) y_lo[off] = l;
                }
            }
        }
    }
}

// dy may carry the forward's output dropout (drop_p/seed over the same [rows,n] offsets).
// dx = LN'(x)^T dy_eff ; dgamma += sum dy_eff*xhat ; dbeta += sum dy_eff
// optional: dx_hi/lo = (dx [* branch dropout mask]) as the A operand of the upstream GEMMs,
//           dcol += column sums of that masked dx (bias gradient of the upstream Linear)
template <int MAXP>
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                     const float* __restrict__ gamma, const float* __restrict__ mean_in,
                                                     const float* __restrict__ rstd_in, float* __restrict__ dx,
                                                     bf16* __restrict__ dx_hi, bf16* __restrict__ dx_lo,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                     float* __restrict__ dcol, int rows, int n, float out_drop_p,
                                                     unsigned long long out_seed, float br_drop_p,
                                                     unsigned long long br_seed, const float* __restrict__ dres) {
    extern __shared__ float sacc[];   // [3][n]
    for (int i = threadIdx.x; i < 3 * n; i += blockDim.x) sacc[i] = 0.0f;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const float inv_n = 1.0f / (float)n;
    const float out_scale = 1.0f / (1.0f - out_drop_p), br_scale = 1.0f / (1.0f - br_drop_p);
    float ag[MAXP], ab[MAXP], ac[MAXP];
#pragma unroll
    for (int i = 0; i < MAXP; ++i) ag[i] = ab[i] = ac[i] = 0.0f;
    for (int row = warp_global; row < rows; row += nwarps) {
        const float mean = mean_in[row], rstd = rstd_in[row];
        float xh[MAXP], g[MAXP];
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            xh[i] = 0.0f;
            g[i] = 0.0f;
            if (c < n) {
                const long off = (long)row * n + c;
                float d = dy[off];
                if (out_drop_p > 0.0f)
                    d = hash_uniform(out_seed, (unsigned long long)off) < out_drop_p ? 0.0f : d * out_scale;
                xh[i] = (x[off] - mean) * rstd;
                ag[i] += d * xh[i];
                ab[i] += d;
                g[i] = d * gamma[c];
                s1 += g[i];
                s2 += g[i] * xh[i];
            }
        }
        s1 = warp_sum(s1) * inv_n;
        s2 = warp_sum(s2) * inv_n;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            if (c < n) {
                const long off = (long)row * n + c;
                float o = rstd * (g[i] - s1 - xh[i] * s2);
                if (dres) o += dres[off];   // + gradient arriving over the residual path around a pre-LN branch
                if (dx) dx[off] = o;
                if (dx_hi || dcol) {
                    float ob = o;
                    if (br_drop_p > 0.0f)
                        ob = hash_uniform(br_seed, (unsigned long long)off) < br_drop_p ? 0.0f : o * br_scale;
                    ac[i] += ob;
                    if (dx_hi) {
                        bf16 h, l;
                        split_bf16(ob, h, l);
                        dx_hi[off] = h;
                        if (dx_lo) dx_lo[off] = l;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
        const int c = lane + 32 * i;
        if (c < n) {
            atomicAdd(&sacc[c], ag[i]);
            atomicAdd(&sacc[n + c], ab[i]);
            atomicAdd(&sacc[2 * n + c], ac[i]);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < n; c += blockDim.x) {
        if (dgamma) atomicAdd(dgamma + c, sacc[c]);
        if (dbeta) atomicAdd(dbeta + c, sacc[n + c]);
        if (dcol) atomicAdd(dcol + c, sacc[2 * n + c]);
    }
}

// ---------------------------------------------------------------------------------------------
// vectorised variants (n % 4 == 0): a lane owns float4 columns lane + 32 i, a warp one row; the
// parameter-gradient column sums of the backward are a separate column-parallel kernel, so the
// row kernel keeps no per-column accumulators (the scalar kernel needs ~250 registers for them).
// ---------------------------------------------------------------------------------------------
template <int VPL>
__global__ void __launch_bounds__(256) ln_fwd_vec_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, float eps, float* __restrict__ y,
                                                         bf16* __restrict__ y_hi, bf16* __restrict__ y_lo,
                                                         float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
                                                         int n, float drop_p, unsigned long long seed) {
    B2T_PDL_TRIGGER();
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const float inv_n = 1.0f / (float)n;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    for (int row = warp_global; row < rows; row += nwarps) {
        const float* xr = x + (long)row * n;
        float4 v[VPL];
        float s = 0.0f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (lane + 32 * i) * 4;
            v[i] = c < n ? *reinterpret_cast<const float4*>(xr + c) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
        }
        const float mean = warp_sum(s) * inv_n;
        float q = 0.0f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            if ((lane + 32 * i) * 4 < n) {
                const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
                q += (a * a + b * b) + (c * c + d * d);
            }
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_n + eps);
        if (lane == 0) {
            if (mean_out) mean_out[row] = mean;
            if (rstd_out) rstd_out[row] = rstd;
        }
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (lane + 32 * i) * 4;
            if (c < n) {
                const float4 ga = *reinterpret_cast<const float4*>(gamma + c), be = *reinterpret_cast<const float4*>(beta + c);
                const long off = (long)row * n + c;
                float o[4] = {(v[i].x - mean) * rstd * ga.x + be.x, (v[i].y - mean) * rstd * ga.y + be.y,
                              (v[i].z - mean) * rstd * ga.z + be.z, (v[i].w - mean) * rstd * ga.w + be.w};
                if (drop_p > 0.0f) {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        o[e] = hash_uniform(seed, (unsigned long long)(off + e)) < drop_p ? 0.0f : o[e] * drop_scale;
                }
                if (y) *reinterpret_cast<float4*>(y + off) = make_float4(o[0], o[1], o[2], o[3]);
                if (y_hi) store_split4(o, y_hi, y_lo, off);
            }
        }
    }
}

template <int VPL>
__global__ void __launch_bounds__(256, 2) ln_bwd_vec_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                         const float* __restrict__ gamma, const float* __restrict__ mean_in,
                                                         const float* __restrict__ rstd_in, float* __restrict__ dx,
                                                         bf16* __restrict__ dx_hi, bf16* __restrict__ dx_lo, int rows, int n,
                                                         float out_drop_p, unsigned long long out_seed, float br_drop_p,
                                                         unsigned long long br_seed, const float* __restrict__ dres) {
    B2T_PDL_TRIGGER();
    const int lane = threadIdx.x & 31;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const float inv_n = 1.0f / (float)n;
    const float out_scale = 1.0f / (1.0f - out_drop_p), br_scale = 1.0f / (1.0f - br_drop_p);
    for (int row = warp_global; row < rows; row += nwarps) {
        const float mean = mean_in[row], rstd = rstd_in[row];
        float xh[VPL][4], g[VPL][4];
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (lane + 32 * i) * 4;
            if (c < n) {
                const long off = (long)row * n + c;
                const float4 d4 = *reinterpret_cast<const float4*>(dy + off);
                const float4 x4 = *reinterpret_cast<const float4*>(x + off);
                const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
                // the residual gradient is only needed after the two row reductions: start its trip from DRAM now
                if (dres && (lane & 7) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(dres + off));
                float d[4] = {d4.x, d4.y, d4.z, d4.w};
                const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
                const float gv[4] = {ga.x, ga.y, ga.z, ga.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (out_drop_p > 0.0f)
                        d[e] = hash_uniform(out_seed, (unsigned long long)(off + e)) < out_drop_p ? 0.0f : d[e] * out_scale;
                    xh[i][e] = (xv[e] - mean) * rstd;
                    g[i][e] = d[e] * gv[e];
                    s1 += g[i][e];
                    s2 += g[i][e] * xh[i][e];
                }
            }
        }
        s1 = warp_sum(s1) * inv_n;
        s2 = warp_sum(s2) * inv_n;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = (lane + 32 * i) * 4;
            if (c < n) {
                const long off = (long)row * n + c;
                float o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) o[e] = rstd * (g[i][e] - s1 - xh[i][e] * s2);
                if (dres) {
                    const float4 r4 = *reinterpret_cast<const float4*>(dres + off);
                    o[0] += r4.x, o[1] += r4.y, o[2] += r4.z, o[3] += r4.w;
                }
                if (dx) *reinterpret_cast<float4*>(dx + off) = make_float4(o[0], o[1], o[2], o[3]);
                if (dx_hi) {
                    if (br_drop_p > 0.0f) {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            o[e] = hash_uniform(br_seed, (unsigned long long)(off + e)) < br_drop_p ? 0.0f : o[e] * br_scale;
                    }
                    store_split4(o, dx_hi, dx_lo, off);
                }
            }
        }
    }
}

// forward twin of ln_bwd_vec2_kernel (below): two warps per row, 4 rows per CTA, mean and variance combined
// through 64 bytes of smem (two-pass variance on the register copy of the row, as in the one-warp kernel)
template <int VPL>
__global__ void __launch_bounds__(256) ln_fwd_vec2_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps, float* __restrict__ y,
                                                          bf16* __restrict__ y_hi, bf16* __restrict__ y_lo,
                                                          float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
                                                          int n, float drop_p, unsigned long long seed) {
    B2T_PDL_TRIGGER();
    __shared__ float red[2][4][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = warp >> 1, half = warp & 1;
    const int row = blockIdx.x * 4 + rl;
    const bool row_ok = row < rows;
    const float inv_n = 1.0f / (float)n;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    float4 v[VPL];
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((half * VPL + i) * 32 + lane) * 4;
        v[i] = (row_ok && c < n) ? *reinterpret_cast<const float4*>(x + (long)row * n + c) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
    s = warp_sum(s);
    if (lane == 0) red[0][rl][half] = s;
    __syncthreads();
    const float mean = (red[0][rl][0] + red[0][rl][1]) * inv_n;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        if (((half * VPL + i) * 32 + lane) * 4 < n) {
            const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
    }
    q = warp_sum(q);
    if (lane == 0) red[1][rl][half] = q;
    __syncthreads();
    const float rstd = rsqrtf((red[1][rl][0] + red[1][rl][1]) * inv_n + eps);
    if (row_ok && half == 0 && lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((half * VPL + i) * 32 + lane) * 4;
        if (row_ok && c < n) {
            const float4 ga = *reinterpret_cast<const float4*>(gamma + c), be = *reinterpret_cast<const float4*>(beta + c);
            const long off = (long)row * n + c;
            float o[4] = {(v[i].x - mean) * rstd * ga.x + be.x, (v[i].y - mean) * rstd * ga.y + be.y,
                          (v[i].z - mean) * rstd * ga.z + be.z, (v[i].w - mean) * rstd * ga.w + be.w};
            if (drop_p > 0.0f) {
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    o[e] = hash_uniform(seed, (unsigned long long)(off + e)) < drop_p ? 0.0f : o[e] * drop_scale;
            }
            if (y) *reinterpret_cast<float4*>(y + off) = make_float4(o[0], o[1], o[2], o[3]);
            if (y_hi) store_split4(o, y_hi, y_lo, off);
        }
    }
}

// Wide rows (n >= 512): TWO warps per row, VPL float4 per lane each, 4 rows per 256-thread CTA.  One warp per
// row left 3776 warps of 164 registers for the whole [3776, 1024] problem (1 CTA/SM, 12 % of the warp slots,
// 13 % of DRAM bandwidth in ncu); half the per-thread footprint doubles the rows in flight per SM.
template <int VPL>
__global__ void __launch_bounds__(256) ln_bwd_vec2_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                          const float* __restrict__ gamma, const float* __restrict__ mean_in,
                                                          const float* __restrict__ rstd_in, float* __restrict__ dx,
                                                          bf16* __restrict__ dx_hi, bf16* __restrict__ dx_lo, int rows, int n,
                                                          float out_drop_p, unsigned long long out_seed, float br_drop_p,
                                                          unsigned long long br_seed, const float* __restrict__ dres) {
    B2T_PDL_TRIGGER();
    __shared__ float red[4][2][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = warp >> 1, half = warp & 1;
    const int row = blockIdx.x * 4 + rl;
    const bool row_ok = row < rows;
    const float inv_n = 1.0f / (float)n;
    const float out_scale = 1.0f / (1.0f - out_drop_p), br_scale = 1.0f / (1.0f - br_drop_p);
    const float mean = row_ok ? mean_in[row] : 0.0f, rstd = row_ok ? rstd_in[row] : 0.0f;
    float xh[VPL][4], g[VPL][4];
    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((half * VPL + i) * 32 + lane) * 4;
#pragma unroll
        for (int e = 0; e < 4; ++e) xh[i][e] = g[i][e] = 0.0f;
        if (row_ok && c < n) {
            const long off = (long)row * n + c;
            const float4 d4 = *reinterpret_cast<const float4*>(dy + off);
            const float4 x4 = *reinterpret_cast<const float4*>(x + off);
            const float4 ga = *reinterpret_cast<const float4*>(gamma + c);
            if (dres && (lane & 7) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(dres + off));
            float d[4] = {d4.x, d4.y, d4.z, d4.w};
            const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
            const float gv[4] = {ga.x, ga.y, ga.z, ga.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (out_drop_p > 0.0f)
                    d[e] = hash_uniform(out_seed, (unsigned long long)(off + e)) < out_drop_p ? 0.0f : d[e] * out_scale;
                xh[i][e] = (xv[e] - mean) * rstd;
                g[i][e] = d[e] * gv[e];
                s1 += g[i][e];
                s2 += g[i][e] * xh[i][e];
            }
        }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) {
        red[rl][half][0] = s1;
        red[rl][half][1] = s2;
    }
    __syncthreads();
    s1 = (red[rl][0][0] + red[rl][1][0]) * inv_n;
    s2 = (red[rl][0][1] + red[rl][1][1]) * inv_n;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((half * VPL + i) * 32 + lane) * 4;
        if (row_ok && c < n) {
            const long off = (long)row * n + c;
            float o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) o[e] = rstd * (g[i][e] - s1 - xh[i][e] * s2);
            if (dres) {
                const float4 r4 = *reinterpret_cast<const float4*>(dres + off);
                o[0] += r4.x, o[1] += r4.y, o[2] += r4.z, o[3] += r4.w;
            }
            if (dx) *reinterpret_cast<float4*>(dx + off) = make_float4(o[0], o[1], o[2], o[3]);
            if (dx_hi) {
                if (br_drop_p > 0.0f) {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        o[e] = hash_uniform(br_seed, (unsigned long long)(off + e)) < br_drop_p ? 0.0f : o[e] * br_scale;
                }
                store_split4(o, dx_hi, dx_lo, off);
            }
        }
    }
}

// One-pass LayerNorm backward for wide rows: ln_bwd_vec2_kernel's row math in a loop over row groups (grid = 2 CTAs per
// SM), with the column sums (dgamma, dbeta, dcol) accumulated in registers - a thread's columns are the same in every
// iteration - and folded through 12 KB of shared memory into one vector red.add per 4 columns and CTA at the end.
// The separate column kernel re-read dy, x and dx (12 of the 32 bytes per element of the two-kernel version).
template <int VPL>
__global__ void __launch_bounds__(256, 2) ln_bwd_fused_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                             const float* __restrict__ gamma, const float* __restrict__ mean_in,
                                                             const float* __restrict__ rstd_in, float* __restrict__ dx,
                                                             bf16* __restrict__ dx_hi, bf16* __restrict__ dx_lo,
                                                             float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                             float* __restrict__ dcol, int rows, int n, float out_drop_p,
                                                             unsigned long long out_seed, float br_drop_p,
                                                             unsigned long long br_seed, const float* __restrict__ dres) {
    B2T_PDL_TRIGGER();
    __shared__ float red[2][4][2][2];
    __shared__ float csum[3][1024];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = warp >> 1, half = warp & 1;
    const float inv_n = 1.0f / (float)n;
    const float out_scale = 1.0f / (1.0f - out_drop_p), br_scale = 1.0f / (1.0f - br_drop_p);
    for (int i = threadIdx.x; i < 3 * 1024; i += 256) (&csum[0][0])[i] = 0.0f;
    float ag[VPL][4], ab[VPL][4], ac[VPL][4];
    float4 ga[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((half * VPL + i) * 32 + lane) * 4;
        ga[i] = c < n ? *reinterpret_cast<const float4*>(gamma + c) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int e = 0; e < 4; ++e) ag[i][e] = ab[i][e] = ac[i][e] = 0.0f;
    }
    const int groups = (rows + 3) / 4;
    int par = 0;
    for (int grp = blockIdx.x; grp < groups; grp += gridDim.x, par ^= 1) {
        const int row = grp * 4 + rl;
        const bool row_ok = row < rows;
        const float mean = row_ok ? mean_in[row] : 0.0f, rstd = row_ok ? rstd_in[row] : 0.0f;
        float xh[VPL][4], g[VPL][4];
        float4 d4[VPL], x4[VPL];
#pragma unroll
        for (int i = 0; i < VPL; ++i) {   // all loads of the row first
            const int c = ((half * VPL + i) * 32 + lane) * 4;
            const long off = (long)row * n + c;
            const bool ok = row_ok && c < n;
            d4[i] = ok ? *reinterpret_cast<const float4*>(dy + off) : make_float4(0.f, 0.f, 0.f, 0.f);
            x4[i] = ok ? *reinterpret_cast<const float4*>(x + off) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok && dres && (lane & 7) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(dres + off));
        }
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = ((half * VPL + i) * 32 + lane) * 4;
            const long off = (long)row * n + c;
            const bool ok = row_ok && c < n;
            float d[4] = {d4[i].x, d4[i].y, d4[i].z, d4[i].w};
            const float xv[4] = {x4[i].x, x4[i].y, x4[i].z, x4[i].w};
            const float gv[4] = {ga[i].x, ga[i].y, ga[i].z, ga[i].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (out_drop_p > 0.0f && ok)
                    d[e] = hash_uniform(out_seed, (unsigned long long)(off + e)) < out_drop_p ? 0.0f : d[e] * out_scale;
                xh[i][e] = ok ? (xv[e] - mean) * rstd : 0.0f;
                g[i][e] = d[e] * gv[e];
                s1 += g[i][e];
                s2 += g[i][e] * xh[i][e];
                ag[i][e] += d[e] * xh[i][e];
                ab[i][e] += d[e];
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (lane == 0) {
            red[par][rl][half][0] = s1;
            red[par][rl][half][1] = s2;
        }
        __syncthreads();
        s1 = (red[par][rl][0][0] + red[par][rl][1][0]) * inv_n;
        s2 = (red[par][rl][0][1] + red[par][rl][1][1]) * inv_n;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = ((half * VPL + i) * 32 + lane) * 4;
            if (row_ok && c < n) {
                const long off = (long)row * n + c;
                float o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) o[e] = rstd * (g[i][e] - s1 - xh[i][e] * s2);
                if (dres) {
                    const float4 r4 = *reinterpret_cast<const float4*>(dres + off);
                    o[0] += r4.x, o[1] += r4.y, o[2] += r4.z, o[3] += r4.w;
                }
                if (dx) *reinterpret_cast<float4*>(dx + off) = make_float4(o[0], o[1], o[2], o[3]);
                if (br_drop_p > 0.0f && (dx_hi || dcol)) {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        o[e] = hash_uniform(br_seed, (unsigned long long)(off + e)) < br_drop_p ? 0.0f : o[e] * br_scale;
                }
                if (dx_hi) store_split4(o, dx_hi, dx_lo, off);
#pragma unroll
                for (int e = 0; e < 4; ++e) ac[i][e] += o[e];
            }
        }
    }
    __syncthreads();   // csum was zeroed by the first loop of the kernel before anyone adds to it
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((half * VPL + i) * 32 + lane) * 4;
        if (c < n) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                atomicAdd(&csum[0][c + e], ag[i][e]);
                atomicAdd(&csum[1][c + e], ab[i][e]);
                atomicAdd(&csum[2][c + e], ac[i][e]);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * (n >> 2); i += 256) {
        const int which = i / (n >> 2), c = (i - which * (n >> 2)) * 4;
        float* dst = which == 0 ? dgamma : (which == 1 ? dbeta : dcol);
        if (!dst) continue;
        const float4 v = *reinterpret_cast<const float4*>(&csum[which][c]);
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + c), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    }
}

// column sums of the LayerNorm backward: dgamma += sum_r dy.xhat, dbeta += sum_r dy, dcol += sum_r mask.dx
// grid (ceil(n / 32), row slabs), block (32, 8): a warp reads 128 contiguous bytes of a row.
constexpr int LNC_ROWS = 128;
__global__ void __launch_bounds__(256) ln_bwd_cols_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                          const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                                                          const float* __restrict__ dx, float* __restrict__ dgamma,
                                                          float* __restrict__ dbeta, float* __restrict__ dcol, int rows, int n,
                                                          float out_drop_p, unsigned long long out_seed, float br_drop_p,
                                                          unsigned long long br_seed) {
    __shared__ float part[3][8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int r0 = blockIdx.y * LNC_ROWS;
    const int r1 = r0 + LNC_ROWS < rows ? r0 + LNC_ROWS : rows;
    const float out_scale = 1.0f / (1.0f - out_drop_p), br_scale = 1.0f / (1.0f - br_drop_p);
    float ag = 0.0f, ab = 0.0f, ac = 0.0f;
    if (c < n) {
        for (int r = r0 + threadIdx.y; r < r1; r += 8) {
            const long off = (long)r * n + c;
            float d = dy[off];
            if (out_drop_p > 0.0f) d = hash_uniform(out_seed, (unsigned long long)off) < out_drop_p ? 0.0f : d * out_scale;
            ag += d * (x[off] - mean_in[r]) * rstd_in[r];
            ab += d;
            if (dcol) {
                float o = dx[off];
                if (br_drop_p > 0.0f) o = hash_uniform(br_seed, (unsigned long long)off) < br_drop_p ? 0.0f : o * br_scale;
                ac += o;
            }
        }
    }
    part[0][threadIdx.y][threadIdx.x] = ag;
    part[1][threadIdx.y][threadIdx.x] = ab;
    part[2][threadIdx.y][threadIdx.x] = ac;
    __syncthreads();
    if (threadIdx.y < 3 && c < n) {
        float sum = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += part[threadIdx.y][j][threadIdx.x];
        float* dst = threadIdx.y == 0 ? dgamma : (threadIdx.y == 1 ? dbeta : dcol);
        if (dst) atomicAdd(dst + c, sum);
    }
}

// float4 variant: a thread owns 4 consecutive columns, a warp reads 512 contiguous bytes of a row, 4 rows of
// (dy, x, dx) are requested before the first is used.  grid (ceil(n / 128), row slabs of 64), block (32, 8).
constexpr int LNCV_ROWS = 64;
__global__ void __launch_bounds__(256) ln_bwd_cols_vec_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                              const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                                                              const float* __restrict__ dx, float* __restrict__ dgamma,
                                                              float* __restrict__ dbeta, float* __restrict__ dcol, int rows, int n,
                                                              float out_drop_p, unsigned long long out_seed, float br_drop_p,
                                                              unsigned long long br_seed) {
    B2T_PDL_TRIGGER();
    __shared__ float4 part[3][8][33];
    const int c = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int r0 = blockIdx.y * LNCV_ROWS;
    const int r1 = r0 + LNCV_ROWS < rows ? r0 + LNCV_ROWS : rows;
    const float out_scale = 1.0f / (1.0f - out_drop_p), br_scale = 1.0f / (1.0f - br_drop_p);
    float ag[4] = {0.f, 0.f, 0.f, 0.f}, ab[4] = {0.f, 0.f, 0.f, 0.f}, ac[4] = {0.f, 0.f, 0.f, 0.f};
    if (c < n) {
        for (int rb = r0 + threadIdx.y; rb < r1; rb += 32) {
            float4 d4[4], x4[4], o4[4];
            float mu[4], rs[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = rb + 8 * i;
                const bool ok = r < r1;
                const long off = (long)(ok ? r : rb) * n + c;
                d4[i] = *reinterpret_cast<const float4*>(dy + off);
                x4[i] = *reinterpret_cast<const float4*>(x + off);
                if (dcol) o4[i] = *reinterpret_cast<const float4*>(dx + off);
                mu[i] = mean_in[ok ? r : rb];
                rs[i] = ok ? rstd_in[r] : 0.0f;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = rb + 8 * i;
                if (r >= r1) continue;
                const long off = (long)r * n + c;
                float d[4] = {d4[i].x, d4[i].y, d4[i].z, d4[i].w};
                const float xv[4] = {x4[i].x, x4[i].y, x4[i].z, x4[i].w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (out_drop_p > 0.0f)
                        d[e] = hash_uniform(out_seed, (unsigned long long)(off + e)) < out_drop_p ? 0.0f : d[e] * out_scale;
                    ag[e] += d[e] * (xv[e] - mu[i]) * rs[i];
                    ab[e] += d[e];
                }
                if (dcol) {
                    float o[4] = {o4[i].x, o4[i].y, o4[i].z, o4[i].w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (br_drop_p > 0.0f)
                            o[e] = hash_uniform(br_seed, (unsigned long long)(off + e)) < br_drop_p ? 0.0f : o[e] * br_scale;
                        ac[e] += o[e];
                    }
                }
            }
        }
    }
    part[0][threadIdx.y][threadIdx.x] = make_float4(ag[0], ag[1], ag[2], ag[3]);
    part[1][threadIdx.y][threadIdx.x] = make_float4(ab[0], ab[1], ab[2], ab[3]);
    part[2][threadIdx.y][threadIdx.x] = make_float4(ac[0], ac[1], ac[2], ac[3]);
    __syncthreads();
    if (threadIdx.y < 3 && c < n) {
        float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 t = part[threadIdx.y][j][threadIdx.x];
            sum.x += t.x, sum.y += t.y, sum.z += t.z, sum.w += t.w;
        }
        float* dst = threadIdx.y == 0 ? dgamma : (threadIdx.y == 1 ? dbeta : dcol);
        if (dst) {
            atomicAdd(dst + c, sum.x);
            atomicAdd(dst + c + 1, sum.y);
            atomicAdd(dst + c + 2, sum.z);
            atomicAdd(dst + c + 3, sum.w);
        }
    }
}

// y = dropout(x) (same hash mask contract as the GEMM epilogue) with optional fp32 / hi / lo outputs;
// p == 0 degenerates to a plain split.  Also the backward (apply to the incoming gradient).
__global__ void dropout_kernel(const float* __restrict__ x, float* __restrict__ y, bf16* __restrict__ hi,
                               bf16* __restrict__ lo, long n, float p, unsigned long long seed) {
    B2T_PDL_TRIGGER();
    const float scale = 1.0f / (1.0f - p);
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        float v = x[i];
        if (p > 0.0f) v = hash_uniform(seed, (unsigned long long)i) < p ? 0.0f : v * scale;
        if (y) y[i] = v;
        if (hi) {
            bf16 h, l;
            split_bf16(v, h, l);
            hi[i] = h;
            if (lo) lo[i] = l;
        }
    }
}

// out[c] += sum_r x[r, c]; x row stride ld.  grid (col tiles of 32, row slabs); 32x8 threads
__global__ void colsum_kernel(const float* __restrict__ x, float* __restrict__ out, long rows, int n, long ld,
                              long rows_per_block) {
    B2T_PDL_TRIGGER();
    __shared__ float part[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long r0 = (long)blockIdx.y * rows_per_block;
    const long r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
    float acc = 0.0f;
    if (c < n)
        for (long r = r0 + threadIdx.y; r < r1; r += 8) acc += x[r * ld + c];
    part[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < n) {
        float s = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += part[j][threadIdx.x];
        atomicAdd(out + c, s);
    }
}

__global__ void axpby_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long n,
                             float alpha, float beta_) {
    B2T_PDL_TRIGGER();
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        out[i] = alpha * a[i] + (b ? beta_ * b[i] : 0.0f);
}

inline int grid_for(long n, int block, int cap_mult = 16) {
    long g = (n + block - 1) / block;
    const long cap = 148L * cap_mult;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

#define LN_DISPATCH(n, CALL)                       \
    do {                                           \
        const int per = ((n) + 31) / 32;           \
        if (per <= 2) { CALL(2); }                 \
        else if (per <= 8) { CALL(8); }            \
        else if (per <= 24) { CALL(24); }          \
        else { CALL(32); }                         \
    } while (0)

extern "C" int b2t_layernorm_fwd(const float* x, const float* gamma, const float* beta, float eps, float* y, void* y_hi,
                                 void* y_lo, float* mean, float* rstd, int rows, int n, float drop_p, uint64_t seed,
                                 b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && gamma && beta && (y || y_hi) && rows > 0 && n > 0 && n <= 1024, "b2t_layernorm_fwd: bad arguments (n=%d)", n);
    B2T_CHECK_ARG(drop_p >= 0.0f && drop_p < 1.0f, "b2t_layernorm_fwd: drop_p");
    const int grid = grid_for((long)rows * 32, 256, 8);
    auto al16 = [](const void* q) { return ((uintptr_t)q & 15) == 0; };
    if (n % 4 == 0 && al16(x) && al16(gamma) && al16(beta) && al16(y) && ((uintptr_t)y_hi & 7) == 0 && ((uintptr_t)y_lo & 7) == 0) {
        const int vpl = (n / 4 + 31) / 32;
#define VCALL(V)                                                                                                         \
    ln_fwd_vec_kernel<V><<<grid, 256, 0, stream>>>(x, gamma, beta, eps, y, (bf16*)y_hi, (bf16*)y_lo, mean, rstd, rows, n, \
                                                   drop_p, (unsigned long long)seed)
        if (n >= 512) {
            const int v2 = (n / 4 + 63) / 64, g2 = (rows + 3) / 4;
#define V2CALL(V)                                                                                                        \
    ln_fwd_vec2_kernel<V><<<g2, 256, 0, stream>>>(x, gamma, beta, eps, y, (bf16*)y_hi, (bf16*)y_lo, mean, rstd, rows, n, \
                                                  drop_p, (unsigned long long)seed)
            if (v2 <= 2) { V2CALL(2); } else if (v2 <= 3) { V2CALL(3); } else { V2CALL(4); }
#undef V2CALL
        } else if (vpl <= 1) { VCALL(1); } else if (vpl <= 2) { VCALL(2); } else if (vpl <= 4) { VCALL(4); } else if (vpl <= 6) { VCALL(6); } else { VCALL(8); }
#undef VCALL
        B2T_CHECK_LAUNCH("b2t_layernorm_fwd");
        return 0;
    }
#define CALL(P)                                                                                                          \
    ln_fwd_kernel<P><<<grid, 256, 0, stream>>>(x, gamma, beta, eps, y, (bf16*)y_hi, (bf16*)y_lo, mean, rstd, rows, n, drop_p, \
                                               (unsigned long long)seed)
    LN_DISPATCH(n, CALL);
#undef CALL
    B2T_CHECK_LAUNCH("b2t_layernorm_fwd");
    return 0;
}

extern "C" int b2t_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                                 float* dx, void* dx_hi, void* dx_lo, float* dgamma, float* dbeta, float* dcol, int rows,
                                 int n, float out_drop_p, uint64_t out_seed, float br_drop_p, uint64_t br_seed,
                                 const float* dres, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(dy && x && gamma && mean && rstd && rows > 0 && n > 0 && n <= 1024, "b2t_layernorm_bwd: bad arguments");
    auto al16 = [](const void* q) { return ((uintptr_t)q & 15) == 0; };
    if (n % 4 == 0 && al16(dy) && al16(x) && al16(gamma) && al16(dx) && al16(dres) && ((uintptr_t)dx_hi & 7) == 0 &&
        ((uintptr_t)dx_lo & 7) == 0 && (dx || !dcol)) {
        const int vgrid = grid_for((long)rows * 32, 256, 8);
        const int vpl = (n / 4 + 31) / 32;
#define VCALL(V)                                                                                                          \
    ln_bwd_vec_kernel<V><<<vgrid, 256, 0, stream>>>(dy, x, gamma, mean, rstd, dx, (bf16*)dx_hi, (bf16*)dx_lo, rows, n, out_drop_p, \
                                                    (unsigned long long)out_seed, br_drop_p, (unsigned long long)br_seed, dres)
        const bool cols16 = (((uintptr_t)dgamma | (uintptr_t)dbeta | (uintptr_t)dcol) & 15) == 0;
        static int fused_on = -1;
        if (fused_on < 0) {
            const char* e = getenv("B2T_LN_BWD_FUSED");
            fused_on = e ? atoi(e) : 1;
        }
        if (n >= 512 && (dgamma || dbeta || dcol) && cols16 && fused_on) {
            // one pass: row gradients and column sums together
            const int v2 = (n / 4 + 63) / 64, groups = (rows + 3) / 4;
            int sms = 148, dev = 0;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            const int gf = groups < 2 * sms ? groups : 2 * sms;
#define FCALL(V)                                                                                                          \
    ln_bwd_fused_kernel<V><<<gf, 256, 0, stream>>>(dy, x, gamma, mean, rstd, dx, (bf16*)dx_hi, (bf16*)dx_lo, dgamma, dbeta, dcol, \
                                                   rows, n, out_drop_p, (unsigned long long)out_seed, br_drop_p,                 \
                                                   (unsigned long long)br_seed, dres)
            if (v2 <= 2) { FCALL(2); } else if (v2 <= 3) { FCALL(3); } else { FCALL(4); }
#undef FCALL
            B2T_CHECK_LAUNCH("b2t_layernorm_bwd (fused)");
            return 0;
        }
        if (n >= 512) {
            const int v2 = (n / 4 + 63) / 64, g2 = (rows + 3) / 4;
#define V2CALL(V)                                                                                                         \
    ln_bwd_vec2_kernel<V><<<g2, 256, 0, stream>>>(dy, x, gamma, mean, rstd, dx, (bf16*)dx_hi, (bf16*)dx_lo, rows, n, out_drop_p, \
                                                  (unsigned long long)out_seed, br_drop_p, (unsigned long long)br_seed, dres)
            if (v2 <= 2) { V2CALL(2); } else if (v2 <= 3) { V2CALL(3); } else { V2CALL(4); }
#undef V2CALL
        } else if (vpl <= 1) { VCALL(1); } else if (vpl <= 2) { VCALL(2); } else if (vpl <= 4) { VCALL(4); } else if (vpl <= 6) { VCALL(6); } else { VCALL(8); }
#undef VCALL
        B2T_CHECK_LAUNCH("b2t_layernorm_bwd");
        if (dgamma || dbeta || dcol) {
            const bool v4 = n % 4 == 0 && (((uintptr_t)dy | (uintptr_t)x | (uintptr_t)dx) & 15) == 0;
            if (v4) {
                dim3 vgrid((n + 127) / 128, (rows + LNCV_ROWS - 1) / LNCV_ROWS);
                ln_bwd_cols_vec_kernel<<<vgrid, dim3(32, 8), 0, stream>>>(dy, x, mean, rstd, dx, dgamma, dbeta, dcol, rows, n,
                                                                          out_drop_p, (unsigned long long)out_seed, br_drop_p,
                                                                          (unsigned long long)br_seed);
                B2T_CHECK_LAUNCH("b2t_layernorm_bwd (column sums)");
                return 0;
            }
            dim3 cgrid((n + 31) / 32, (rows + LNC_ROWS - 1) / LNC_ROWS);
            ln_bwd_cols_kernel<<<cgrid, dim3(32, 8), 0, stream>>>(dy, x, mean, rstd, dx, dgamma, dbeta, dcol, rows, n, out_drop_p,
                                                                  (unsigned long long)out_seed, br_drop_p,
                                                                  (unsigned long long)br_seed);
            B2T_CHECK_LAUNCH("b2t_layernorm_bwd (column sums)");
        }
        return 0;
    }
    const int grid = grid_for((long)rows * 32, 256, 2);
    const size_t smem = 3 * (size_t)n * sizeof(float);
#define CALL(P)                                                                                                       \
    ln_bwd_kernel<P><<<grid, 256, smem, stream>>>(dy, x, gamma, mean, rstd, dx, (bf16*)dx_hi, (bf16*)dx_lo, dgamma, dbeta, \
                                                  dcol, rows, n, out_drop_p, (unsigned long long)out_seed, br_drop_p,  \
                                                  (unsigned long long)br_seed, dres)
    LN_DISPATCH(n, CALL);
#undef CALL
    B2T_CHECK_LAUNCH("b2t_layernorm_bwd");
    return 0;
}

extern "C" int b2t_dropout(const float* x, float* y, void* hi, void* lo, int64_t n, float p, uint64_t seed,
                           b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && (y || hi) && n > 0 && p >= 0.0f && p < 1.0f, "b2t_dropout: bad arguments");
    dropout_kernel<<<grid_for(n, 256), 256, 0, stream>>>(x, y, (bf16*)hi, (bf16*)lo, n, p, (unsigned long long)seed);
    B2T_CHECK_LAUNCH("b2t_dropout");
    return 0;
}

extern "C" int b2t_colsum(const float* x, float* out, int64_t rows, int n, int64_t ld, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && out && rows > 0 && n > 0 && ld >= n, "b2t_colsum: bad arguments");
    const long rpb = 256;
    dim3 grid((n + 31) / 32, (unsigned)((rows + rpb - 1) / rpb));
    colsum_kernel<<<grid, dim3(32, 8), 0, stream>>>(x, out, rows, n, ld, rpb);
    B2T_CHECK_LAUNCH("b2t_colsum");
    return 0;
}

extern "C" int b2t_axpby(const float* a, const float* b, float* out, int64_t n, float alpha, float beta, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(a && out && n > 0, "b2t_axpby: bad arguments");
    axpby_kernel<<<grid_for(n, 256), 256, 0, stream>>>(a, b, out, n, alpha, beta);
    B2T_CHECK_LAUNCH("b2t_axpby");
    return 0;
}
