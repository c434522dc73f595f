// Conformer-specific kernels of the wav2vec2-conformer (RoPE) encoder block
// (transformers modeling_wav2vec2_conformer.py:568-631): everything of the block that is not a
// GEMM, LayerNorm or attention call.
//
//   R1  LayerNorm + rotary embedding in one pass (:466-471,509-527): the layer-normed state is
//       written twice as bf16 hi/lo GEMM operands — rotated (A operand of the Q/K projection)
//       and plain (A operand of the V projection) — without an fp32 round trip.
//   V1  convolution module (:398-417) between the two pointwise GEMMs:
//       GLU -> depthwise conv (k=31, pad 15) -> BatchNorm1d (train: batch statistics over all
//       B*T' frames) -> SiLU, forward and backward.  HBM-bound: one CTA = one trial x 32
//       channels x a time tile staged in shared memory with its halo.
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

constexpr int DW_MAX_K = 33;     // depthwise kernel length supported (31 on the path)
constexpr int DW_TT = 128;       // output frames per time tile
constexpr int DW_CH = 32;        // channels per CTA (= warp width, lane = channel)
constexpr int DW_ROWS = DW_TT + DW_MAX_K - 1;

// cos/sin tables of Wav2Vec2ConformerRotaryPositionalEmbedding (:125-156): emb = cat(f, f), f = t * inv_freq
__global__ void rope_table_kernel(const float* __restrict__ inv_freq, float* __restrict__ cos_t, float* __restrict__ sin_t,
                                  int T, int half) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T * half) return;
    const int t = i / half, j = i - t * half;
    const float f = (float)t * inv_freq[j];
    const float c = cosf(f), s = sinf(f);
    cos_t[t * 2 * half + j] = c;
    cos_t[t * 2 * half + half + j] = c;
    sin_t[t * 2 * half + j] = s;
    sin_t[t * 2 * half + half + j] = s;
}

// one warp per row; the normalised row is parked in shared memory so that the rotary partner
// (p +- hd/2 inside the head) can be read back whatever the head size
template <int MAXP>
__global__ void __launch_bounds__(256) ln_rope_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps,
                                                          bf16* __restrict__ y_hi, bf16* __restrict__ y_lo,
                                                          bf16* __restrict__ r_hi, bf16* __restrict__ r_lo,
                                                          float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                          const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                                          int rows, int n, int T, int hd) {
    B2T_PDL_TRIGGER();
    extern __shared__ float srow[];   // [warps][n]
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    float* buf = srow + (size_t)wib * n;
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const float inv_n = 1.0f / (float)n;
    const int half = hd >> 1;
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
            mean_out[row] = mean;
            rstd_out[row] = rstd;
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            if (c < n) {
                v[i] = (v[i] - mean) * rstd * gamma[c] + beta[c];
                buf[c] = v[i];
            }
        }
        __syncwarp();
        const int t = row % T;
#pragma unroll
        for (int i = 0; i < MAXP; ++i) {
            const int c = lane + 32 * i;
            if (c < n) {
                const long off = (long)row * n + c;
                bf16 h, l;
                split_bf16(v[i], h, l);
                y_hi[off] = h;
                if (y_lo) y_lo[off] = l;
                const int p = c % hd;
                const float rot = p < half ? -buf[c + half] : buf[c - half];
                const float o = v[i] * cos_t[t * hd + p] + rot * sin_t[t * hd + p];
                split_bf16(o, h, l);
                r_hi[off] = h;
                if (r_lo) r_lo[off] = l;
            }
        }
        __syncwarp();
    }
}

// float4 variants of the two rotary kernels (n % 4 == 0, head_dim % 8 == 0, 16-byte aligned rows).  The scalar kernels
// above / below moved one element per lane and instruction - 2-byte bf16 stores, an integer division and two modulos per
// element - and ran at 0.23 / 0.33 of HBM in the conformer step (31 / 29 us for [3776, 1024]).
// Two warps per row (as ln_fwd_vec2_kernel): VPL float4 per lane, the normalised row parked in smem for the rotary partner.
template <int VPL>
__global__ void __launch_bounds__(256) ln_rope_fwd_vec_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, float eps,
                                                              bf16* __restrict__ y_hi, bf16* __restrict__ y_lo,
                                                              bf16* __restrict__ r_hi, bf16* __restrict__ r_lo,
                                                              float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                              const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                                              int rows, int n, int T, int hd) {
    B2T_PDL_TRIGGER();
    extern __shared__ float srow[];   // [4 rows][n]
    __shared__ float red[2][4][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = warp >> 1, hf = warp & 1;
    const int row = blockIdx.x * 4 + rl;
    const bool row_ok = row < rows;
    float* buf = srow + (size_t)rl * n;
    const float inv_n = 1.0f / (float)n;
    const int half = hd >> 1;
    float4 v[VPL];
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((hf * VPL + i) * 32 + lane) * 4;
        v[i] = (row_ok && c < n) ? *reinterpret_cast<const float4*>(x + (long)row * n + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
    s = warp_sum(s);
    if (lane == 0) red[0][rl][hf] = s;
    __syncthreads();
    const float mean = (red[0][rl][0] + red[0][rl][1]) * inv_n;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        if (((hf * VPL + i) * 32 + lane) * 4 < n) {
            const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
    }
    q = warp_sum(q);
    if (lane == 0) red[1][rl][hf] = q;
    __syncthreads();
    const float rstd = rsqrtf((red[1][rl][0] + red[1][rl][1]) * inv_n + eps);
    if (row_ok && hf == 0 && lane == 0) {
        mean_out[row] = mean;
        rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((hf * VPL + i) * 32 + lane) * 4;
        if (c < n) {
            const float4 ga = *reinterpret_cast<const float4*>(gamma + c), be = *reinterpret_cast<const float4*>(beta + c);
            v[i] = make_float4((v[i].x - mean) * rstd * ga.x + be.x, (v[i].y - mean) * rstd * ga.y + be.y,
                               (v[i].z - mean) * rstd * ga.z + be.z, (v[i].w - mean) * rstd * ga.w + be.w);
            *reinterpret_cast<float4*>(buf + c) = v[i];
        }
    }
    __syncthreads();   // the partner of a column may have been normalised by the row's other warp
    if (!row_ok) return;
    const int t = row % T;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = ((hf * VPL + i) * 32 + lane) * 4;
        if (c < n) {
            const long off = (long)row * n + c;
            const float o[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
            store_split4(o, y_hi, y_lo, off);
            const int p = c % hd;   // 4 consecutive columns stay inside one half of the head (half % 4 == 0)
            const float4 pr = *reinterpret_cast<const float4*>(buf + (p < half ? c + half : c - half));
            const float sg = p < half ? -1.0f : 1.0f;
            const float4 cs = *reinterpret_cast<const float4*>(cos_t + (long)t * hd + p);
            const float4 sn = *reinterpret_cast<const float4*>(sin_t + (long)t * hd + p);
            const float r[4] = {o[0] * cs.x + sg * pr.x * sn.x, o[1] * cs.y + sg * pr.y * sn.y, o[2] * cs.z + sg * pr.z * sn.z,
                                o[3] * cs.w + sg * pr.w * sn.w};
            store_split4(r, r_hi, r_lo, off);
        }
    }
}

// one CTA per row group: thread = 4 consecutive columns
__global__ void __launch_bounds__(256) rope_bwd_add_vec_kernel(const float* __restrict__ d_rot, long ld_rot,
                                                               const float* __restrict__ d_plain, long ld_plain,
                                                               float* __restrict__ out, const float* __restrict__ cos_t,
                                                               const float* __restrict__ sin_t, long rows, int n, int T, int hd) {
    B2T_PDL_TRIGGER();
    const int half = hd >> 1;
    const int n4 = n >> 2;
    for (long row = blockIdx.x; row < rows; row += gridDim.x) {
        const int t = (int)(row % T);
        const float* dr = d_rot + row * ld_rot;
        const float* cs = cos_t + (long)t * hd;
        const float* sn = sin_t + (long)t * hd;
        for (int c4 = threadIdx.x; c4 < n4; c4 += blockDim.x) {
            const int c = c4 * 4;
            const int p = c % hd;
            const bool lo = p < half;
            const float4 a = *reinterpret_cast<const float4*>(dr + c);
            const float4 b = *reinterpret_cast<const float4*>(dr + (lo ? c + half : c - half));
            const float4 cc = *reinterpret_cast<const float4*>(cs + p);
            const float4 ss = *reinterpret_cast<const float4*>(sn + (lo ? p + half : p - half));
            const float sg = lo ? 1.0f : -1.0f;
            float4 o = make_float4(a.x * cc.x + sg * b.x * ss.x, a.y * cc.y + sg * b.y * ss.y, a.z * cc.z + sg * b.z * ss.z,
                                   a.w * cc.w + sg * b.w * ss.w);
            if (d_plain) {
                const float4 d = *reinterpret_cast<const float4*>(d_plain + row * ld_plain + c);
                o.x += d.x, o.y += d.y, o.z += d.z, o.w += d.w;
            }
            *reinterpret_cast<float4*>(out + row * n + c) = o;
        }
    }
}

// du = d_rot * cos + R^T(d_rot * sin) + d_plain,  R^T(y)[p] = y[p + half] (p < half), -y[p - half] (p >= half)
__global__ void rope_bwd_add_kernel(const float* __restrict__ d_rot, long ld_rot, const float* __restrict__ d_plain,
                                    long ld_plain, float* __restrict__ out, const float* __restrict__ cos_t,
                                    const float* __restrict__ sin_t, long rows, int n, int T, int hd) {
    B2T_PDL_TRIGGER();
    const int half = hd >> 1;
    const long total = rows * n;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long row = i / n;
        const int c = (int)(i - row * n);
        const int t = (int)(row % T);
        const int p = c % hd;
        const float* dr = d_rot + row * ld_rot;
        const float* cs = cos_t + (long)t * hd;
        const float* sn = sin_t + (long)t * hd;
        float o = dr[c] * cs[p];
        o += p < half ? dr[c + half] * sn[p + half] : -dr[c - half] * sn[p - half];
        if (d_plain) o += d_plain[row * ld_plain + c];
        out[i] = o;
    }
}

// ex2.approx + rcp.approx (~1e-7 relative): the IEEE division's slow-path branch serialised the unrolled row batches
// ex2.approx.ftz: no denormal-range fix-up code around the MUFU (a flushed result gives sigmoid = 1, inf gives 0)
__device__ __forceinline__ float ex2_ftz(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sigmoid_acc(float x) { return rcp_approx(1.0f + ex2_ftz(-1.4426950408889634f * x)); }
__device__ __forceinline__ float silu_acc(float x) { return x * sigmoid_acc(x); }
__device__ __forceinline__ float silu_acc_grad(float x) {
    const float s = sigmoid_acc(x);
    return s * (1.0f + x * (1.0f - s));
}

// GLU + depthwise conv forward + per-channel sum / sum of squares (BatchNorm batch statistics).
// g [B*T, 2D] (pointwise_conv1 output: value | gate), w [D, K], out [B*T, D].
// grid (ceil(D/32), B, ceil(T/DW_TT)), 256 threads: lane = channel, warp = frame phase.
__global__ void __launch_bounds__(256) glu_dwconv_fwd_kernel(const float* __restrict__ g, const float* __restrict__ w,
                                                             float* __restrict__ out, double* __restrict__ ssum,
                                                             double* __restrict__ ssq, int T, int D, int K) {
    B2T_PDL_TRIGGER();
    __shared__ float s[DW_ROWS][DW_CH];
    __shared__ float sw[DW_MAX_K][DW_CH];
    __shared__ float red[2][8][DW_CH];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * DW_CH + lane;
    const int b = blockIdx.y;
    const int t0 = blockIdx.z * DW_TT;
    const int pad = (K - 1) / 2;
    const bool c_ok = c < D;
    for (int k = warp; k < K; k += 8) sw[k][lane] = c_ok ? w[(long)c * K + k] : 0.0f;
    const int nrows = DW_TT + K - 1;
    // rows in batches of 4 per warp with every global load issued before the first sigmoid (as one loop the
    // 19 iterations of a warp each waited for their own two loads)
    for (int rb = warp; rb < nrows; rb += 32) {
        float ga[4], gb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = rb + 8 * j, t = t0 - pad + r;
            const bool ok = c_ok && r < nrows && t >= 0 && t < T;
            const float* gr = g + ((long)b * T + (ok ? t : 0)) * 2 * D;
            ga[j] = ok ? gr[c] : 0.0f;
            gb[j] = ok ? gr[D + c] : 0.0f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = rb + 8 * j;
            if (r < nrows) s[r][lane] = ga[j] * sigmoid_acc(gb[j]);
        }
    }
    __syncthreads();
    float sum = 0.0f, sq = 0.0f;
    if (K == 31) {
        // register-blocked over time (the path's kernel length): a thread keeps the 31 taps and computes 8
        // consecutive frames from 38 window values - 69 smem loads per 8 outputs instead of 496; taps are still
        // applied in k = 0..30 order per output (same rounding as the plain loop)
        constexpr int KK = 31, FR = 8;
        float wk[KK];
#pragma unroll
        for (int k = 0; k < KK; ++k) wk[k] = sw[k][lane];
        for (int rg = warp * FR; rg < DW_TT && t0 + rg < T; rg += 8 * FR) {
            float acc[FR];
#pragma unroll
            for (int o = 0; o < FR; ++o) acc[o] = 0.0f;
#pragma unroll
            for (int j = 0; j < FR + KK - 1; ++j) {
                const float v = s[rg + j][lane];
#pragma unroll
                for (int o = 0; o < FR; ++o) {
                    const int k = j - o;
                    if (k >= 0 && k < KK) acc[o] = fmaf(wk[k], v, acc[o]);
                }
            }
#pragma unroll
            for (int o = 0; o < FR; ++o) {
                if (t0 + rg + o < T) {
                    if (c_ok) out[((long)b * T + t0 + rg + o) * D + c] = acc[o];
                    sum += acc[o];
                    sq += acc[o] * acc[o];
                }
            }
        }
    } else {
        for (int r = warp; r < DW_TT && t0 + r < T; r += 8) {
            float acc = 0.0f;
            for (int k = 0; k < K; ++k) acc = fmaf(sw[k][lane], s[r + k][lane], acc);
            if (c_ok) out[((long)b * T + t0 + r) * D + c] = acc;
            sum += acc;
            sq += acc * acc;
        }
    }
    if (ssum) {
        red[0][warp][lane] = sum;
        red[1][warp][lane] = sq;
        __syncthreads();
        if (warp == 0 && c_ok) {
            double a = 0.0, q = 0.0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                a += (double)red[0][j][lane];
                q += (double)red[1][j][lane];
            }
            atomicAdd(ssum + c, a);
            atomicAdd(ssq + c, q);
        }
    }
}

// BatchNorm1d (train: batch statistics, running-stat update; eval: running statistics) + SiLU,
// output as bf16 hi/lo operand of pointwise_conv2.  thread = channel, block rows [by*RPB, ...).
constexpr int BN_RPB = 32;
__global__ void __launch_bounds__(128) bn_silu_fwd_kernel(const float* __restrict__ x, const double* __restrict__ ssum,
                                                          const double* __restrict__ ssq, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps, float momentum,
                                                          float* __restrict__ running_mean, float* __restrict__ running_var,
                                                          long long* __restrict__ nbt, float* __restrict__ mean_out,
                                                          float* __restrict__ rstd_out, bf16* __restrict__ y_hi,
                                                          bf16* __restrict__ y_lo, long M, int D, int training) {
    const int c = blockIdx.x * 128 + threadIdx.x;
    if (c >= D) return;
    float mean, rstd;
    if (training) {
        const double m = ssum[c] / (double)M;
        double var = ssq[c] / (double)M - m * m;
        var = var < 0.0 ? 0.0 : var;
        mean = (float)m;
        rstd = (float)(1.0 / sqrt(var + (double)eps));
        if (blockIdx.y == 0) {
            const double unb = M > 1 ? var * (double)M / (double)(M - 1) : var;
            running_mean[c] = (1.0f - momentum) * running_mean[c] + momentum * mean;
            running_var[c] = (1.0f - momentum) * running_var[c] + momentum * (float)unb;
            if (c == 0 && nbt) *nbt += 1;
        }
    } else {
        mean = running_mean[c];
        rstd = 1.0f / sqrtf(running_var[c] + eps);
    }
    if (blockIdx.y == 0) {
        mean_out[c] = mean;
        rstd_out[c] = rstd;
    }
    const float ga = gamma[c] * rstd, be = beta[c] - mean * gamma[c] * rstd;
    const long r0 = (long)blockIdx.y * BN_RPB;
    const long r1 = r0 + BN_RPB < M ? r0 + BN_RPB : M;
    for (long rb = r0; rb < r1; rb += 8) {   // 8 rows per batch, loads first
        float xv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) xv[j] = rb + j < r1 ? x[(rb + j) * D + c] : 0.0f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (rb + j >= r1) break;
            const long off = (rb + j) * D + c;
            const float y = silu_acc(fmaf(xv[j], ga, be));
            bf16 h, l;
            split_bf16(y, h, l);
            y_hi[off] = h;
            if (y_lo) y_lo[off] = l;
        }
    }
}

// backward, pass 1: dbn = ds * silu'(bn(x));  dbeta[c] += sum dbn,  dgamma[c] += sum dbn * xhat
__global__ void __launch_bounds__(128) bn_silu_bwd_stats_kernel(const float* __restrict__ ds, const float* __restrict__ x,
                                                                const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                float* __restrict__ dgamma, float* __restrict__ dbeta, long M,
                                                                int D) {
    const int c = blockIdx.x * 128 + threadIdx.x;
    if (c >= D) return;
    const float mu = mean[c], rs = rstd[c], ga = gamma[c], be = beta[c];
    const long r0 = (long)blockIdx.y * BN_RPB;
    const long r1 = r0 + BN_RPB < M ? r0 + BN_RPB : M;
    float sb = 0.0f, sg = 0.0f;
    for (long rb = r0; rb < r1; rb += 8) {   // 8 rows per batch, all 16 loads first
        float xv[8], dv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const bool ok = rb + j < r1;
            xv[j] = ok ? x[(rb + j) * D + c] : 0.0f;
            dv[j] = ok ? ds[(rb + j) * D + c] : 0.0f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (rb + j >= r1) break;
            const float xh = (xv[j] - mu) * rs;
            const float dbn = dv[j] * silu_acc_grad(fmaf(xh, ga, be));
            sb += dbn;
            sg += dbn * xh;
        }
    }
    atomicAdd(dbeta + c, sb);
    atomicAdd(dgamma + c, sg);
}

// float4 forms of the two kernels above (D % 4 == 0): lane = 4 channels, warp = 8 rows in one batch of 16-byte loads,
// CTA = 128 channels x 32 rows; 8-byte bf16 stores; the backward sums its 4 warps in smem before the atomics.
__global__ void __launch_bounds__(128) bn_silu_fwd_vec_kernel(const float* __restrict__ x, const double* __restrict__ ssum,
                                                              const double* __restrict__ ssq, const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, float eps, float momentum,
                                                              float* __restrict__ running_mean, float* __restrict__ running_var,
                                                              long long* __restrict__ nbt, float* __restrict__ mean_out,
                                                              float* __restrict__ rstd_out, bf16* __restrict__ y_hi,
                                                              bf16* __restrict__ y_lo, long M, int D, int training) {
    B2T_PDL_TRIGGER();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * 128 + lane * 4;
    if (c >= D) return;
    const long r0 = (long)blockIdx.y * BN_RPB + warp * 8;
    float4 xv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
        xv[j] = r0 + j < M ? *reinterpret_cast<const float4*>(x + (r0 + j) * D + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    const bool writer = blockIdx.y == 0 && warp == 0;
    float ga[4], be[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float mean, rstd;
        if (training) {
            const double m = ssum[c + q] / (double)M;
            double var = ssq[c + q] / (double)M - m * m;
            var = var < 0.0 ? 0.0 : var;
            mean = (float)m;
            rstd = (float)(1.0 / sqrt(var + (double)eps));
            if (writer) {
                const double unb = M > 1 ? var * (double)M / (double)(M - 1) : var;
                running_mean[c + q] = (1.0f - momentum) * running_mean[c + q] + momentum * mean;
                running_var[c + q] = (1.0f - momentum) * running_var[c + q] + momentum * (float)unb;
                if (c + q == 0 && nbt) *nbt += 1;
            }
        } else {
            mean = running_mean[c + q];
            rstd = 1.0f / sqrtf(running_var[c + q] + eps);
        }
        if (writer) {
            mean_out[c + q] = mean;
            rstd_out[c + q] = rstd;
        }
        ga[q] = gamma[c + q] * rstd;
        be[q] = beta[c + q] - mean * gamma[c + q] * rstd;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (r0 + j >= M) break;
        const float o[4] = {silu_acc(fmaf(xv[j].x, ga[0], be[0])), silu_acc(fmaf(xv[j].y, ga[1], be[1])),
                            silu_acc(fmaf(xv[j].z, ga[2], be[2])), silu_acc(fmaf(xv[j].w, ga[3], be[3]))};
        store_split4(o, y_hi, y_lo, (r0 + j) * D + c);
    }
}

__global__ void __launch_bounds__(128) bn_silu_bwd_stats_vec_kernel(const float* __restrict__ ds, const float* __restrict__ x,
                                                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                    float* __restrict__ dgamma, float* __restrict__ dbeta, long M,
                                                                    int D) {
    B2T_PDL_TRIGGER();
    __shared__ float red[2][4][128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * 128 + lane * 4;
    const bool c_ok = c < D;
    const long r0 = (long)blockIdx.y * BN_RPB + warp * 8;
    float4 xv[8], dv[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const bool ok = c_ok && r0 + j < M;
        xv[j] = ok ? *reinterpret_cast<const float4*>(x + (r0 + j) * D + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        dv[j] = ok ? *reinterpret_cast<const float4*>(ds + (r0 + j) * D + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float mu[4], rs[4], ga[4], be[4], sb[4] = {0.f, 0.f, 0.f, 0.f}, sg[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        mu[q] = c_ok ? mean[c + q] : 0.f, rs[q] = c_ok ? rstd[c + q] : 0.f;
        ga[q] = c_ok ? gamma[c + q] : 0.f, be[q] = c_ok ? beta[c + q] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {   // masked rows carry ds = 0
        const float xs[4] = {xv[j].x, xv[j].y, xv[j].z, xv[j].w}, dd4[4] = {dv[j].x, dv[j].y, dv[j].z, dv[j].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float xh = (xs[q] - mu[q]) * rs[q];
            const float dbn = dd4[q] * silu_acc_grad(fmaf(xh, ga[q], be[q]));
            sb[q] += dbn;
            sg[q] = fmaf(dbn, xh, sg[q]);
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        red[0][warp][lane * 4 + q] = sb[q];
        red[1][warp][lane * 4 + q] = sg[q];
    }
    __syncthreads();
    const int cc = blockIdx.x * 128 + threadIdx.x;
    if (cc < D) {
        const int t = threadIdx.x;
        atomicAdd(dbeta + cc, (red[0][0][t] + red[0][1][t]) + (red[0][2][t] + red[0][3][t]));
        atomicAdd(dgamma + cc, (red[1][0][t] + red[1][1][t]) + (red[1][2][t] + red[1][3][t]));
    }
}

// backward, pass 2 (fused): BatchNorm+SiLU backward -> depthwise-conv backward (input and weight
// gradients) -> GLU backward, written as the bf16 hi/lo operand dg [B*T, 2D] of the
// pointwise_conv1 wgrad / dgrad GEMMs.  Same tiling as the forward kernel.
__global__ void __launch_bounds__(256) glu_dwconv_bwd_kernel(const float* __restrict__ ds, const float* __restrict__ x,
                                                             const float* __restrict__ g, const float* __restrict__ w,
                                                             const float* __restrict__ mean, const float* __restrict__ rstd,
                                                             const float* __restrict__ gamma, const float* __restrict__ beta,
                                                             const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                                                             bf16* __restrict__ dg_hi, bf16* __restrict__ dg_lo,
                                                             float* __restrict__ dw, int T, int D, int K, long M,
                                                             int training) {
    B2T_PDL_TRIGGER();
    extern __shared__ float dyn[];
    float (*dd)[DW_CH] = reinterpret_cast<float (*)[DW_CH]>(dyn);                       // d(dwconv out), with halo
    float (*s)[DW_CH] = reinterpret_cast<float (*)[DW_CH]>(dyn + DW_ROWS * DW_CH);      // glu, with halo
    float (*sw)[DW_CH] = reinterpret_cast<float (*)[DW_CH]>(dyn + 2 * DW_ROWS * DW_CH);
    float (*acc_w)[DW_CH] = reinterpret_cast<float (*)[DW_CH]>(dyn + (2 * DW_ROWS + DW_MAX_K) * DW_CH);
    // GLU value and gate sigmoid of the tile's own frames, kept for the GLU backward at the end (no second read of g)
    float (*sa)[DW_CH] = reinterpret_cast<float (*)[DW_CH]>(dyn + (2 * DW_ROWS + 2 * DW_MAX_K) * DW_CH);
    float (*ssg)[DW_CH] = reinterpret_cast<float (*)[DW_CH]>(dyn + (2 * DW_ROWS + 2 * DW_MAX_K + DW_TT) * DW_CH);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * DW_CH + lane;
    const int b = blockIdx.y;
    const int t0 = blockIdx.z * DW_TT;
    const int pad = (K - 1) / 2;
    const bool c_ok = c < D;
    // per-CTA base pointers, 32-bit offsets inside one trial (T * 2D < 2^31)
    const float* __restrict__ xb = x + (long)b * T * D + c;
    const float* __restrict__ dsb = ds + (long)b * T * D + c;
    const float* __restrict__ gb = g + (long)b * T * 2 * D + c;
    for (int k = warp; k < K; k += 8) {
        sw[k][lane] = c_ok ? w[(long)c * K + k] : 0.0f;
        acc_w[k][lane] = 0.0f;
    }
    float mu = 0.0f, rs = 0.0f, ga = 0.0f, be = 0.0f, mdb = 0.0f, mdg = 0.0f;
    if (c_ok) {
        mu = mean[c], rs = rstd[c], ga = gamma[c], be = beta[c];
        if (training) {
            mdb = dbeta[c] / (float)M;
            mdg = dgamma[c] / (float)M;
        }
    }
    // rows r <-> frame t = t0 - (K-1-pad) + r for dd (needs [t - (K-1-pad), t + pad]) ; keep one
    // common origin t0 - pad' with pad' = K - 1 - pad... for odd K pad' == pad.
    const int nrows = DW_TT + K - 1;
    for (int rb = warp; rb < nrows; rb += 32) {   // 4 rows per warp and batch, all 16 loads first
        float xv[4], dsv4[4], g0[4], g1[4];
        bool okv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = rb + 8 * j, t = t0 - pad + r;
            okv[j] = c_ok && r < nrows && t >= 0 && t < T;
            const int to = okv[j] ? t : 0;
            xv[j] = okv[j] ? xb[to * D] : 0.0f;
            dsv4[j] = okv[j] ? dsb[to * D] : 0.0f;
            g0[j] = okv[j] ? gb[to * 2 * D] : 0.0f;
            g1[j] = okv[j] ? gb[to * 2 * D + D] : 0.0f;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {   // branch-free: masked frames carry zeros through the math
            const int r = rb + 8 * j;
            const float xh = (xv[j] - mu) * rs;
            const float dbn = dsv4[j] * silu_acc_grad(fmaf(xh, ga, be));
            const float dv = okv[j] ? ga * rs * (dbn - mdb - xh * mdg) : 0.0f;
            const float sg = sigmoid_acc(g1[j]);
            if (r < nrows) {   // warp-uniform
                dd[r][lane] = dv;
                s[r][lane] = g0[j] * sg;
                const int rc = r - pad;
                if (rc >= 0 && rc < DW_TT) {
                    sa[rc][lane] = g0[j];
                    ssg[rc][lane] = sg;
                }
            }
        }
    }
    __syncthreads();
    // out[t] = sum_k w[k] s[t + k - pad]  =>  d s[u] = sum_k w[k] d out[u - k + pad]
    //                                        d w[k] = sum_t d out[t] s[t + k - pad]
    float wacc[DW_MAX_K];
#pragma unroll
    for (int k = 0; k < DW_MAX_K; ++k) wacc[k] = 0.0f;
    bf16* __restrict__ dgh = dg_hi + ((long)b * T + t0) * 2 * D + c;
    bf16* __restrict__ dgl = dg_lo ? dg_lo + ((long)b * T + t0) * 2 * D + c : nullptr;
    auto emit = [&](int r, float dsv) {
        const float a = sa[r][lane], sg = ssg[r][lane];
        const float da = dsv * sg, dgate = dsv * a * sg * (1.0f - sg);
        const int off = r * 2 * D;
        bf16 h, l;
        split_bf16(da, h, l);
        dgh[off] = h;
        if (dgl) dgl[off] = l;
        split_bf16(dgate, h, l);
        dgh[off + D] = h;
        if (dgl) dgl[off + D] = l;
    };
    if (K == 31) {
        // register-blocked over 8 consecutive frames (see the forward kernel): the taps and the 31 weight-gradient
        // accumulators stay in registers, each window value of dd / s is read once per 8 outputs
        constexpr int KK = 31, FR = 8, PD = 15;
        float wk[KK];
#pragma unroll
        for (int k = 0; k < KK; ++k) wk[k] = sw[k][lane];
        for (int rg = warp * FR; rg < DW_TT && t0 + rg < T; rg += 8 * FR) {
            float dsv[FR], dout[FR];
#pragma unroll
            for (int o = 0; o < FR; ++o) {
                dsv[o] = 0.0f;
                dout[o] = t0 + rg + o < T ? dd[rg + o + PD][lane] : 0.0f;
            }
#pragma unroll
            for (int j = FR + KK - 2; j >= 0; --j) {   // descending j = ascending tap index per output
                const float v = dd[rg + j][lane];
#pragma unroll
                for (int o = 0; o < FR; ++o) {
                    const int k = o + 2 * PD - j;
                    if (k >= 0 && k < KK) dsv[o] = fmaf(wk[k], v, dsv[o]);
                }
            }
#pragma unroll
            for (int j = 0; j < FR + KK - 1; ++j) {
                const float v = s[rg + j][lane];
#pragma unroll
                for (int o = 0; o < FR; ++o) {
                    const int k = j - o;
                    if (k >= 0 && k < KK) wacc[k] = fmaf(dout[o], v, wacc[k]);
                }
            }
            if (c_ok) {
#pragma unroll
                for (int o = 0; o < FR; ++o)
                    if (t0 + rg + o < T) emit(rg + o, dsv[o]);
            }
        }
    } else {
        for (int r = warp; r < DW_TT && t0 + r < T; r += 8) {
            // frame t = t0 + r sits at smem row r + pad
            float dsv = 0.0f;
            const float dout = dd[r + pad][lane];
#pragma unroll
            for (int k = 0; k < DW_MAX_K; ++k) {
                if (k < K) {
                    dsv = fmaf(sw[k][lane], dd[r + 2 * pad - k][lane], dsv);
                    wacc[k] = fmaf(dout, s[r + k][lane], wacc[k]);
                }
            }
            if (c_ok) emit(r, dsv);
        }
    }
#pragma unroll
    for (int k = 0; k < DW_MAX_K; ++k)
        if (k < K) atomicAdd(&acc_w[k][lane], wacc[k]);
    __syncthreads();
    if (c_ok)
        for (int k = warp; k < K; k += 8) atomicAdd(dw + (long)c * K + k, acc_w[k][lane]);
}

constexpr size_t DW_BWD_SMEM = (size_t)(2 * DW_ROWS + 2 * DW_MAX_K + 2 * DW_TT) * DW_CH * sizeof(float);

inline int grid_for(long n, int block, int cap_mult = 16) {
    long g = (n + block - 1) / block;
    const long cap = 148L * cap_mult;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

extern "C" int b2t_rope_table(const float* inv_freq, float* cos_t, float* sin_t, int T, int head_dim, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(inv_freq && cos_t && sin_t && T > 0 && head_dim > 0 && head_dim % 2 == 0, "b2t_rope_table: bad arguments");
    const int half = head_dim / 2;
    rope_table_kernel<<<(T * half + 127) / 128, 128, 0, stream>>>(inv_freq, cos_t, sin_t, T, half);
    B2T_CHECK_LAUNCH("b2t_rope_table");
    return 0;
}

extern "C" int b2t_layernorm_rope_fwd(const float* x, const float* gamma, const float* beta, float eps, void* y_hi, void* y_lo,
                                      void* r_hi, void* r_lo, float* mean, float* rstd, const float* cos_t,
                                      const float* sin_t, int rows, int n, int T, int head_dim, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && gamma && beta && y_hi && r_hi && mean && rstd && cos_t && sin_t, "b2t_layernorm_rope_fwd: null argument");
    B2T_CHECK_ARG(rows > 0 && n > 0 && n <= 1024 && T > 0 && rows % T == 0, "b2t_layernorm_rope_fwd: bad sizes rows=%d n=%d T=%d",
                  rows, n, T);
    B2T_CHECK_ARG(head_dim > 0 && head_dim % 2 == 0 && n % head_dim == 0, "b2t_layernorm_rope_fwd: head_dim %d must be even and divide n", head_dim);
    B2T_CHECK_ARG((y_lo == nullptr) == (r_lo == nullptr), "b2t_layernorm_rope_fwd: y_lo / r_lo must both be set or both NULL");
    {
        auto al = [](const void* q, int b) { return q == nullptr || ((uintptr_t)q & (b - 1)) == 0; };
        if (n % 4 == 0 && n >= 256 && head_dim % 8 == 0 && al(x, 16) && al(gamma, 16) && al(beta, 16) && al(cos_t, 16) && al(sin_t, 16) &&
            al(y_hi, 8) && al(y_lo, 8) && al(r_hi, 8) && al(r_lo, 8)) {
            const int v2 = (n / 4 + 63) / 64, g2 = (rows + 3) / 4;
            const size_t sm = 4 * (size_t)n * sizeof(float);
#define VCALL(V)                                                                                                        \
    ln_rope_fwd_vec_kernel<V><<<g2, 256, sm, stream>>>(x, gamma, beta, eps, (bf16*)y_hi, (bf16*)y_lo, (bf16*)r_hi, (bf16*)r_lo, \
                                                       mean, rstd, cos_t, sin_t, rows, n, T, head_dim)
            if (v2 <= 1) { VCALL(1); } else if (v2 <= 2) { VCALL(2); } else if (v2 <= 3) { VCALL(3); } else { VCALL(4); }
#undef VCALL
            B2T_CHECK_LAUNCH("b2t_layernorm_rope_fwd");
            return 0;
        }
    }
    const int grid = grid_for((long)rows * 32, 256, 8);
    const size_t smem = 8 * (size_t)n * sizeof(float);
    const int per = (n + 31) / 32;
#define CALL(P)                                                                                                         \
    ln_rope_fwd_kernel<P><<<grid, 256, smem, stream>>>(x, gamma, beta, eps, (bf16*)y_hi, (bf16*)y_lo, (bf16*)r_hi, (bf16*)r_lo, \
                                                       mean, rstd, cos_t, sin_t, rows, n, T, head_dim)
    if (per <= 2) { CALL(2); } else if (per <= 8) { CALL(8); } else if (per <= 24) { CALL(24); } else { CALL(32); }
#undef CALL
    B2T_CHECK_LAUNCH("b2t_layernorm_rope_fwd");
    return 0;
}

extern "C" int b2t_rope_bwd_add(const float* d_rot, int64_t ld_rot, const float* d_plain, int64_t ld_plain, float* out,
                                const float* cos_t, const float* sin_t, int64_t rows, int n, int T, int head_dim,
                                b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(d_rot && out && cos_t && sin_t && rows > 0 && n > 0 && T > 0 && head_dim % 2 == 0 && n % head_dim == 0,
                  "b2t_rope_bwd_add: bad arguments");
    {
        auto al16 = [](const void* q) { return q == nullptr || ((uintptr_t)q & 15) == 0; };
        if (n % 4 == 0 && head_dim % 8 == 0 && ld_rot % 4 == 0 && (d_plain == nullptr || ld_plain % 4 == 0) && al16(d_rot) &&
            al16(d_plain) && al16(out) && al16(cos_t) && al16(sin_t)) {
            const long g = rows < 148 * 8 ? rows : 148 * 8;
            rope_bwd_add_vec_kernel<<<(unsigned)g, n / 4 < 256 ? ((n / 4 + 31) / 32) * 32 : 256, 0, stream>>>(
                d_rot, ld_rot, d_plain, ld_plain, out, cos_t, sin_t, rows, n, T, head_dim);
            B2T_CHECK_LAUNCH("b2t_rope_bwd_add");
            return 0;
        }
    }
    rope_bwd_add_kernel<<<grid_for(rows * n, 256), 256, 0, stream>>>(d_rot, ld_rot, d_plain, ld_plain, out, cos_t, sin_t, rows, n,
                                                                    T, head_dim);
    B2T_CHECK_LAUNCH("b2t_rope_bwd_add");
    return 0;
}

extern "C" int b2t_glu_dwconv_fwd(const float* g, const float* w, float* out, double* stats, int B, int T, int D, int K,
                                  b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(g && w && out && B > 0 && T > 0 && D > 0, "b2t_glu_dwconv_fwd: bad arguments");
    B2T_CHECK_ARG(K > 0 && K <= DW_MAX_K && K % 2 == 1, "b2t_glu_dwconv_fwd: depthwise kernel %d must be odd and <= %d", K, DW_MAX_K);
    dim3 grid((D + DW_CH - 1) / DW_CH, B, (T + DW_TT - 1) / DW_TT);
    glu_dwconv_fwd_kernel<<<grid, 256, 0, stream>>>(g, w, out, stats, stats ? stats + D : nullptr, T, D, K);
    B2T_CHECK_LAUNCH("b2t_glu_dwconv_fwd");
    return 0;
}

extern "C" int b2t_bn_silu_fwd(const float* x, const double* stats, const float* gamma, const float* beta, float eps,
                               float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                               float* mean_out, float* rstd_out, void* y_hi, void* y_lo, int64_t M, int D, int training,
                               b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && gamma && beta && running_mean && running_var && mean_out && rstd_out && y_hi && M > 0 && D > 0,
                  "b2t_bn_silu_fwd: bad arguments");
    B2T_CHECK_ARG(!training || stats, "b2t_bn_silu_fwd: training mode needs the batch statistics buffer");
    dim3 grid((D + 127) / 128, (unsigned)((M + BN_RPB - 1) / BN_RPB));
    const bool vec = D % 4 == 0 && ((uintptr_t)x % 16 == 0) && ((uintptr_t)y_hi % 8 == 0) && (!y_lo || (uintptr_t)y_lo % 8 == 0);
    if (vec)
        bn_silu_fwd_vec_kernel<<<grid, 128, 0, stream>>>(x, stats, stats ? stats + D : nullptr, gamma, beta, eps, momentum,
                                                         running_mean, running_var, (long long*)num_batches_tracked, mean_out,
                                                         rstd_out, (bf16*)y_hi, (bf16*)y_lo, M, D, training);
    else
        bn_silu_fwd_kernel<<<grid, 128, 0, stream>>>(x, stats, stats ? stats + D : nullptr, gamma, beta, eps, momentum, running_mean,
                                                     running_var, (long long*)num_batches_tracked, mean_out, rstd_out, (bf16*)y_hi,
                                                     (bf16*)y_lo, M, D, training);
    B2T_CHECK_LAUNCH("b2t_bn_silu_fwd");
    return 0;
}

extern "C" int b2t_bn_silu_bwd_stats(const float* ds, const float* x, const float* mean, const float* rstd, const float* gamma,
                                     const float* beta, float* dgamma, float* dbeta, int64_t M, int D, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(ds && x && mean && rstd && gamma && beta && dgamma && dbeta && M > 0 && D > 0, "b2t_bn_silu_bwd_stats: bad arguments");
    dim3 grid((D + 127) / 128, (unsigned)((M + BN_RPB - 1) / BN_RPB));
    if (D % 4 == 0 && (uintptr_t)x % 16 == 0 && (uintptr_t)ds % 16 == 0)
        bn_silu_bwd_stats_vec_kernel<<<grid, 128, 0, stream>>>(ds, x, mean, rstd, gamma, beta, dgamma, dbeta, M, D);
    else
        bn_silu_bwd_stats_kernel<<<grid, 128, 0, stream>>>(ds, x, mean, rstd, gamma, beta, dgamma, dbeta, M, D);
    B2T_CHECK_LAUNCH("b2t_bn_silu_bwd_stats");
    return 0;
}

extern "C" int b2t_glu_dwconv_bwd(const float* ds, const float* x, const float* g, const float* w, const float* mean,
                                  const float* rstd, const float* gamma, const float* beta, const float* dgamma,
                                  const float* dbeta, void* dg_hi, void* dg_lo, float* dw, int B, int T, int D, int K,
                                  int training, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(ds && x && g && w && mean && rstd && gamma && beta && dgamma && dbeta && dg_hi && dw && B > 0 && T > 0 && D > 0,
                  "b2t_glu_dwconv_bwd: bad arguments");
    B2T_CHECK_ARG(K > 0 && K <= DW_MAX_K && K % 2 == 1, "b2t_glu_dwconv_bwd: depthwise kernel %d must be odd and <= %d", K, DW_MAX_K);
    static bool attr_set = false;
    if (!attr_set) {
        B2T_CUDA(cudaFuncSetAttribute(glu_dwconv_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DW_BWD_SMEM));
        attr_set = true;
    }
    dim3 grid((D + DW_CH - 1) / DW_CH, B, (T + DW_TT - 1) / DW_TT);
    glu_dwconv_bwd_kernel<<<grid, 256, DW_BWD_SMEM, stream>>>(ds, x, g, w, mean, rstd, gamma, beta, dgamma, dbeta, (bf16*)dg_hi,
                                                             (bf16*)dg_lo, dw, T, D, K, (long)B * T, training);
    B2T_CHECK_LAUNCH("b2t_glu_dwconv_bwd");
    return 0;
}
