// Tensor-core fused self-attention for the single-tile case of this path (T' <= 128 keys, head
// size 16/32/64): one CTA per (trial, head), 8 warps x 16 query rows, warp-level mma.sync
// (m16n8k16, bf16 operands, fp32 accumulate) in the bf16x3 form  hi.hi + hi.lo + lo.hi  so the
// scores keep ~fp32 accuracy.  Q, K, V are converted fp32 -> bf16 hi/lo once into
// shared memory; S = QK^T, softmax, dropout and PV stay in registers (the accumulator fragment of
// S is re-packed as the A fragment of PV); only the probabilities needed by backward touch HBM.
// The per-head problem is 128 x 128 x 64: tensor work is ~1 % of the step, the kernel is bound by
// the softmax / fragment latency chain, which is why it is a warp-level kernel rather than a
// tcgen05 one (a 128-row UMMA would serialise the softmax behind a TMEM round trip per head).
// Longer sequences fall back to the SIMT kernels of attention.cu.
#include "common.cuh"
#include "../../include/b2t_sm100.h"
#include <math_constants.h>
#include <stdlib.h>

namespace attn_mma {

constexpr int TP = 128;          // padded sequence tile (queries and keys)
constexpr int THREADS = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// c += (ah + al) . (bh + bl) without the al.bl term
__device__ __forceinline__ void mma3(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], uint32_t bh0,
                                     uint32_t bh1, uint32_t bl0, uint32_t bl1) {
    mma16816(c, ah, bh0, bh1);
    mma16816(c, ah, bl0, bl1);
    mma16816(c, al, bh0, bh1);
}
// same values as split_bf16 per element; two values per conversion (F2FP.BF16.PACK_AB on the ALU - the scalar
// cvt.rn.bf16 is an F2F on the quarter-rate pipe and was the top long-scoreboard stall of the prologue in ncu)
__device__ __forceinline__ void pack_split(float x, float y, uint32_t& hi, uint32_t& lo) {
    __nv_bfloat162 h = __floats2bfloat162_rn(x, y);
    const float2 hf = __bfloat1622float2(h);
    __nv_bfloat162 l = __floats2bfloat162_rn(x - hf.x, y - hf.y);
    hi = *reinterpret_cast<uint32_t*>(&h);
    lo = *reinterpret_cast<uint32_t*>(&l);
}

// fp32 [T, D] slice (row stride ld) -> bf16 hi/lo [TP][D + 8] in shared memory, rows >= T zeroed.
// Two steps so that every global load of a tile (or of several tiles) is in flight before the first
// conversion: as one loop ptxas kept load -> split -> store order, i.e. 8 serialised L2/DRAM round trips
// per tile and 24 per kernel prologue with a single CTA per SM to hide them.
template <int D>
struct TileRegs {
    static constexpr int ITER = TP * (D / 4) / THREADS;
    float4 v[ITER];
};
template <int D>
__device__ __forceinline__ void load_tile_regs(const float* __restrict__ src, long ld, int T, bool vec, TileRegs<D>& r) {
#pragma unroll
    for (int j = 0; j < TileRegs<D>::ITER; ++j) {
        const int i = threadIdx.x + j * THREADS;
        const int t = i / (D / 4), c4 = (i - t * (D / 4)) * 4;
        float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (t < T) {
            const float* p = src + (long)t * ld + c4;
            if (vec) {
                v = *reinterpret_cast<const float4*>(p);
            } else {
                v.x = p[0], v.y = p[1], v.z = p[2], v.w = p[3];
            }
        }
        r.v[j] = v;
    }
}
template <int D>
__device__ __forceinline__ void store_split_regs(const TileRegs<D>& r, float scale, bf16* hi, bf16* lo) {
    constexpr int PITCH = D + 8;
#pragma unroll
    for (int j = 0; j < TileRegs<D>::ITER; ++j) {
        const int i = threadIdx.x + j * THREADS;
        const int t = i / (D / 4), c4 = (i - t * (D / 4)) * 4;
        uint2 h, l;
        pack_split(r.v[j].x * scale, r.v[j].y * scale, h.x, l.x);
        pack_split(r.v[j].z * scale, r.v[j].w * scale, h.y, l.y);
        *reinterpret_cast<uint2*>(hi + t * PITCH + c4) = h;
        *reinterpret_cast<uint2*>(lo + t * PITCH + c4) = l;
    }
}
template <int D>
__device__ __forceinline__ void load_split_tile(const float* __restrict__ src, long ld, int T, float scale, bf16* hi,
                                                bf16* lo, bool vec) {
    TileRegs<D> r;
    load_tile_regs<D>(src, ld, T, vec, r);
    store_split_regs<D>(r, scale, hi, lo);
}
// three tiles at once (the kernels' prologue): all loads first
template <int D>
__device__ __forceinline__ void load_split_tiles3(const float* __restrict__ s0, long ld0, bool vec0, float scale0, bf16* h0, bf16* l0,
                                                  const float* __restrict__ s1, const float* __restrict__ s2, long ld, bool vec,
                                                  bf16* h1, bf16* l1, bf16* h2, bf16* l2, int T) {
    TileRegs<D> r0, r1, r2;
    load_tile_regs<D>(s0, ld0, T, vec0, r0);
    load_tile_regs<D>(s1, ld, T, vec, r1);
    load_tile_regs<D>(s2, ld, T, vec, r2);
    store_split_regs<D>(r0, scale0, h0, l0);
    store_split_regs<D>(r1, 1.0f, h1, l1);
    store_split_regs<D>(r2, 1.0f, h2, l2);
}

// A fragment (16 rows x 16 k) of a row-major [rows][pitch] tile
__device__ __forceinline__ uint32_t a_addr(const bf16* tile, int pitch, int row0, int k0, int lane) {
    return smem_u32(tile + (row0 + (lane & 15)) * pitch + k0 + (lane >> 4) * 8);
}
// B fragments of two n-tiles from a tile stored [n][k] (B = tile^T): regs {b0,b1 | b0,b1}
__device__ __forceinline__ uint32_t bn_addr(const bf16* tile, int pitch, int n0, int k0, int lane) {
    const int mi = lane >> 3;
    return smem_u32(tile + (n0 + (mi >> 1) * 8 + (lane & 7)) * pitch + k0 + (mi & 1) * 8);
}
// B fragments of two n-tiles from a tile stored [k][n] (ldmatrix.trans): regs {b0,b1 | b0,b1}
__device__ __forceinline__ uint32_t bk_addr(const bf16* tile, int pitch, int k0, int n0, int lane) {
    const int mi = lane >> 3;
    return smem_u32(tile + (k0 + (mi & 1) * 8 + (lane & 7)) * pitch + n0 + (mi >> 1) * 8);
}
// A fragment (16 m x 16 k) of A = tile^T with the tile stored [k][m] (ldmatrix.trans)
__device__ __forceinline__ uint32_t at_addr(const bf16* tile, int pitch, int m0, int k0, int lane) {
    const int mi = lane >> 3;
    return smem_u32(tile + (k0 + (mi >> 1) * 8 + (lane & 7)) * pitch + m0 + (mi & 1) * 8);
}

template <int D>
struct Smem {
    static constexpr int PITCH = D + 8;
    static constexpr int TILE = TP * PITCH;            // elements of one bf16 tile
    static constexpr int PPITCH = TP + 8;
    static constexpr int PTILE = TP * PPITCH;
    static constexpr size_t FWD_BYTES = (size_t)6 * TILE * sizeof(bf16);
    static constexpr size_t BWD_BYTES = (size_t)(6 * TILE + 2 * PTILE) * sizeof(bf16);
};

// two CTAs per SM (110 KB of shared memory each at D = 64, 128 registers): ncu showed one CTA's 8 warps issuing on
// 28 % of the cycles, stalled on fixed-latency dependencies and instruction fetch of the fully unrolled body
template <int D>
__global__ void __launch_bounds__(THREADS, 2)
attn_fwd_mma_kernel(const float* __restrict__ qkv, long ld, int q_off, int k_off, int v_off, const int* __restrict__ kv_lens,
                    float* __restrict__ probs, float* __restrict__ ctx, bf16* __restrict__ ctx_hi, bf16* __restrict__ ctx_lo,
                    long ld_ctx, int T, int H, float scale, float drop_p, unsigned long long seed, int vec, uint32_t drop_thr) {
    B2T_PDL_TRIGGER();
    using S = Smem<D>;
    constexpr int PITCH = S::PITCH;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* Qh = reinterpret_cast<bf16*>(smem_raw);
    bf16* Ql = Qh + S::TILE;
    bf16* Kh = Ql + S::TILE;
    bf16* Kl = Kh + S::TILE;
    bf16* Vh = Kl + S::TILE;
    bf16* Vl = Vh + S::TILE;
    const int b = blockIdx.x / H, h = blockIdx.x % H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* base = qkv + (long)b * T * ld + h * D;
    load_split_tiles3<D>(base + q_off, ld, vec != 0, scale, Qh, Ql, base + k_off, base + v_off, ld, vec != 0, Kh, Kl, Vh, Vl, T);
    __syncthreads();
    const int row0 = warp * 16;
    if (row0 >= T) return;
    const int kv_len = kv_lens ? min(kv_lens[b], T) : T;
    const int g = lane >> 2, c2 = (lane & 3) * 2;

    // ---- S = (scale Q) K^T ----
    float s[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.0f;
#pragma unroll
    for (int kk = 0; kk < D / 16; ++kk) {
        uint32_t ah[4], al[4];
        ldsm_x4(ah, a_addr(Qh, PITCH, row0, kk * 16, lane));
        ldsm_x4(al, a_addr(Ql, PITCH, row0, kk * 16, lane));
#pragma unroll
        for (int np = 0; np < 8; ++np) {
            if (np * 16 < T) {   // warp-uniform
                uint32_t bh[4], bl[4];
                ldsm_x4(bh, bn_addr(Kh, PITCH, np * 16, kk * 16, lane));
                ldsm_x4(bl, bn_addr(Kl, PITCH, np * 16, kk * 16, lane));
                mma3(s[2 * np], ah, al, bh[0], bh[1], bl[0], bl[1]);
                mma3(s[2 * np + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
            }
        }
    }
    // ---- softmax over keys (rows g and g + 8 of this warp's 16) ----
    float mx0 = -CUDART_INF_F, mx1 = -CUDART_INF_F;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const bool ok = nt * 8 + c2 + e < kv_len;
            s[nt][e] = ok ? s[nt][e] : -CUDART_INF_F;
            s[nt][2 + e] = ok ? s[nt][2 + e] : -CUDART_INF_F;
            mx0 = fmaxf(mx0, s[nt][e]);
            mx1 = fmaxf(mx1, s[nt][2 + e]);
        }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    float sum0 = 0.0f, sum1 = 0.0f;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const float e0 = s[nt][e] == -CUDART_INF_F ? 0.0f : expf(s[nt][e] - mx0);
            const float e1 = s[nt][2 + e] == -CUDART_INF_F ? 0.0f : expf(s[nt][2 + e] - mx1);
            s[nt][e] = e0;
            s[nt][2 + e] = e1;
            sum0 += e0;
            sum1 += e1;
        }
    }
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 1);
    sum0 += __shfl_xor_sync(0xffffffffu, sum0, 2);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 1);
    sum1 += __shfl_xor_sync(0xffffffffu, sum1, 2);
    const float inv0 = sum0 > 0.0f ? 1.0f / sum0 : 0.0f, inv1 = sum1 > 0.0f ? 1.0f / sum1 : 0.0f;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    const int r0 = row0 + g, r1 = row0 + g + 8;
    const long prow0 = (((long)b * H + h) * T + r0) * T, prow1 = (((long)b * H + h) * T + r1) * T;
    const uint32_t thr = drop_thr, s_lo = (uint32_t)seed, s_hi = (uint32_t)(seed >> 32);
    const bool pair_ok = (T & 1) == 0;   // (row * T + even column) is 8-byte aligned: one store per column pair
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
        const int col = nt * 8 + c2;
        float p00 = s[nt][0] * inv0, p01 = s[nt][1] * inv0, p10 = s[nt][2] * inv1, p11 = s[nt][3] * inv1;
        if (col < T) {
            const bool c1 = col + 1 < T;
            if (r0 < T) {
                if (probs) {
                    if (pair_ok) {
                        *reinterpret_cast<float2*>(probs + prow0 + col) = make_float2(p00, p01);
                    } else {
                        probs[prow0 + col] = p00;
                        if (c1) probs[prow0 + col + 1] = p01;
                    }
                }
                if (drop_p > 0.0f) {
                    p00 = hash_drop(hash_fold((unsigned long long)(prow0 + col)), s_lo, s_hi, thr) ? 0.0f : p00 * drop_scale;
                    if (c1) p01 = hash_drop(hash_fold((unsigned long long)(prow0 + col + 1)), s_lo, s_hi, thr) ? 0.0f : p01 * drop_scale;
                }
            }
            if (r1 < T) {
                if (probs) {
                    if (pair_ok) {
                        *reinterpret_cast<float2*>(probs + prow1 + col) = make_float2(p10, p11);
                    } else {
                        probs[prow1 + col] = p10;
                        if (c1) probs[prow1 + col + 1] = p11;
                    }
                }
                if (drop_p > 0.0f) {
                    p10 = hash_drop(hash_fold((unsigned long long)(prow1 + col)), s_lo, s_hi, thr) ? 0.0f : p10 * drop_scale;
                    if (c1) p11 = hash_drop(hash_fold((unsigned long long)(prow1 + col + 1)), s_lo, s_hi, thr) ? 0.0f : p11 * drop_scale;
                }
            }
        }
        s[nt][0] = p00, s[nt][1] = p01, s[nt][2] = p10, s[nt][3] = p11;
    }
    // ---- O = P V ----
    float o[D / 8][4];
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.0f;
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
        if (kt * 16 < T) {   // warp-uniform
            uint32_t ah[4], al[4];
            pack_split(s[2 * kt][0], s[2 * kt][1], ah[0], al[0]);
            pack_split(s[2 * kt][2], s[2 * kt][3], ah[1], al[1]);
            pack_split(s[2 * kt + 1][0], s[2 * kt + 1][1], ah[2], al[2]);
            pack_split(s[2 * kt + 1][2], s[2 * kt + 1][3], ah[3], al[3]);
#pragma unroll
            for (int np = 0; np < D / 16; ++np) {
                uint32_t bh[4], bl[4];
                ldsm_x4_t(bh, bk_addr(Vh, PITCH, kt * 16, np * 16, lane));
                ldsm_x4_t(bl, bk_addr(Vl, PITCH, kt * 16, np * 16, lane));
                mma3(o[2 * np], ah, al, bh[0], bh[1], bl[0], bl[1]);
                mma3(o[2 * np + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
            }
        }
    }
#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) {
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            const int r = hr ? r1 : r0;
            if (r >= T) continue;
            const float x = o[nt][2 * hr], y = o[nt][2 * hr + 1];
            const long off = ((long)b * T + r) * ld_ctx + h * D + nt * 8 + c2;
            if (ctx) *reinterpret_cast<float2*>(ctx + off) = make_float2(x, y);   // ld_ctx, h * D, c2 are even
            if (ctx_hi) {
                uint32_t hh, ll;
                pack_split(x, y, hh, ll);
                *reinterpret_cast<uint32_t*>(ctx_hi + off) = hh;
                if (ctx_lo) *reinterpret_cast<uint32_t*>(ctx_lo + off) = ll;
            }
        }
    }
}

// store a 16 x D accumulator tile (rows row0 + g / g + 8, this warp) into the packed dqkv buffers
template <int D>
__device__ __forceinline__ void store_dtile(const float (&acc)[D / 8][4], float* __restrict__ dst, bf16* __restrict__ dst_hi,
                                            bf16* __restrict__ dst_lo, long ld, int row0, int T, int lane) {
    const int g = lane >> 2, c2 = (lane & 3) * 2;
#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) {
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            const int r = row0 + g + 8 * hr;
            if (r >= T) continue;
            const float x = acc[nt][2 * hr], y = acc[nt][2 * hr + 1];
            const long off = (long)r * ld + nt * 8 + c2;
            *reinterpret_cast<float2*>(dst + off) = make_float2(x, y);   // ld and the column offsets are even
            if (dst_hi) {
                uint32_t hh, ll;
                pack_split(x, y, hh, ll);
                *reinterpret_cast<uint32_t*>(dst_hi + off) = hh;
                if (dst_lo) *reinterpret_cast<uint32_t*>(dst_lo + off) = ll;
            }
        }
    }
}

// write a 16 x 128 register tile (accumulator layout) as bf16 hi/lo into a [TP][TP + 8] smem tile
__device__ __forceinline__ void store_ptile(const float (&v)[16][4], bf16* hi, bf16* lo, int pitch, int row0, int lane) {
    const int g = lane >> 2, c2 = (lane & 3) * 2;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
        uint32_t h0, l0, h1, l1;
        pack_split(v[nt][0], v[nt][1], h0, l0);
        pack_split(v[nt][2], v[nt][3], h1, l1);
        const int o0 = (row0 + g) * pitch + nt * 8 + c2, o1 = o0 + 8 * pitch;
        *reinterpret_cast<uint32_t*>(hi + o0) = h0;
        *reinterpret_cast<uint32_t*>(lo + o0) = l0;
        *reinterpret_cast<uint32_t*>(hi + o1) = h1;
        *reinterpret_cast<uint32_t*>(lo + o1) = l1;
    }
}

// C[16 keys x D] = X^T . Y  with X a [query][key] bf16 hi/lo smem tile (pitch PP) and Y a [query][D] tile
template <int D>
__device__ __forceinline__ void mma_xt_y(float (&acc)[D / 8][4], const bf16* Xh, const bf16* Xl, int pp, const bf16* Yh,
                                         const bf16* Yl, int pitch, int m0, int T, int lane) {
#pragma unroll
    for (int i = 0; i < D / 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
        if (kt * 16 < T) {   // warp-uniform; query rows >= T of X are zero
            uint32_t ah[4], al[4];
            ldsm_x4_t(ah, at_addr(Xh, pp, m0, kt * 16, lane));
            ldsm_x4_t(al, at_addr(Xl, pp, m0, kt * 16, lane));
#pragma unroll
            for (int np = 0; np < D / 16; ++np) {
                uint32_t bh[4], bl[4];
                ldsm_x4_t(bh, bk_addr(Yh, pitch, kt * 16, np * 16, lane));
                ldsm_x4_t(bl, bk_addr(Yl, pitch, kt * 16, np * 16, lane));
                mma3(acc[2 * np], ah, al, bh[0], bh[1], bl[0], bl[1]);
                mma3(acc[2 * np + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
            }
        }
    }
}

template <int D>
__global__ void __launch_bounds__(THREADS, 1)
attn_bwd_mma_kernel(const float* __restrict__ qkv, long ld, int q_off, int k_off, int v_off, const float* __restrict__ probs,
                    const float* __restrict__ dctx, long ld_ctx, float* __restrict__ dqkv, bf16* __restrict__ dqkv_hi,
                    bf16* __restrict__ dqkv_lo, int T, int H, float scale, float drop_p, unsigned long long seed, int vec,
                    int vec_ctx, uint32_t drop_thr) {
    B2T_PDL_TRIGGER();
    using S = Smem<D>;
    constexpr int PITCH = S::PITCH, PP = S::PPITCH;
    extern __shared__ __align__(16) uint8_t smem_raw[];
    bf16* Oh = reinterpret_cast<bf16*>(smem_raw);   // dO
    bf16* Ol = Oh + S::TILE;
    bf16* Vh = Ol + S::TILE;                        // V, later Q
    bf16* Vl = Vh + S::TILE;
    bf16* Kh = Vl + S::TILE;
    bf16* Kl = Kh + S::TILE;
    bf16* Ph = Kl + S::TILE;                        // dropped P, later dS   [query][key]
    bf16* Pl = Ph + S::PTILE;
    const int b = blockIdx.x / H, h = blockIdx.x % H;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* base = qkv + (long)b * T * ld + h * D;
    const long obase = (long)b * T * ld + h * D;
    load_split_tiles3<D>(dctx + (long)b * T * ld_ctx + h * D, ld_ctx, vec_ctx != 0, 1.0f, Oh, Ol, base + v_off, base + k_off, ld,
                         vec != 0, Vh, Vl, Kh, Kl, T);
    __syncthreads();
    const int row0 = warp * 16;
    const bool active = row0 < T;
    const int g = lane >> 2, c2 = (lane & 3) * 2;
    const float drop_scale = 1.0f / (1.0f - drop_p);
    float ds[16][4];
    if (active) {
        // ---- dP = dO V^T ----
#pragma unroll
        for (int i = 0; i < 16; ++i) ds[i][0] = ds[i][1] = ds[i][2] = ds[i][3] = 0.0f;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
            uint32_t ah[4], al[4];
            ldsm_x4(ah, a_addr(Oh, PITCH, row0, kk * 16, lane));
            ldsm_x4(al, a_addr(Ol, PITCH, row0, kk * 16, lane));
#pragma unroll
            for (int np = 0; np < 8; ++np) {
                if (np * 16 < T) {
                    uint32_t bh[4], bl[4];
                    ldsm_x4(bh, bn_addr(Vh, PITCH, np * 16, kk * 16, lane));
                    ldsm_x4(bl, bn_addr(Vl, PITCH, np * 16, kk * 16, lane));
                    mma3(ds[2 * np], ah, al, bh[0], bh[1], bl[0], bl[1]);
                    mma3(ds[2 * np + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
                }
            }
        }
        // ---- dropout on dP, delta = sum_j dP_ij P_ij, dS = scale P (dP - delta); dropped P -> smem ----
        const int r0 = row0 + g, r1 = r0 + 8;
        const long prow0 = (((long)b * H + h) * T + r0) * T, prow1 = (((long)b * H + h) * T + r1) * T;
        float delta0 = 0.0f, delta1 = 0.0f;
        float pd[16][4];
        // all probability loads first (independent; one 8-byte load per column pair when the rows are 8-byte aligned),
        // then the mask / delta arithmetic
        const bool pair_ok = (T & 1) == 0;
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) {
            const int col = nt * 8 + c2;
            float2 a0 = make_float2(0.0f, 0.0f), a1 = make_float2(0.0f, 0.0f);
            if (col < T) {
                if (pair_ok) {
                    if (r0 < T) a0 = *reinterpret_cast<const float2*>(probs + prow0 + col);
                    if (r1 < T) a1 = *reinterpret_cast<const float2*>(probs + prow1 + col);
                } else {
                    if (r0 < T) a0 = make_float2(probs[prow0 + col], col + 1 < T ? probs[prow0 + col + 1] : 0.0f);
                    if (r1 < T) a1 = make_float2(probs[prow1 + col], col + 1 < T ? probs[prow1 + col + 1] : 0.0f);
                }
            }
            pd[nt][0] = a0.x, pd[nt][1] = a0.y, pd[nt][2] = a1.x, pd[nt][3] = a1.y;
        }
        // keep / drop decision of every entry ONCE (bit nt * 4 + j of `keep`); it is used for dP and for the dropped P
        unsigned long long keep = ~0ull;
        if (drop_p > 0.0f) {
            const uint32_t s_lo = (uint32_t)seed, s_hi = (uint32_t)(seed >> 32);
#pragma unroll
            for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = nt * 8 + c2 + (j & 1);
                    const bool in = col < T && ((j >> 1) ? r1 : r0) < T;
                    const long off = ((j >> 1) ? prow1 : prow0) + col;
                    if (in && hash_drop(hash_fold((unsigned long long)off), s_lo, s_hi, drop_thr)) keep &= ~(1ull << (nt * 4 + j));
                }
            }
        }
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float m = (keep >> (nt * 4 + j)) & 1ull ? drop_scale : 0.0f;
                const float d = ds[nt][j] * (drop_p > 0.0f ? m : 1.0f);
                if (j < 2) delta0 += d * pd[nt][j]; else delta1 += d * pd[nt][j];
                ds[nt][j] = d;
            }
        }
        delta0 += __shfl_xor_sync(0xffffffffu, delta0, 1);
        delta0 += __shfl_xor_sync(0xffffffffu, delta0, 2);
        delta1 += __shfl_xor_sync(0xffffffffu, delta1, 1);
        delta1 += __shfl_xor_sync(0xffffffffu, delta1, 2);
#pragma unroll
        for (int nt = 0; nt < 16; ++nt) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float pv = pd[nt][j];
                ds[nt][j] = pv * (ds[nt][j] - (j < 2 ? delta0 : delta1)) * scale;
                if (drop_p > 0.0f) pd[nt][j] = (keep >> (nt * 4 + j)) & 1ull ? pv * drop_scale : 0.0f;
            }
        }
        store_ptile(pd, Ph, Pl, PP, row0, lane);
    }
    // Q (needed for dK at the end) is read into registers now - the 64 registers of the probability tile are free -
    // so that its memory round trip runs under the dQ MMAs instead of after them
    TileRegs<D> rq;
    load_tile_regs<D>(base + q_off, ld, T, vec != 0, rq);
    if (active) {
        // ---- dQ = dS K ----
        float acc[D / 8][4];
#pragma unroll
        for (int i = 0; i < D / 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
#pragma unroll
        for (int kt = 0; kt < 8; ++kt) {
            if (kt * 16 < T) {
                uint32_t ah[4], al[4];
                pack_split(ds[2 * kt][0], ds[2 * kt][1], ah[0], al[0]);
                pack_split(ds[2 * kt][2], ds[2 * kt][3], ah[1], al[1]);
                pack_split(ds[2 * kt + 1][0], ds[2 * kt + 1][1], ah[2], al[2]);
                pack_split(ds[2 * kt + 1][2], ds[2 * kt + 1][3], ah[3], al[3]);
#pragma unroll
                for (int np = 0; np < D / 16; ++np) {
                    uint32_t bh[4], bl[4];
                    ldsm_x4_t(bh, bk_addr(Kh, PITCH, kt * 16, np * 16, lane));
                    ldsm_x4_t(bl, bk_addr(Kl, PITCH, kt * 16, np * 16, lane));
                    mma3(acc[2 * np], ah, al, bh[0], bh[1], bl[0], bl[1]);
                    mma3(acc[2 * np + 1], ah, al, bh[2], bh[3], bl[2], bl[3]);
                }
            }
        }
        store_dtile<D>(acc, dqkv + obase + q_off, dqkv_hi ? dqkv_hi + obase + q_off : nullptr,
                       dqkv_lo ? dqkv_lo + obase + q_off : nullptr, ld, row0, T, lane);
    }
    __syncthreads();   // dropped P complete; V and K tiles are dead
    store_split_regs<D>(rq, 1.0f, Vh, Vl);   // Q (unscaled) into the V region
    // ---- dV = Pdrop^T dO (this warp: keys row0 .. row0 + 15) ----
    if (active) {
        float acc[D / 8][4];
        mma_xt_y<D>(acc, Ph, Pl, PP, Oh, Ol, PITCH, row0, T, lane);
        store_dtile<D>(acc, dqkv + obase + v_off, dqkv_hi ? dqkv_hi + obase + v_off : nullptr,
                       dqkv_lo ? dqkv_lo + obase + v_off : nullptr, ld, row0, T, lane);
    }
    __syncthreads();   // everyone is done reading the dropped P; Q is in place
    if (active) store_ptile(ds, Ph, Pl, PP, row0, lane);
    __syncthreads();
    // ---- dK = dS^T Q ----
    if (active) {
        float acc[D / 8][4];
        mma_xt_y<D>(acc, Ph, Pl, PP, Vh, Vl, PITCH, row0, T, lane);
        store_dtile<D>(acc, dqkv + obase + k_off, dqkv_hi ? dqkv_hi + obase + k_off : nullptr,
                       dqkv_lo ? dqkv_lo + obase + k_off : nullptr, ld, row0, T, lane);
    }
}

template <int D>
int launch_fwd(const float* qkv, long ld, int q_off, int k_off, int v_off, const int* kv_lens, float* probs, float* ctx,
               bf16* ctx_hi, bf16* ctx_lo, long ld_ctx, int B, int T, int H, float scale, float drop_p,
               unsigned long long seed, int vec, cudaStream_t stream) {
    auto kern = attn_fwd_mma_kernel<D>;
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem<D>::FWD_BYTES));
        attr = true;
    }
    kern<<<B * H, THREADS, Smem<D>::FWD_BYTES, stream>>>(qkv, ld, q_off, k_off, v_off, kv_lens, probs, ctx, ctx_hi, ctx_lo, ld_ctx,
                                                         T, H, scale, drop_p, seed, vec, drop_threshold(drop_p));
    B2T_CHECK_LAUNCH("b2t_attention_fwd (mma)");
    return 0;
}

template <int D>
int launch_bwd(const float* qkv, long ld, int q_off, int k_off, int v_off, const float* probs, const float* dctx, long ld_ctx,
               float* dqkv, bf16* dqkv_hi, bf16* dqkv_lo, int B, int T, int H, float scale, float drop_p,
               unsigned long long seed, int vec, int vec_ctx, cudaStream_t stream) {
    auto kern = attn_bwd_mma_kernel<D>;
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem<D>::BWD_BYTES));
        attr = true;
    }
    kern<<<B * H, THREADS, Smem<D>::BWD_BYTES, stream>>>(qkv, ld, q_off, k_off, v_off, probs, dctx, ld_ctx, dqkv, dqkv_hi, dqkv_lo,
                                                         T, H, scale, drop_p, seed, vec, vec_ctx, drop_threshold(drop_p));
    B2T_CHECK_LAUNCH("b2t_attention_bwd (mma)");
    return 0;
}

static bool aligned_for(const void* p, long ld, int o0, int o1, int o2, int d, int n) {
    return ((uintptr_t)p % (4 * n)) == 0 && ld % n == 0 && o0 % n == 0 && o1 % n == 0 && o2 % n == 0 && d % n == 0;
}

}  // namespace attn_mma

// returns 1 when the shape is outside the single-tile tensor-core kernel (caller falls back to SIMT)
int b2t_attn_mma_supported(int T, int d, long ld, long ld_ctx, int q_off, int k_off, int v_off) {
    if (getenv("B2T_ATTN_SIMT")) return 0;
    return T <= attn_mma::TP && (d == 16 || d == 32 || d == 64) && ld % 2 == 0 && ld_ctx % 2 == 0 && q_off % 2 == 0 &&
           k_off % 2 == 0 && v_off % 2 == 0;
}

int b2t_attn_mma_fwd(const float* qkv, long ld, int q_off, int k_off, int v_off, const int* kv_lens, float* probs, float* ctx,
                     void* ctx_hi, void* ctx_lo, long ld_ctx, int B, int T, int H, int d, float scale, float drop_p,
                     unsigned long long seed, cudaStream_t stream) {
    using namespace attn_mma;
    const int vec = aligned_for(qkv, ld, q_off, k_off, v_off, d, 4);
#define GO(DD) return launch_fwd<DD>(qkv, ld, q_off, k_off, v_off, kv_lens, probs, ctx, (bf16*)ctx_hi, (bf16*)ctx_lo, ld_ctx, B, T, H, scale, drop_p, seed, vec, stream)
    if (d == 64) GO(64);
    if (d == 32) GO(32);
    GO(16);
#undef GO
}

int b2t_attn_mma_bwd(const float* qkv, long ld, int q_off, int k_off, int v_off, const float* probs, const float* dctx,
                     long ld_ctx, float* dqkv, void* dqkv_hi, void* dqkv_lo, int B, int T, int H, int d, float scale,
                     float drop_p, unsigned long long seed, cudaStream_t stream) {
    using namespace attn_mma;
    const int vec = aligned_for(qkv, ld, q_off, k_off, v_off, d, 4);
    const int vec_ctx = aligned_for(dctx, ld_ctx, 0, 0, 0, d, 4);
#define GO(DD) return launch_bwd<DD>(qkv, ld, q_off, k_off, v_off, probs, dctx, ld_ctx, dqkv, (bf16*)dqkv_hi, (bf16*)dqkv_lo, B, T, H, scale, drop_p, seed, vec, vec_ctx, stream)
    if (d == 64) GO(64);
    if (d == 32) GO(32);
    GO(16);
#undef GO
}
