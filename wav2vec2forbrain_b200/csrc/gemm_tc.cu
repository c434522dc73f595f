// tcgen05 + TMA + TMEM GEMM for sm_100a.
//
//   D[batch][m, n] = epilogue( alpha * sum_t  A_t[m, :] . B_t[n, :] )
//
// terms: 1 (plain bf16 operands) or 3 (bf16x3: Ahi.Bhi + Ahi.Blo + Alo.Bhi, fp32 accumulate in
// TMEM -> ~fp32 accuracy at 3x the MMA cost).
//
// Structure (persistent, warp specialised, one CTA per SM):
//   warp 0      TMA producer: cp.async.bulk.tensor.3d of the A/B (hi, lo) tiles into a
//               128B-swizzled smem ring, completion on "full" mbarriers
//   warp 1      allocates TMEM; one lane issues tcgen05.mma (M=128, N=BN, K=16) per k-step and
//               term into a double-buffered fp32 accumulator in TMEM; tcgen05.commit releases
//               smem slots ("empty" barriers) and publishes finished accumulators ("tmem_full")
//   warps 2-9   epilogue: tcgen05.ld 32 lanes x 32 columns -> registers; a thread owns one output
//               row and 32 consecutive columns, so every global access is a 16-byte vector ->
//               fused bias / activation / act'(aux) / dropout / residual / bf16 hi-lo split /
//               red.add.v4 / column sums (shuffle transpose-reduce)
// Operands may be K-major (row-major [rows, K]) or MN-major (row-major [K, rows]); both are
// described as 3-D strided views so that (a) batches never share a tile row, (b) the GRU
// layer-0 / pos-conv "unfold" is an overlapping-stride view instead of an im2col copy.
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;          // 64 bf16 = 128 B = one swizzle-128B row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 320;       // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue
constexpr int NUM_EPI_WARPS = 8;
constexpr int SMEM_LIMIT = 227 * 1024;

struct KParams {
    int m_per_batch, batches, n, num_k_blocks, k_chunks_per_slice;
    int k_spans_batches, a_batched, b_batched;
    const int* b_batch_idx;
    const int* c_batch_idx;
    float* out_f32;
    float* out_pre;
    bf16* out_hi;
    bf16* out_lo;
    long long ldc, c_batch_stride, bias_batch_stride;
    const float* bias;
    const float* residual;
    const float* aux;
    float* colsum;
    int act, aux_mode, atomic, split_k;
    float alpha;
    float drop_p, drop_scale, out_scale;
    int vec4;
    unsigned long long drop_seed;
    int m_tiles, n_tiles, total_items;
};

// ------------------------------------------------------------------------------------------
// PTX wrappers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// bounded spin: a wrong descriptor / byte count must trap, not hang the box
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (spin > (1u << 22)) {
            printf("b2t_gemm: mbarrier timeout block %d thread %d bar %u parity %u\n", blockIdx.x, threadIdx.x, bar,
                   parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// UMMA shared-memory descriptor (sm_100): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46),
// version=1 [46,48), layout SWIZZLE_128B=2 [61,64).
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// instruction descriptor, kind::f16: D=f32 (bits 4-5 = 1), A=B=bf16 (bits 7-9, 10-12 = 1),
// a_major bit 15, b_major bit 16 (1 = MN-major), N>>3 bits 17-22, M>>4 bits 24-28
__host__ __device__ constexpr uint32_t make_idesc(int m, int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

template <int NTERMS, int BN>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = (NTERMS == 3 ? 2 : 1) * (A_BYTES + B_BYTES);
    static constexpr int EPI_BYTES = 0;      // the epilogue works from registers
    static constexpr int AUX_BYTES = 1024;   // barriers + tmem pointer
    static constexpr int STAGES_RAW = (SMEM_LIMIT - 1024 - EPI_BYTES - AUX_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + EPI_BYTES + AUX_BYTES;
    static constexpr int TMEM_COLS = 2 * BN >= 512 ? 512 : (2 * BN >= 256 ? 256 : (2 * BN >= 128 ? 128 : 64));
    static_assert(STAGES >= 2, "need a pipeline");
};

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == 1) return gelu_erf(v);
    if (act == 2) return silu_f(v);
    if (act == 3) return v / (1.0f + fabsf(v));
    return v;
}
__device__ __forceinline__ float apply_dact(float a, int mode) {
    if (mode == 1) return gelu_erf_grad(a);
    if (mode == 2) return silu_grad(a);
    if (mode == 3) {
        const float d = 1.0f + fabsf(a);
        return 1.0f / (d * d);
    }
    return 1.0f;
}

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Column sums of an 8-column register slab over the 32 rows (lanes) of a warp: recursive halving,
// 9 shuffles.  Returns the total of column `col_of_lane` (valid in lanes with (lane & 3) == 0).
__device__ __forceinline__ float warp_colsum8(const float (&v)[8], int lane, int& col_of_lane) {
    float a[4], b[2];
    const bool u16 = (lane & 16) != 0, u8 = (lane & 8) != 0, u4 = (lane & 4) != 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float keep = u16 ? v[i + 4] : v[i], send = u16 ? v[i] : v[i + 4];
        a[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float keep = u8 ? a[i + 2] : a[i], send = u8 ? a[i] : a[i + 2];
        b[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    float c = (u4 ? b[1] : b[0]) + __shfl_xor_sync(0xffffffffu, u4 ? b[0] : b[1], 4);
    c += __shfl_xor_sync(0xffffffffu, c, 1);
    c += __shfl_xor_sync(0xffffffffu, c, 2);
    col_of_lane = (u16 ? 4 : 0) + (u8 ? 2 : 0) + (u4 ? 1 : 0);
    return c;
}

// Fused epilogue on 8 consecutive columns of one output row held in registers.  Every stage is its
// own short loop with the mode branch hoisted out of it: the executed instruction footprint stays
// small (a fully unrolled 32-column body with all modes inline thrashed the instruction cache).
// off0 = element offset of (row, col0); nvalid = columns inside N (8, or 4 at the edge on the VEC
// path; anything in 1..8 on the scalar path).
template <bool VEC>
__device__ __forceinline__ void epi_cols8(const KParams& p, float (&v)[8], long long off0, int col0, int nvalid, bool row_ok,
                                          const float* bias_ptr, int lane) {
    if (row_ok) {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = e < nvalid ? v[e] * p.alpha : 0.0f;
        if (bias_ptr) {
#pragma unroll
            for (int e = 0; e < 8; ++e)
                if (e < nvalid) v[e] += bias_ptr[col0 + e];
        }
        if (p.out_pre) {
            if (VEC) {
                *reinterpret_cast<float4*>(p.out_pre + off0) = make_float4(v[0], v[1], v[2], v[3]);
                if (nvalid > 4) *reinterpret_cast<float4*>(p.out_pre + off0 + 4) = make_float4(v[4], v[5], v[6], v[7]);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (e < nvalid) p.out_pre[off0 + e] = v[e];
            }
        }
        if (p.act == 1) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = gelu_erf(v[e]);
        } else if (p.act == 2) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = silu_f(v[e]);
        } else if (p.act == 3) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = v[e] / (1.0f + fabsf(v[e]));
        }
        if (p.aux_mode) {
            float a[8];
            if (VEC) {
                const float4 t0 = *reinterpret_cast<const float4*>(p.aux + off0);
                const float4 t1 = nvalid > 4 ? *reinterpret_cast<const float4*>(p.aux + off0 + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                a[0] = t0.x, a[1] = t0.y, a[2] = t0.z, a[3] = t0.w, a[4] = t1.x, a[5] = t1.y, a[6] = t1.z, a[7] = t1.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) a[e] = e < nvalid ? p.aux[off0 + e] : 0.0f;
            }
            if (p.aux_mode == 1) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] *= gelu_erf_grad(a[e]);
            } else if (p.aux_mode == 2) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] *= silu_grad(a[e]);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float d = 1.0f + fabsf(a[e]);
                    v[e] *= 1.0f / (d * d);
                }
            }
        }
        if (p.drop_p > 0.0f) {
#pragma unroll
            for (int e = 0; e < 8; ++e)
                v[e] = hash_uniform(p.drop_seed, (unsigned long long)(off0 + e)) < p.drop_p ? 0.0f : v[e] * p.drop_scale;
        }
        if (p.out_scale != 1.0f) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] *= p.out_scale;
        }
        if (p.residual) {
            if (VEC) {
                const float4 t0 = *reinterpret_cast<const float4*>(p.residual + off0);
                const float4 t1 = nvalid > 4 ? *reinterpret_cast<const float4*>(p.residual + off0 + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                v[0] += t0.x, v[1] += t0.y, v[2] += t0.z, v[3] += t0.w, v[4] += t1.x, v[5] += t1.y, v[6] += t1.z, v[7] += t1.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (e < nvalid) v[e] += p.residual[off0 + e];
            }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = e < nvalid ? v[e] : 0.0f;
        if (p.out_f32) {
            float* o = p.out_f32 + off0;
            if (VEC) {
                if (p.atomic) {
                    red_add_v4(o, v[0], v[1], v[2], v[3]);
                    if (nvalid > 4) red_add_v4(o + 4, v[4], v[5], v[6], v[7]);
                } else {
                    *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
                    if (nvalid > 4) *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    if (e < nvalid) {
                        if (p.atomic)
                            atomicAdd(o + e, v[e]);
                        else
                            o[e] = v[e];
                    }
                }
            }
        }
        if (p.out_hi) {
            bf16 h[8], l[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) split_bf16(v[e], h[e], l[e]);
            if (VEC) {
                auto pk = [](bf16 a, bf16 b) {
                    __nv_bfloat162 t = __halves2bfloat162(a, b);
                    return *reinterpret_cast<uint32_t*>(&t);
                };
                const uint2 h0 = make_uint2(pk(h[0], h[1]), pk(h[2], h[3])), h1 = make_uint2(pk(h[4], h[5]), pk(h[6], h[7]));
                *reinterpret_cast<uint2*>(p.out_hi + off0) = h0;
                if (nvalid > 4) *reinterpret_cast<uint2*>(p.out_hi + off0 + 4) = h1;
                if (p.out_lo) {
                    const uint2 l0 = make_uint2(pk(l[0], l[1]), pk(l[2], l[3])), l1 = make_uint2(pk(l[4], l[5]), pk(l[6], l[7]));
                    *reinterpret_cast<uint2*>(p.out_lo + off0) = l0;
                    if (nvalid > 4) *reinterpret_cast<uint2*>(p.out_lo + off0 + 4) = l1;
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    if (e < nvalid) {
                        p.out_hi[off0 + e] = h[e];
                        if (p.out_lo) p.out_lo[off0 + e] = l[e];
                    }
                }
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = 0.0f;   // outside the matrix: contributes nothing to column sums
    }
    if (p.colsum) {   // warp-uniform
        int cl;
        const float s = warp_colsum8(v, lane, cl);
        if ((lane & 3) == 0 && cl < nvalid) atomicAdd(p.colsum + col0 + cl, s);
    }
}

template <int NTERMS, bool A_MN, bool B_MN, int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
               const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl, const KParams p) {
    using C = Cfg<NTERMS, BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + C::STAGES * C::STAGE_BYTES + C::EPI_BYTES;
    // barrier layout: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then tmem ptr
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
    volatile uint32_t* tmem_ptr_gen =
        reinterpret_cast<volatile uint32_t*>(smem_gen + C::STAGES * C::STAGE_BYTES + C::EPI_BYTES + 8 * (2 * C::STAGES + 4));
    const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * C::STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAh) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmBh) : "memory");
        if (NTERMS == 3) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAl) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmBl) : "memory");
        }
        for (int s = 0; s < C::STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull_bar(s), 1);
            mbar_init(tempty_bar(s), NUM_EPI_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_ptr_addr),
                     "r"((uint32_t)C::TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_gen;

    auto decode_item = [&](int item, int& split, int& batch, int& m_blk, int& n_blk) {
        n_blk = item % p.n_tiles;
        int t = item / p.n_tiles;
        m_blk = t % p.m_tiles;
        t /= p.m_tiles;
        batch = t % p.batches;
        split = t / p.batches;
    };
    auto k_range = [&](int split, int& kb0, int& kb1) {
        kb0 = (int)(((long long)split * p.num_k_blocks) / p.split_k);
        kb1 = (int)(((long long)(split + 1) * p.num_k_blocks) / p.split_k);
    };

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x) {
                int split, batch, m_blk, n_blk, kb0, kb1;
                decode_item(item, split, batch, m_blk, n_blk);
                k_range(split, kb0, kb1);
                const int bbatch = p.b_batch_idx ? p.b_batch_idx[batch] : (p.b_batched ? batch : 0);
                const int abatch = p.a_batched ? batch : 0;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    const uint32_t fb = full_bar(stage);
                    mbar_expect_tx(fb, (uint32_t)C::STAGE_BYTES);
                    const uint32_t sA = smem_base + stage * C::STAGE_BYTES;
                    const uint32_t sB = sA + (NTERMS == 3 ? 2 : 1) * C::A_BYTES;
                    int krow = kb * BK, kslice = 0;
                    if (p.k_spans_batches) {
                        kslice = kb / p.k_chunks_per_slice;
                        krow = (kb - kslice * p.k_chunks_per_slice) * BK;
                    }
#pragma unroll
                    for (int t = 0; t < (NTERMS == 3 ? 2 : 1); ++t) {
                        const CUtensorMap* ma = t == 0 ? &tmAh : &tmAl;
                        const CUtensorMap* mb = t == 0 ? &tmBh : &tmBl;
                        if (!A_MN) {
                            tma_load_3d(sA + t * C::A_BYTES, ma, fb, kb * BK, m_blk * BM, abatch);
                        } else {
#pragma unroll
                            for (int i = 0; i < BM / 64; ++i)
                                tma_load_3d(sA + t * C::A_BYTES + i * (BK * 128), ma, fb, m_blk * BM + i * 64, krow,
                                            p.k_spans_batches ? kslice : abatch);
                        }
                        if (!B_MN) {
                            tma_load_3d(sB + t * C::B_BYTES, mb, fb, kb * BK, n_blk * BN, bbatch);
                        } else {
#pragma unroll
                            for (int i = 0; i < BN / 64; ++i)
                                tma_load_3d(sB + t * C::B_BYTES + i * (BK * 128), mb, fb, n_blk * BN + i * 64, krow,
                                            p.k_spans_batches ? kslice : bbatch);
                        }
                    }
                    if (++stage == C::STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BM, BN, A_MN, B_MN);
            // K-major SW128: 8-row groups 1024 B apart (SBO), k-step = +32 B inside the swizzle row
            // MN-major SW128: 64-element MN atoms BK*128 B apart (LBO), 8 k-rows = 1024 B (SBO),
            //                 k-step of 16 rows = +2048 B
            constexpr uint32_t A_LBO = A_MN ? BK * 128 : 0, B_LBO = B_MN ? BK * 128 : 0;
            constexpr uint32_t A_KSTEP = A_MN ? UMMA_K * 128 : UMMA_K * 2, B_KSTEP = B_MN ? UMMA_K * 128 : UMMA_K * 2;
            int stage = 0;
            uint32_t phase = 0;
            int local_iter = 0;
            for (int item = blockIdx.x; item < p.total_items; item += gridDim.x, ++local_iter) {
                int split, batch, m_blk, n_blk, kb0, kb1;
                decode_item(item, split, batch, m_blk, n_blk);
                k_range(split, kb0, kb1);
                const int acc = local_iter & 1;
                const uint32_t acc_phase = (uint32_t)(local_iter >> 1) & 1u;
                mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
                uint32_t accumulate = 0;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(full_bar(stage), phase);
                    tc_fence_after();
                    const uint32_t sA = smem_base + stage * C::STAGE_BYTES;
                    const uint32_t sB = sA + (NTERMS == 3 ? 2 : 1) * C::A_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        const uint64_t dAh = make_desc(sA + k * A_KSTEP, A_LBO, 1024);
                        const uint64_t dBh = make_desc(sB + k * B_KSTEP, B_LBO, 1024);
                        tc_mma_bf16(tmem_d, dAh, dBh, idesc, accumulate);
                        accumulate = 1;
                        if (NTERMS == 3) {
                            const uint64_t dAl = make_desc(sA + C::A_BYTES + k * A_KSTEP, A_LBO, 1024);
                            const uint64_t dBl = make_desc(sB + C::B_BYTES + k * B_KSTEP, B_LBO, 1024);
                            tc_mma_bf16(tmem_d, dAh, dBl, idesc, 1);
                            tc_mma_bf16(tmem_d, dAl, dBh, idesc, 1);
                        }
                    }
                    tc_commit(empty_bar(stage));
                    if (kb == kb1 - 1) tc_commit(tfull_bar(acc));
                    if (++stage == C::STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else {
        // ================================ epilogue ====================================
        // 8 warps: TMEM lane quadrant q = warp & 3 (hardware restriction), the two warps of a
        // quadrant take alternate 32-column chunks.  A thread owns ONE output row (its TMEM lane)
        // and 32 consecutive columns of it: all global traffic is 16-byte vectors on that row.
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        int local_iter = 0;
        for (int item = blockIdx.x; item < p.total_items; item += gridDim.x, ++local_iter) {
            int split, batch, m_blk, n_blk;
            decode_item(item, split, batch, m_blk, n_blk);
            const int acc = local_iter & 1;
            const uint32_t acc_phase = (uint32_t)(local_iter >> 1) & 1u;
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            const int cbatch = p.c_batch_idx ? p.c_batch_idx[batch] : batch;
            const int row = m_blk * BM + q * 32 + lane;
            const bool row_ok = row < p.m_per_batch;
            const long long row_off = (long long)cbatch * p.c_batch_stride + (long long)row * p.ldc;
            const int bslot = p.b_batch_idx ? p.b_batch_idx[batch] : (p.b_batched ? batch : 0);
            const float* bias_ptr =
                (p.bias != nullptr && split == 0) ? p.bias + (long long)bslot * p.bias_batch_stride : nullptr;
#pragma unroll 1
            for (int c = half; c < BN / 32; c += 2) {
#pragma unroll 1
                for (int s8 = 0; s8 < 4; ++s8) {
                    float v[8];
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + c * 32 + s8 * 8);
                    tmem_ld_32x8(taddr, v);
                    if (c + 2 >= BN / 32 && s8 == 3) {
                        // this warp's share of the accumulator is in registers: hand the TMEM stage back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty_bar(acc));
                    }
                    const int col0 = n_blk * BN + c * 32 + s8 * 8;
                    if (col0 >= p.n) continue;   // warp-uniform
                    const int nvalid = p.n - col0 < 8 ? p.n - col0 : 8;
                    if (p.vec4)
                        epi_cols8<true>(p, v, row_off + col0, col0, nvalid, row_ok, bias_ptr, lane);
                    else
                        epi_cols8<false>(p, v, row_off + col0, col0, nvalid, row_ok, bias_ptr, lane);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                     : "memory");
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

// 3-D bf16 map, d0 contiguous, box {64, box1, 1}, 128B swizzle, zero OOB fill
int make_map(CUtensorMap* map, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t s1, int64_t s2, int box1,
             const char* what) {
    EncodeTiledFn fn = get_encode_fn();
    B2T_CHECK_ARG(fn != nullptr, "b2t_gemm: cuTensorMapEncodeTiled entry point not found");
    B2T_CHECK_ARG(((uintptr_t)ptr & 15) == 0, "b2t_gemm: operand %s not 16-byte aligned", what);
    B2T_CHECK_ARG(s1 % 8 == 0 && s2 % 8 == 0, "b2t_gemm: operand %s strides (%lld, %lld) must be multiples of 8 elements",
                  what, (long long)s1, (long long)s2);
    cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
    cuuint64_t strides[2] = {(cuuint64_t)s1 * 2, (cuuint64_t)s2 * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)box1, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B2T_CHECK_ARG(r == CUDA_SUCCESS,
                  "b2t_gemm: cuTensorMapEncodeTiled(%s) failed with %d (dims %lld %lld %lld strides %lld %lld box1 %d)", what,
                  (int)r, (long long)d0, (long long)d1, (long long)d2, (long long)s1, (long long)s2, box1);
    return 0;
}

int num_sms() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    }
    return n;
}

template <int NTERMS, bool A_MN, bool B_MN, int BN>
int launch(const b2t_gemm_args* a, const KParams& kp, cudaStream_t stream) {
    using C = Cfg<NTERMS, BN>;
    CUtensorMap mAh, mAl, mBh, mBl;
    const int a_box1 = A_MN ? BK : BM;
    const int b_box1 = B_MN ? BK : BN;
    int rc;
    if ((rc = make_map(&mAh, a->a_hi, a->a_d0, a->a_d1, a->a_d2, a->a_s1, a->a_s2, a_box1, "a_hi"))) return rc;
    if ((rc = make_map(&mBh, a->b_hi, a->b_d0, a->b_d1, a->b_d2, a->b_s1, a->b_s2, b_box1, "b_hi"))) return rc;
    if (NTERMS == 3) {
        if ((rc = make_map(&mAl, a->a_lo, a->a_d0, a->a_d1, a->a_d2, a->a_s1, a->a_s2, a_box1, "a_lo"))) return rc;
        if ((rc = make_map(&mBl, a->b_lo, a->b_d0, a->b_d1, a->b_d2, a->b_s1, a->b_s2, b_box1, "b_lo"))) return rc;
    } else {
        mAl = mAh;
        mBl = mBh;
    }
    auto kern = gemm_tc_kernel<NTERMS, A_MN, B_MN, BN>;
    static bool attr_set = false;
    if (!attr_set) {
        B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        attr_set = true;
    }
    int grid = kp.total_items < num_sms() ? kp.total_items : num_sms();
    kern<<<grid, NUM_THREADS, C::SMEM_BYTES, stream>>>(mAh, mAl, mBh, mBl, kp);
    B2T_CHECK_LAUNCH("b2t_gemm");
    return 0;
}

template <int NTERMS, int BN>
int dispatch_major(const b2t_gemm_args* a, const KParams& kp, cudaStream_t s) {
    if (a->a_mn_major) {
        if (a->b_mn_major) return launch<NTERMS, true, true, BN>(a, kp, s);
        return launch<NTERMS, true, false, BN>(a, kp, s);
    }
    if (a->b_mn_major) return launch<NTERMS, false, true, BN>(a, kp, s);
    return launch<NTERMS, false, false, BN>(a, kp, s);
}

}  // namespace

extern "C" int b2t_gemm(const b2t_gemm_args* a, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(a != nullptr, "b2t_gemm: null args");
    B2T_CHECK_ARG(a->a_hi && a->b_hi, "b2t_gemm: a_hi / b_hi are required");
    B2T_CHECK_ARG((a->a_lo == nullptr) == (a->b_lo == nullptr), "b2t_gemm: a_lo and b_lo must both be set or both NULL");
    B2T_CHECK_ARG(a->m_per_batch > 0 && a->batches > 0 && a->n > 0 && a->k > 0, "b2t_gemm: bad sizes m=%d batches=%d n=%d k=%d",
                  a->m_per_batch, a->batches, a->n, a->k);
    B2T_CHECK_ARG(a->split_k >= 1, "b2t_gemm: split_k must be >= 1");
    B2T_CHECK_ARG(a->split_k == 1 || (a->atomic && a->act == 0 && !a->out_hi && !a->out_pre && !a->aux_mode && !a->residual && a->drop_p == 0.0f &&
                                      !a->colsum && a->out_f32),
                  "b2t_gemm: split_k > 1 needs atomic fp32 output and a linear epilogue");
    B2T_CHECK_ARG(!a->k_spans_batches || (a->a_mn_major && a->b_mn_major),
                  "b2t_gemm: k_spans_batches needs both operands MN-major");
    B2T_CHECK_ARG(!a->aux_mode || a->aux, "b2t_gemm: aux_mode without aux");
    B2T_CHECK_ARG(a->out_f32 || a->out_hi || a->out_pre || a->colsum, "b2t_gemm: no output");

    KParams kp;
    memset(&kp, 0, sizeof(kp));
    kp.m_per_batch = a->m_per_batch;
    kp.batches = a->batches;
    kp.n = a->n;
    kp.k_spans_batches = a->k_spans_batches;
    if (a->k_spans_batches) {
        kp.k_chunks_per_slice = ceil_div_i(a->k, BK);
        kp.num_k_blocks = kp.k_chunks_per_slice * (int)a->a_d2;
        B2T_CHECK_ARG(a->a_d2 == a->b_d2 && a->a_d1 == a->b_d1, "b2t_gemm: k_spans_batches operand extents differ");
    } else {
        kp.k_chunks_per_slice = 1;
        kp.num_k_blocks = ceil_div_i(a->k, BK);
    }
    B2T_CHECK_ARG(a->split_k <= kp.num_k_blocks, "b2t_gemm: split_k %d > k blocks %d", a->split_k, kp.num_k_blocks);
    kp.a_batched = a->a_batched;
    kp.b_batched = a->b_batched;
    kp.b_batch_idx = a->b_batch_idx;
    kp.c_batch_idx = a->c_batch_idx;
    kp.out_f32 = a->out_f32;
    kp.out_pre = a->out_pre;
    kp.out_hi = (bf16*)a->out_hi;
    kp.out_lo = (bf16*)a->out_lo;
    kp.ldc = a->ldc;
    kp.c_batch_stride = a->c_batch_stride;
    kp.bias_batch_stride = a->bias_batch_stride;
    kp.bias = a->bias;
    kp.residual = a->residual;
    kp.aux = a->aux;
    kp.colsum = a->colsum;
    kp.act = a->act;
    kp.aux_mode = a->aux_mode;
    kp.atomic = a->atomic;
    kp.split_k = a->split_k;
    kp.alpha = a->alpha;
    B2T_CHECK_ARG(a->drop_p >= 0.0f && a->drop_p < 1.0f, "b2t_gemm: drop_p must be in [0, 1)");
    kp.drop_p = a->drop_p;
    kp.drop_scale = 1.0f / (1.0f - a->drop_p);
    kp.drop_seed = a->drop_seed;
    kp.out_scale = a->out_scale == 0.0f ? 1.0f : a->out_scale;
    {
        auto al16 = [](const void* q) { return q == nullptr || ((uintptr_t)q & 15) == 0; };
        auto al8 = [](const void* q) { return q == nullptr || ((uintptr_t)q & 7) == 0; };
        kp.vec4 = (a->n % 4 == 0) && (a->ldc % 4 == 0) && (a->c_batch_stride % 4 == 0) && al16(a->out_f32) &&
                  al16(a->out_pre) && al16(a->residual) && al16(a->aux) && al8(a->out_hi) && al8(a->out_lo);
    }

    int bn = a->block_n;
    if (bn == 0) bn = a->n <= 64 ? 64 : 128;
    B2T_CHECK_ARG(bn == 64 || bn == 128 || bn == 256, "b2t_gemm: block_n must be 64, 128 or 256");
    kp.m_tiles = ceil_div_i(a->m_per_batch, BM);
    kp.n_tiles = ceil_div_i(a->n, bn);
    kp.total_items = kp.m_tiles * kp.n_tiles * kp.batches * a->split_k;

    const bool three = a->a_lo != nullptr;
    if (three) {
        if (bn == 64) return dispatch_major<3, 64>(a, kp, stream);
        if (bn == 128) return dispatch_major<3, 128>(a, kp, stream);
        return dispatch_major<3, 256>(a, kp, stream);
    }
    if (bn == 64) return dispatch_major<1, 64>(a, kp, stream);
    if (bn == 128) return dispatch_major<1, 128>(a, kp, stream);
    return dispatch_major<1, 256>(a, kp, stream);
}
