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
// warp group 0 = warp 0 TMA, warp 1 MMA, warps 2-3 idle; warp groups 1-2 = the 8 epilogue warps.  Roles are cut on warp-group
// boundaries so that setmaxnreg can move registers from the control group (REGS_CTRL each) to the epilogue (REGS_EPI
// each): with 10 warps the fullest scheduler held 3 and the whole kernel was capped at 168 registers, which the
// row-per-thread epilogue (32 accumulators + a prefetched 32-register input chunk + 32 packed outputs) spilled.
constexpr int NUM_THREADS = 384;
constexpr int EPI_WARP0 = 4;
constexpr int REGS_CTRL = 56, REGS_EPI = 224;   // 128 * 56 + 256 * 224 = 64512 <= 65536
constexpr int NUM_EPI_WARPS = 8;
constexpr int SMEM_LIMIT = 227 * 1024;
constexpr int SK_FLAG_BYTES = 1024;    // stream-K workspace: 256 int flags, then gridDim.x fp32 tile slots

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
    int pre_mode;   // 1: out_pre = act'(x) * dropout factor * out_scale (the backward factor) instead of x
    float alpha;
    float drop_p, drop_scale, out_scale;
    uint32_t drop_thr;   // drop_threshold(drop_p): keep / drop as an integer compare (computed on the host, it needs fp64)
    int vec4;
    int debug;   // B2T_GEMM_DEBUG (profiling aid): 1 skip the epilogue body, 2 skip MMA issue, 4 skip TMA loads
    uint32_t flags;
    unsigned long long drop_seed;
    int m_tiles, n_tiles, total_items;
    // stream-K: the tiles' k-blocks form one line of `sk_units` units cut into gridDim.x equal ranges
    int dynamic;       // work items are handed out by cluster launch control (one cluster launched per item, see Sched)
    int stream_k;
    int sk_tile0;      // tiles [0, sk_tile0) run as whole tiles, tiles [sk_tile0, tiles) form the stream-K line
    long long sk_units;
    int sk_G;          // units (CTAs / pairs) the line is cut over = min(grid units, sk_units): every one holds >= 1 k-block
    float* sk_ws;      // [gridDim.x][BM * BN] fp32 partial accumulators
    int* sk_flags;     // [gridDim.x] arrivals of partials at the tile's owner CTA (self-cleaning)
};

// One unit of work of a CTA.  mode 0: whole tile (or a legacy split-K slice, reduced with red.add);
// stream-K: mode 1 = contributor (k-range does not start the tile: raw partial -> workspace, signal the
// owner), mode 2 = owner (k-range starts the tile but does not finish it: waits for the n_contrib
// following CTAs' partials, adds them, runs the fused epilogue).
struct Item {
    int split, batch, m_blk, n_blk, kb0, kb1, mode, n_contrib, owner;
};

// `idx` / `G`: this CTA's position among, and the number of, scheduling units of the grid - CTAs for cta_group::1,
// CTA pairs for cta_group::2 (both CTAs of a pair walk the same item list; the pair's tile is 2 x BM rows, the
// CTA of cluster rank r owns row block m_blk = 2 * (pair row block) + r, kept in it.m_blk).
//
// Dynamic schedule (p.dynamic; opt-in, b2t_gemm_set_dynamic_schedule / B2T_GEMM_CLC=1): the grid has ONE cluster per work item and the hardware's cluster launch
// control hands the items out.  A resident cluster processes its own item, and while it does, asks to CANCEL the launch
// of a not-yet-launched cluster (clusterlaunchcontrol.try_cancel): if that succeeds it owns that cluster's item next,
// if it fails nothing is left and the cluster finishes.  Clusters that cannot become resident - SMs held by an NCCL
// kernel, by the GEMM of the other stream, by the previous kernel's tail under programmatic dependent launch - simply
// have their items taken by the resident ones; with a fixed persistent grid and statically strided item lists those CTAs
// started late and then ran their WHOLE list (up to twice the kernel's time, measured as exposed all-reduce time).
// Responses arrive through a ring of two 16-byte slots per CTA (multicast to both CTAs of a pair): a query is issued as
// soon as every consumer role of the cluster (TMA producers, MMA issuer, epilogue warps) has read the slot's previous
// response, i.e. a cluster is committed to at most two items beyond the one its epilogue works on.
constexpr int CLC_OFF = 256;        // byte offset of {clc_full[2], clc_empty[2], pad, resp[2]} inside the barrier block
constexpr int CLC_RESP_OFF = 320;
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity);
__device__ __forceinline__ void mbar_arrive(uint32_t bar);
__device__ __forceinline__ void mbar_arrive_cluster_release(uint32_t cluster_bar);
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank);
// (valid, first CTA id of the cancelled cluster) of a response slot
__device__ __forceinline__ bool clc_read(uint32_t resp_addr, int& ctaid_x) {
    uint32_t x = 0, y, z, valid;
    asm volatile(
        "{\n\t.reg .pred p1;\n\t.reg .b128 r;\n\t"
        "ld.shared.b128 r, [%4];\n\t"
        "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p1, r;\n\t"
        "selp.u32 %3, 1, 0, p1;\n\t"
        "@p1 clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%0, %1, %2, _}, r;\n\t}"
        : "+r"(x), "=r"(y), "=r"(z), "=r"(valid)
        : "r"(resp_addr)
        : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the slot is rewritten by the async proxy
    ctaid_x = (int)x;
    return valid != 0;
}
__device__ __forceinline__ void clc_issue(uint32_t resp_addr, uint32_t bar) {
    asm volatile(
        "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.multicast::cluster::all.b128 [%0], [%1];"
        ::"r"(resp_addr), "r"(bar)
        : "memory");
}

struct Sched {
    const KParams& p;
    const int idx, G, cg, rank;
    int item;
    long long u, u_end;
    // dynamic schedule: this consumer's position in the response ring
    uint32_t clc_base, clc_empty_leader, use;
    bool first, warp_consumer;
    // bar_base: shared address of the CTA's barrier block; warp_consumer: next() is executed by all 32 lanes of a warp
    __device__ __forceinline__ Sched(const KParams& p_, int cg_, int rank_, uint32_t bar_base, bool warp_consumer_)
        : p(p_), idx((int)blockIdx.x / cg_), G((int)gridDim.x / cg_), cg(cg_), rank(rank_), item((int)blockIdx.x / cg_), u(0), u_end(0),
          clc_base(bar_base + CLC_OFF), use(0), first(true), warp_consumer(warp_consumer_) {
        clc_empty_leader = cg_ == 2 ? mapa_u32(bar_base + CLC_OFF + 16, 0) : bar_base + CLC_OFF + 16;
        if (p.stream_k && idx < p.sk_G) {
            u = (long long)idx * p.sk_units / p.sk_G;
            u_end = (long long)(idx + 1) * p.sk_units / p.sk_G;
        }
    }
    __device__ __forceinline__ void decode_tile(int t, Item& it) const {
        it.n_blk = t % p.n_tiles;
        t /= p.n_tiles;
        it.m_blk = (t % p.m_tiles) * cg + rank;   // p.m_tiles counts pair row blocks when cg == 2
        it.batch = t / p.m_tiles;
    }
    __device__ __forceinline__ bool next(Item& it) {
        if (p.dynamic) {
            if (first) {
                first = false;
            } else {
                const uint32_t slot = use & 1u;
                mbar_wait(clc_base + 8u * slot, (use >> 1) & 1u);
                int ctaid;
                const bool valid = clc_read(clc_base + (CLC_RESP_OFF - CLC_OFF) + 16u * slot, ctaid);
                if (warp_consumer) __syncwarp();
                if (!warp_consumer || (threadIdx.x & 31) == 0) {   // this role has read the slot
                    if (cg == 2)
                        mbar_arrive_cluster_release(clc_empty_leader + 8u * slot);
                    else
                        mbar_arrive(clc_empty_leader + 8u * slot);
                }
                ++use;
                if (!valid) return false;
                item = ctaid / cg;
            }
        }
        if (!p.stream_k) {
            if (item >= p.total_items) return false;
            const int per_split = p.n_tiles * p.m_tiles * p.batches;
            it.split = item / per_split;
            decode_tile(item - it.split * per_split, it);
            it.kb0 = (int)(((long long)it.split * p.num_k_blocks) / p.split_k);
            it.kb1 = (int)(((long long)(it.split + 1) * p.num_k_blocks) / p.split_k);
            it.mode = 0, it.n_contrib = 0, it.owner = 0;
            if (!p.dynamic) item += G;
            return true;
        }
        // Hybrid schedule: the last (tiles mod G) tiles would be a mostly empty wave, so THEIR k-blocks are cut into G
        // equal ranges (stream-K) and done first - the owners' fix-up epilogues then run under the MMAs of the whole
        // tiles that follow - and the remaining multiple of G tiles run as whole tiles, one per unit and wave.
        it.split = 0;
        if (u < u_end) {
            const long long KB = p.num_k_blocks;
            const long long t = u / KB, t0 = t * KB;
            const long long e = u_end < t0 + KB ? u_end : t0 + KB;
            decode_tile(p.sk_tile0 + (int)t, it);
            it.kb0 = (int)(u - t0);
            it.kb1 = (int)(e - t0);
            it.n_contrib = 0, it.owner = 0;
            if (it.kb0 == 0 && it.kb1 == KB) {
                it.mode = 0;
            } else if (it.kb0 == 0) {
                it.mode = 2;   // the unit holding k-block unit x is ((x + 1) * G - 1) / units
                it.n_contrib = (int)(((t0 + KB) * p.sk_G - 1) / p.sk_units) - idx;
            } else {
                it.mode = 1;
                it.owner = (int)(((t0 + 1) * p.sk_G - 1) / p.sk_units);
            }
            u = e;
            return true;
        }
        if (item >= p.sk_tile0) return false;
        decode_tile(item, it);
        it.kb0 = 0, it.kb1 = p.num_k_blocks;
        it.mode = 0, it.n_contrib = 0, it.owner = 0;
        item += G;
        return true;
    }
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
// ---- cta_group::2 (CTA pair) forms ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on a barrier of another CTA of the cluster (address from mapa)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
    // relaxed: the TMEM reads this arrive publishes are ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync; a
    // release at cluster scope would first drain the warp's outstanding global stores (MEMBAR, seen as stall_membar in ncu)
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster_release(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
// the load lands in THIS CTA's shared memory, its bytes are counted on the pair leader's barrier
__device__ __forceinline__ void tma_load_3d_pair(uint32_t dst, const CUtensorMap* map, uint32_t cluster_bar, int c0, int c1,
                                                 int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(cluster_bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {   // arrives on `bar` of BOTH CTAs of the pair
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
        "h"((uint16_t)3)
        : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                 uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
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

// CG = 2: the tile is 2*BM x BN, computed by a CTA pair with tcgen05 cta_group::2 (UMMA M = 256).  Each CTA stages its
// own BM rows of A and HALF of the B tile (BN/2 rows) and holds its BM x BN half of the accumulator in its own TMEM,
// so per MMA a CTA's shared memory is read for 4 KB of A + BN/2 * 32 B of B instead of 4 KB + BN * 32 B: at BN = 256
// that is 8 KB per 128 tensor clocks (50 % of the 128 B/clk port) where the single-CTA 128x128 tile needs 8 KB per 64.
template <int NTERMS, int BN, int CG>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_ROWS = BN / CG;
    static constexpr int B_BYTES = B_ROWS * BK * 2;
    static constexpr int STAGE_BYTES = (NTERMS == 3 ? 2 : 1) * (A_BYTES + B_BYTES);
    static constexpr int EPI_BYTES = NUM_EPI_WARPS * 32 * 32 * 4;   // one 32x32 fp32 transpose buffer per epilogue warp
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

__device__ __forceinline__ void prefetch_l2(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, float (&a)[8], float (&b)[8]) {
    uint32_t* r = reinterpret_cast<uint32_t*>(a);
    uint32_t* q = reinterpret_cast<uint32_t*>(b);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(q[0]),
          "=r"(q[1]), "=r"(q[2]), "=r"(q[3]), "=r"(q[4]), "=r"(q[5]), "=r"(q[6]), "=r"(q[7])
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

__device__ __forceinline__ void ld_f8_cg(const float* a, float (&t)[8]) {
    const float4 t0 = __ldcg(reinterpret_cast<const float4*>(a));
    const float4 t1 = __ldcg(reinterpret_cast<const float4*>(a + 4));
    t[0] = t0.x, t[1] = t0.y, t[2] = t0.z, t[3] = t0.w, t[4] = t1.x, t[5] = t1.y, t[6] = t1.z, t[7] = t1.w;
}

// Fused epilogue on 8 consecutive columns of one output row held in registers.  Every stage is its
// own short loop with the mode branch hoisted out of it: the executed instruction footprint stays
// small (a fully unrolled 32-column body with all modes inline thrashed the instruction cache).
// off0 = element offset of (row, col0); nvalid = columns inside N (8, or 4 at the edge on the VEC
// path; anything in 1..8 on the scalar path).
// epilogue stage bits (KParams::flags, held in ONE register for the whole kernel: testing p.<field> per
// slab cost an LDCU + UISETP + branch per stage per 8 columns)
enum : uint32_t {
    F_ALPHA = 1u << 0, F_BIAS = 1u << 1, F_PRE = 1u << 2, F_ACT = 1u << 3, F_AUX = 1u << 4, F_DROP = 1u << 5,
    F_SCALE = 1u << 6, F_RES = 1u << 7, F_F32 = 1u << 8, F_ATOMIC = 1u << 9, F_HI = 1u << 10, F_LO = 1u << 11,
    F_COLSUM = 1u << 12, F_VEC = 1u << 13,
    F_PREFAC = 1u << 15,   // out_pre holds the backward factor (pre_mode 1)
    F_TMA = 1u << 14,   // aligned AND rows / batches of the output do not overlap in memory: TMA-store epilogue
};

__device__ __forceinline__ void st_f8(float* o, const float (&v)[8], int nvalid) {
    *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
    if (nvalid > 4) *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void ld_f8(const float* a, float (&t)[8], int nvalid) {
    const float4 t0 = *reinterpret_cast<const float4*>(a);
    const float4 t1 = nvalid > 4 ? *reinterpret_cast<const float4*>(a + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    t[0] = t0.x, t[1] = t0.y, t[2] = t0.z, t[3] = t0.w, t[4] = t1.x, t[5] = t1.y, t[6] = t1.z, t[7] = t1.w;
}

// Fused epilogue on 8 consecutive columns of one output row held in registers.  Every stage is its
// own short loop behind one register-bit test: the executed instruction footprint stays small (a fully
// unrolled 32-column body with all modes inline thrashed the instruction cache).
// off0 = element offset of (row, col0); nvalid = columns inside N (8, or 4 at the edge on the VEC
// path; anything in 1..8 on the scalar path).
template <bool VEC>
__device__ __forceinline__ void epi_cols8(const KParams& p, const uint32_t F, float (&v)[8], long long off0, int col0,
                                          int nvalid, bool row_ok, const float* bias_ptr, int lane) {
    if (row_ok) {
        if (F & F_ALPHA) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] *= p.alpha;
        }
        if (bias_ptr) {
            if (VEC) {
                float t[8];
                ld_f8(bias_ptr + col0, t, nvalid);
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] += t[e];
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (e < nvalid) v[e] += bias_ptr[col0 + e];
            }
        }
        if (F & F_PRE) {
            if (VEC) {
                st_f8(p.out_pre + off0, v, nvalid);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (e < nvalid) p.out_pre[off0 + e] = v[e];
            }
        }
        if (F & F_ACT) {
            if (p.act == 1) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = gelu_erf(v[e]);
            } else if (p.act == 2) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = silu_f(v[e]);
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = v[e] / (1.0f + fabsf(v[e]));
            }
        }
        if (F & F_AUX) {
            float a[8];
            if (VEC) {
                ld_f8(p.aux + off0, a, nvalid);
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
        if (F & F_DROP) {
#pragma unroll
            for (int e = 0; e < 8; ++e)
                v[e] = hash_uniform(p.drop_seed, (unsigned long long)(off0 + e)) < p.drop_p ? 0.0f : v[e] * p.drop_scale;
        }
        if (F & F_SCALE) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] *= p.out_scale;
        }
        if (F & F_RES) {
            if (VEC) {
                float t[8];
                ld_f8(p.residual + off0, t, nvalid);
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] += t[e];
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e)
                    if (e < nvalid) v[e] += p.residual[off0 + e];
            }
        }
        if (nvalid < 8) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = e < nvalid ? v[e] : 0.0f;
        }
        if (F & F_F32) {
            float* o = p.out_f32 + off0;
            if (VEC) {
                if (F & F_ATOMIC) {
                    red_add_v4(o, v[0], v[1], v[2], v[3]);
                    if (nvalid > 4) red_add_v4(o + 4, v[4], v[5], v[6], v[7]);
                } else {
                    st_f8(o, v, nvalid);
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    if (e < nvalid) {
                        if (F & F_ATOMIC)
                            atomicAdd(o + e, v[e]);
                        else
                            o[e] = v[e];
                    }
                }
            }
        }
        if (F & F_HI) {
            bf16 h[8], l[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) split_bf16(v[e], h[e], l[e]);
            if (VEC) {
                auto pk = [](bf16 a, bf16 b) {
                    __nv_bfloat162 t = __halves2bfloat162(a, b);
                    return *reinterpret_cast<uint32_t*>(&t);
                };
                if (nvalid > 4) {
                    *reinterpret_cast<uint4*>(p.out_hi + off0) = make_uint4(pk(h[0], h[1]), pk(h[2], h[3]), pk(h[4], h[5]), pk(h[6], h[7]));
                    if (F & F_LO)
                        *reinterpret_cast<uint4*>(p.out_lo + off0) =
                            make_uint4(pk(l[0], l[1]), pk(l[2], l[3]), pk(l[4], l[5]), pk(l[6], l[7]));
                } else {
                    *reinterpret_cast<uint2*>(p.out_hi + off0) = make_uint2(pk(h[0], h[1]), pk(h[2], h[3]));
                    if (F & F_LO) *reinterpret_cast<uint2*>(p.out_lo + off0) = make_uint2(pk(l[0], l[1]), pk(l[2], l[3]));
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    if (e < nvalid) {
                        p.out_hi[off0 + e] = h[e];
                        if (F & F_LO) p.out_lo[off0 + e] = l[e];
                    }
                }
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = 0.0f;   // outside the matrix: contributes nothing to column sums
    }
    if (F & F_COLSUM) {   // warp-uniform
        int cl;
        const float s = warp_colsum8(v, lane, cl);
        if ((lane & 3) == 0 && cl < nvalid) atomicAdd(p.colsum + col0 + cl, s);
    }
}

// ------------------------------------------------------------------------------------------
// Coalesced epilogue.  tcgen05.ld hands every thread ONE accumulator row; written to global memory that way
// each warp instruction touches 32 different 128-byte lines with 16 bytes each (measured: every extra
// 62 MB output / residual / aux stream of the FFN GEMM cost 16 us = 3.9 TB/s).  Instead a warp transposes
// its 32-row x 32-column chunk through a 4 KB XOR-swizzled smem buffer, after which 8 lanes cover 128
// contiguous bytes of one row: every global access of the fused stages is a full line, the bias is loaded
// once per chunk, and column sums are accumulated in registers over the chunk's rows.
// ------------------------------------------------------------------------------------------
constexpr int EPI_CHUNK_FLOATS = 32 * 32;

template <bool VEC>
__device__ __forceinline__ void ld4(const float* a, float (&t)[4], int nv) {
    if (VEC) {
        const float4 q = *reinterpret_cast<const float4*>(a);
        t[0] = q.x, t[1] = q.y, t[2] = q.z, t[3] = q.w;
    } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) t[e] = e < nv ? a[e] : 0.0f;
    }
}
template <bool VEC>
__device__ __forceinline__ void st4(float* a, const float (&t)[4], int nv) {
    if (VEC) {
        *reinterpret_cast<float4*>(a) = make_float4(t[0], t[1], t[2], t[3]);
    } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (e < nv) a[e] = t[e];
    }
}

// st: this warp's 4 KB staging buffer.  v: the lane's row (TMEM lane) x 32 columns.  row0: output row of
// lane 0 inside the batch; base_off: element offset of the batch slot; col_base: first column of the chunk.
template <bool VEC, bool PREFAC_OK>
__device__ __forceinline__ void epi_chunk32(const KParams& p, const uint32_t F_, float* st, const float (&v)[32], int lane,
                                            int row0, long long base_off, int col_base, const float* bias_ptr) {
    const uint32_t F = PREFAC_OK ? F_ : (F_ & ~F_PREFAC);
#pragma unroll
    for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(st + lane * 32 + ((j ^ (lane & 7)) << 2)) =
            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int c4 = lane & 7, rsub = lane >> 3;
    const int col = col_base + c4 * 4;
    const int nv = p.n - col >= 4 ? 4 : (p.n - col > 0 ? p.n - col : 0);   // VEC: 0 or 4
    float b4[4] = {0.f, 0.f, 0.f, 0.f};
    if (bias_ptr && nv > 0) ld4<VEC>(bias_ptr + col, b4, nv);
    float cs[4] = {0.f, 0.f, 0.f, 0.f};
#define EPI_ALL(stmt)                         \
    _Pragma("unroll") for (int i = 0; i < 4; ++i) \
        _Pragma("unroll") for (int e = 0; e < 4; ++e) { stmt; }
    const int last_row = p.m_per_batch - 1;
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
        // Stage-major over 4 rows x 4 columns: 16 independent dependency chains per stage (the MUFU / erf
        // latency of the activations is hidden inside the thread), rows past the matrix are computed on a
        // clamped address and only their stores are predicated off.
        float x[4][4], ax[4][4], rs[4][4];
        long long off[4];
        bool ok[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = (h * 4 + i) * 4 + rsub;
            const int row = row0 + r;
            ok[i] = row <= last_row && nv > 0;
            off[i] = base_off + (long long)(row <= last_row ? row : last_row) * p.ldc + (nv > 0 ? col : 0);
        }
        if (F & F_AUX) {
#pragma unroll
            for (int i = 0; i < 4; ++i) ld4<VEC>(p.aux + off[i], ax[i], nv > 0 ? nv : 4);
        }
        if (F & F_RES) {
#pragma unroll
            for (int i = 0; i < 4; ++i) ld4<VEC>(p.residual + off[i], rs[i], nv > 0 ? nv : 4);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = (h * 4 + i) * 4 + rsub;
            const float4 q = *reinterpret_cast<const float4*>(st + r * 32 + ((c4 ^ (r & 7)) << 2));
            x[i][0] = q.x, x[i][1] = q.y, x[i][2] = q.z, x[i][3] = q.w;
        }
        if (F & F_ALPHA) { EPI_ALL(x[i][e] *= p.alpha) }
        EPI_ALL(x[i][e] += b4[e])
        if ((F & F_PRE) && !(F & F_PREFAC)) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (ok[i]) st4<VEC>(p.out_pre + off[i], x[i], nv);
        }
        float dd[4][4];   // pre_mode 1: d(value)/d(pre-activation), stored in place of the pre-activation
        if (F & F_PREFAC) {
            if (p.act == 1) { EPI_ALL(dd[i][e] = gelu_erf_grad(x[i][e])) }
            else if (p.act == 2) { EPI_ALL(dd[i][e] = silu_grad(x[i][e])) }
            else { EPI_ALL(const float d = 1.0f + fabsf(x[i][e]); dd[i][e] = rcp_approx(d * d)) }
        }
        if (F & F_ACT) {
            if (p.act == 1) { EPI_ALL(x[i][e] = gelu_erf(x[i][e])) }
            else if (p.act == 2) { EPI_ALL(x[i][e] = silu_f(x[i][e])) }
            else { EPI_ALL(x[i][e] = x[i][e] * rcp_approx(1.0f + fabsf(x[i][e]))) }
        }
        if (F & F_AUX) {
            if (p.aux_mode == 1) { EPI_ALL(x[i][e] *= gelu_erf_grad(ax[i][e])) }
            else if (p.aux_mode == 2) { EPI_ALL(x[i][e] *= silu_grad(ax[i][e])) }
            else if (p.aux_mode == 4) { EPI_ALL(x[i][e] *= ax[i][e]) }
            else { EPI_ALL(const float d = 1.0f + fabsf(ax[i][e]); x[i][e] *= rcp_approx(d * d)) }
        }
        if (F & F_DROP) {
            EPI_ALL(const float kf = hash_uniform(p.drop_seed, (unsigned long long)(off[i] + e)) < p.drop_p ? 0.0f : p.drop_scale;
                    x[i][e] *= kf; if (F & F_PREFAC) dd[i][e] *= kf)
        }
        if (F & F_SCALE) { EPI_ALL(x[i][e] *= p.out_scale; if (F & F_PREFAC) dd[i][e] *= p.out_scale) }
        if (F & F_PREFAC) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (ok[i]) st4<VEC>(p.out_pre + off[i], dd[i], nv);
        }
        if (F & F_RES) { EPI_ALL(x[i][e] += rs[i][e]) }
        EPI_ALL(x[i][e] = (ok[i] && e < nv) ? x[i][e] : 0.0f)
        if (F & F_F32) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (!ok[i]) continue;
                float* o = p.out_f32 + off[i];
                if (F & F_ATOMIC) {
                    if (VEC) {
                        red_add_v4(o, x[i][0], x[i][1], x[i][2], x[i][3]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e)
                            if (e < nv) atomicAdd(o + e, x[i][e]);
                    }
                } else {
                    st4<VEC>(o, x[i], nv);
                }
            }
        }
        if (F & F_HI) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                bf16 hh[4], ll[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) split_bf16(x[i][e], hh[e], ll[e]);
                if (!ok[i]) continue;
                if (VEC) {
                    auto pk = [](bf16 a, bf16 b) {
                        __nv_bfloat162 t = __halves2bfloat162(a, b);
                        return *reinterpret_cast<uint32_t*>(&t);
                    };
                    *reinterpret_cast<uint2*>(p.out_hi + off[i]) = make_uint2(pk(hh[0], hh[1]), pk(hh[2], hh[3]));
                    if (F & F_LO) *reinterpret_cast<uint2*>(p.out_lo + off[i]) = make_uint2(pk(ll[0], ll[1]), pk(ll[2], ll[3]));
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (e < nv) {
                            p.out_hi[off[i] + e] = hh[e];
                            if (F & F_LO) p.out_lo[off[i] + e] = ll[e];
                        }
                    }
                }
            }
        }
        EPI_ALL(cs[e] += x[i][e])
    }
#undef EPI_ALL
    if (F & F_COLSUM) {   // warp-uniform: rows of the chunk are summed over the 4 lane groups
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            cs[e] += __shfl_xor_sync(0xffffffffu, cs[e], 8);
            cs[e] += __shfl_xor_sync(0xffffffffu, cs[e], 16);
        }
        if (rsub == 0) {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (e < nv) atomicAdd(p.colsum + col + e, cs[e]);
        }
    }
    __syncwarp();   // the buffer is rewritten by the next chunk
}


// ------------------------------------------------------------------------------------------
// TMA-store epilogue (aligned outputs: F_VEC).  ncu on the transposing epilogue above (round 2, pair kernel,
// profiles/ncu_gemm_pair_r02_summary.txt): 690 warp instructions per 32 x 32 chunk at one issue per ~6 clocks and two
// warps per scheduler = 22 k clocks per 128 x 256 tile - as long as the tile's MMAs, and fully exposed behind the
// last tile of every launch.  Here a thread keeps the accumulator row tcgen05.ld gave it (32 independent dependency
// chains per stage, no per-element address arithmetic or bounds predicates), writes it once into the warp's 4 KB
// staging buffer in the swizzle pattern of the output tensor map, and one lane issues a bulk tensor store (or
// reduce-add) - the TMA unit clips rows / columns outside the matrix and writes whole lines.  The aux / residual
// inputs are still read with coalesced 16-byte loads (8 lanes per 128-byte row segment), issued before the TMEM
// load, and turned into rows through the same buffer.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map), "r"(src), "r"(c0),
                 "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void tma_reduce_add_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
    asm volatile("cp.reduce.async.bulk.tensor.3d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
                 "r"(src), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct EpiMaps {
    const CUtensorMap* f32;
    const CUtensorMap* pre;
    const CUtensorMap* hi;
    const CUtensorMap* lo;
};

// coalesced read of a 32 x 32 fp32 chunk: instruction i covers rows 4 i + (lane >> 3), lane & 7 picks the 16-byte unit
__device__ __forceinline__ void chunk_ldg(const KParams& p, const float* src, long long base_off, int row0, int col_base,
                                          int lane, float4 (&t)[8]) {
    const int c4 = lane & 7, rsub = lane >> 3;
    const int col = col_base + c4 * 4;
    const bool col_ok = col < p.n;   // n % 4 == 0 on this path
    const int last_row = p.m_per_batch - 1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = row0 + i * 4 + rsub;
        const long long off = base_off + (long long)(row <= last_row ? row : last_row) * p.ldc + (col_ok ? col : 0);
        t[i] = __ldcg(reinterpret_cast<const float4*>(src + off));
    }
}
// ... and its transposition into "lane = row": through the staging buffer, same XOR swizzle as the stores
__device__ __forceinline__ void chunk_to_rows_store(float* st, int lane, const float4 (&t)[8]) {
    const int c4 = lane & 7, rsub = lane >> 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = i * 4 + rsub;
        *reinterpret_cast<float4*>(st + r * 32 + ((c4 ^ (r & 7)) << 2)) = t[i];
    }
}
__device__ __forceinline__ float4 chunk_row_unit(const float* st, int lane, int j) {
    return *reinterpret_cast<const float4*>(st + lane * 32 + ((j ^ (lane & 7)) << 2));
}

__device__ __forceinline__ uint32_t pack_bf16x2(bf16 a, bf16 b) {
    __nv_bfloat162 t = __halves2bfloat162(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
}

// v: the lane's accumulator row (row0 + lane inside the batch), 32 columns from col_base.  st / st_addr: this warp's
// 4 KB staging buffer (generic pointer / shared address, 1024-byte aligned).  `pending`: a bulk store of this warp may
// still be reading the buffer (lane 0 is the only lane that commits and waits).
template <int EMODE>
__device__ __forceinline__ void epi_row32(const KParams& p, const uint32_t F_, float* st, uint32_t st_addr, float (&v)[32],
                                          int lane, int row0, int cbatch, long long base_off, int col_base,
                                          const float* bias_ptr, const EpiMaps& maps, float4 (&t)[8]) {
    // compile-time dead stages: no input stages without EPI_IN; the backward-factor output (pre_mode 1, forward GEMMs
    // only) exists in the input-free instantiation - the two sets of temporaries never share a register budget
    const uint32_t F = EMODE == 1 ? (F_ & ~F_PREFAC) : (EMODE == 2 ? (F_ & ~(F_AUX | F_RES)) : F_);
#define EPI32(stmt) _Pragma("unroll") for (int e = 0; e < 32; ++e) { stmt; }
    auto buffer_free = [&]() {   // every earlier bulk store of this warp has read the buffer, every lane is past its own reads
        if (lane == 0 && !(p.debug & 64)) bulk_wait_read0();
        __syncwarp();
    };
    const int row = row0 + lane;
    const bool row_ok = row < p.m_per_batch;
    const long long off_row = base_off + (long long)(row_ok ? row : p.m_per_batch - 1) * p.ldc + col_base;   // dropout counter
    if (F & F_ALPHA) { EPI32(v[e] *= p.alpha) }
    if (bias_ptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (col_base + 4 * j < p.n) {   // warp-uniform
                const float4 b = *reinterpret_cast<const float4*>(bias_ptr + col_base + 4 * j);
                v[4 * j] += b.x, v[4 * j + 1] += b.y, v[4 * j + 2] += b.z, v[4 * j + 3] += b.w;
            }
        }
    }
    auto store_pre = [&](const float (&w)[32]) {
        buffer_free();
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(st + lane * 32 + ((j ^ (lane & 7)) << 2)) = make_float4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_3d(maps.pre, st_addr, col_base, row0, cbatch);
            bulk_commit();
        }
    };
    if ((F & F_PRE) && !(F & F_PREFAC)) store_pre(v);
    if (F & F_PREFAC) {
        // activation and its derivative from one sigmoid / erf; the keep / drop decision scales both; the factor goes out
        // in place of the pre-activation, so that the backward GEMM's epilogue is one multiply per element.  Two halves
        // of 16 columns: only 16 factor registers are live next to the 32 accumulators and the prefetched input chunk.
        const uint32_t thr = p.drop_thr, s_lo = (uint32_t)p.drop_seed, s_hi = (uint32_t)(p.drop_seed >> 32);
        const unsigned long long o0 = (unsigned long long)off_row;
        const bool no_carry = (uint32_t)o0 <= 0xFFFFFFFFu - 32u;
        const uint32_t f0 = hash_fold(o0);
        const float oscale = (F & F_SCALE) ? p.out_scale : 1.0f;
        buffer_free();
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            float d[16];
            // one mode branch around 16 independent chains
#define PREFAC16(stmt) _Pragma("unroll") for (int i = 0; i < 16; ++i) { const int e = hf * 16 + i; const float x = v[e]; float a, g; stmt; v[e] = a; d[i] = g; }
            if (p.act == 1) {
                PREFAC16(const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
                         g = cdf + x * 0.39894228040143267794f * __expf(-0.5f * x * x); a = x * cdf)
            } else if (p.act == 2) {
                PREFAC16(const float sg = sigmoidf_(x); g = sg * (1.0f + x * (1.0f - sg)); a = x * sg)
            } else {
                PREFAC16(const float r = rcp_approx(1.0f + fabsf(x)); g = r * r; a = x * r)
            }
#undef PREFAC16
            if (F & F_DROP) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int e = hf * 16 + i;
                    const float k = hash_drop(no_carry ? f0 + (uint32_t)e : hash_fold(o0 + e), s_lo, s_hi, thr) ? 0.0f : p.drop_scale * oscale;
                    v[e] *= k;
                    d[i] *= k;
                }
            } else if (F & F_SCALE) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[hf * 16 + i] *= oscale, d[i] *= oscale;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
                *reinterpret_cast<float4*>(st + lane * 32 + (((hf * 4 + j) ^ (lane & 7)) << 2)) =
                    make_float4(d[4 * j], d[4 * j + 1], d[4 * j + 2], d[4 * j + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_3d(maps.pre, st_addr, col_base, row0, cbatch);
            bulk_commit();
        }
    } else if (F & F_ACT) {
        if (p.act == 1) { EPI32(v[e] = gelu_erf(v[e])) }
        else if (p.act == 2) { EPI32(v[e] = silu_f(v[e])) }
        else { EPI32(v[e] = v[e] * rcp_approx(1.0f + fabsf(v[e]))) }
    }
    if (F & F_AUX) {
        // t holds the aux chunk (coalesced layout, loaded one chunk ahead)
        buffer_free();
        chunk_to_rows_store(st, lane, t);
        __syncwarp();
        float a[32];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 q = chunk_row_unit(st, lane, j);
            a[4 * j] = q.x, a[4 * j + 1] = q.y, a[4 * j + 2] = q.z, a[4 * j + 3] = q.w;
        }
        __syncwarp();
        // one mode branch around 32 independent chains (inside the unit loop the three variants were interleaved eight
        // times: 4 chains per segment and an instruction stream scattered over 1600 slots - stall_no_inst in ncu)
        if (p.aux_mode == 1) { EPI32(v[e] *= gelu_erf_grad(a[e])) }
        else if (p.aux_mode == 2) { EPI32(v[e] *= silu_grad(a[e])) }
        else if (p.aux_mode == 4) { EPI32(v[e] *= a[e]) }
        else { EPI32(const float d = 1.0f + fabsf(a[e]); v[e] *= rcp_approx(d * d)) }
        if (F & F_RES) chunk_ldg(p, p.residual, base_off, row0, col_base, lane, t);   // both inputs: the second read is exposed
    }
    if ((F & F_DROP) && !(F & F_PREFAC)) {
        // same keep / drop decision as hash_uniform(seed, offset) < p, with the 64-bit index folded once per row segment
        const uint32_t thr = p.drop_thr, s_lo = (uint32_t)p.drop_seed, s_hi = (uint32_t)(p.drop_seed >> 32);
        const unsigned long long o0 = (unsigned long long)off_row;
        if ((uint32_t)o0 <= 0xFFFFFFFFu - 32u) {   // no carry into the high word inside the segment
            const uint32_t f0 = hash_fold(o0);
            EPI32(v[e] = hash_drop(f0 + (uint32_t)e, s_lo, s_hi, thr) ? 0.0f : v[e] * p.drop_scale)
        } else {
            EPI32(v[e] = hash_drop(hash_fold(o0 + e), s_lo, s_hi, thr) ? 0.0f : v[e] * p.drop_scale)
        }
    }
    if ((F & F_SCALE) && !(F & F_PREFAC)) { EPI32(v[e] *= p.out_scale) }
    if (F & F_RES) {
        buffer_free();
        chunk_to_rows_store(st, lane, t);
        __syncwarp();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 a = chunk_row_unit(st, lane, j);
            v[4 * j] += a.x, v[4 * j + 1] += a.y, v[4 * j + 2] += a.z, v[4 * j + 3] += a.w;
        }
        __syncwarp();
    }
    if (F & (F_F32 | F_COLSUM)) {
        buffer_free();
        if (F & F_COLSUM) {
            // rows / columns outside the matrix must not reach the column sums (the stores clip them by themselves)
            EPI32(v[e] = (row_ok && col_base + e < p.n) ? v[e] : 0.0f)
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(st + lane * 32 + ((j ^ (lane & 7)) << 2)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        if (F & F_F32) fence_proxy_async_smem();
        __syncwarp();
        if ((F & F_F32) && lane == 0) {
            if (F & F_ATOMIC)
                tma_reduce_add_3d(maps.f32, st_addr, col_base, row0, cbatch);
            else
                tma_store_3d(maps.f32, st_addr, col_base, row0, cbatch);
            bulk_commit();
        }
        if (F & F_COLSUM) {   // lane = column: 32 conflict-free reads down the staged chunk
            float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int r = 0; r < 32; ++r) s4[r & 3] += st[r * 32 + ((((lane >> 2) ^ (r & 7)) << 2) | (lane & 3))];
            const float s = (s4[0] + s4[1]) + (s4[2] + s4[3]);
            if (col_base + lane < p.n) atomicAdd(p.colsum + col_base + lane, s);
            __syncwarp();
        }
    }
    if (F & F_HI) {
        // bf16 hi / lo rows: 64 bytes each, 64-byte swizzle (16-byte unit ^ ((row >> 1) & 3)), hi at +0, lo at +2 KB
        uint32_t h[16], l[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) {
            // two values per conversion (F2FP.BF16.PACK_AB on the ALU; a scalar cvt.rn.bf16 is an F2F on the quarter-rate pipe)
            __nv_bfloat162 hh = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            const float2 hf = __bfloat1622float2(hh);
            __nv_bfloat162 ll = __floats2bfloat162_rn(v[2 * e] - hf.x, v[2 * e + 1] - hf.y);
            h[e] = *reinterpret_cast<uint32_t*>(&hh);
            l[e] = *reinterpret_cast<uint32_t*>(&ll);
        }
        buffer_free();
        uint8_t* sb = reinterpret_cast<uint8_t*>(st);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int o = lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
            *reinterpret_cast<uint4*>(sb + o) = make_uint4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
            if (F & F_LO) *reinterpret_cast<uint4*>(sb + 2048 + o) = make_uint4(l[4 * j], l[4 * j + 1], l[4 * j + 2], l[4 * j + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_3d(maps.hi, st_addr, col_base, row0, cbatch);
            if (F & F_LO) tma_store_3d(maps.lo, st_addr + 2048, col_base, row0, cbatch);
            bulk_commit();
        }
    }
#undef EPI32
}

// EMODE: which epilogue stages are compiled in.  0 = all (single-CTA tiles); the pair kernel, where the register budget
// of the row-per-thread epilogue decides whether its 32 independent chains per stage fit without spilling, has two:
// 1 = with a second input (aux / residual: the prefetched input chunk, 64 registers) but without the backward-factor
// output, 2 = without inputs (forward GEMMs; pre_mode 1 lives here).
template <int NTERMS, bool A_MN, bool B_MN, int BN, int CG, int EMODE>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
               const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
               const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmPre,
               const __grid_constant__ CUtensorMap tmHi, const __grid_constant__ CUtensorMap tmLo, const KParams p) {
    using C = Cfg<NTERMS, BN, CG>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
    const uint32_t bar_base = smem_base + C::STAGES * C::STAGE_BYTES + C::EPI_BYTES;
    // barrier layout: full[STAGES], empty[STAGES], tmem_full[2], tmem_empty[2], then tmem ptr
    // CG == 2: every CTA of the pair has the whole set at the same offsets; `full` and `tmem_empty` are used in the
    // leader (cluster rank 0) only - it issues the MMAs - while `empty` and `tmem_full` are signalled in both CTAs by
    // multicast tcgen05.commit.
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
    volatile uint32_t* tmem_ptr_gen =
        reinterpret_cast<volatile uint32_t*>(smem_gen + C::STAGES * C::STAGE_BYTES + C::EPI_BYTES + 8 * (2 * C::STAGES + 4));
    const uint32_t tmem_ptr_addr = bar_base + 8u * (2 * C::STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int rank = CG == 2 ? (int)cluster_ctarank() : 0;

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
            mbar_init(tempty_bar(s), CG * NUM_EPI_WARPS);
            // cluster-launch-control ring (Sched): full = 16 response bytes; empty (used in the leader) = every consumer role
            // of the cluster: its 2 + NUM_EPI_WARPS, plus the peer's 1 + NUM_EPI_WARPS
            mbar_init(bar_base + CLC_OFF + 8u * s, 1);
            mbar_init(bar_base + CLC_OFF + 16u + 8u * s, CG == 2 ? 2 * NUM_EPI_WARPS + 3 : NUM_EPI_WARPS + 2);
        }
        if (p.dynamic) {
            mbar_expect_tx(bar_base + CLC_OFF, 16);
            mbar_expect_tx(bar_base + CLC_OFF + 8u, 16);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if (CG == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_ptr_addr),
                         "r"((uint32_t)C::TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_ptr_addr),
                         "r"((uint32_t)C::TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    if (CG == 2)
        cluster_sync_all();   // the peer's barriers are initialised before any remote arrive / complete_tx reaches them
    else
        __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_gen;
    // Programmatic dependent launch: everything above (barrier init, TMEM allocation, tensor-map prefetch) touches no
    // memory a predecessor writes, so under the programmatic-stream-serialization launch attribute it runs while the
    // previous kernel of the stream drains; from here on the predecessor is complete and its writes are visible.  The
    // trigger lets the NEXT kernel start its own prologue as soon as SMs free up behind this grid's last tiles.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp < EPI_WARP0) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS_CTRL));
    if (warp == 0) {
        // ================================ TMA producer ================================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            Sched sch(p, CG, rank, bar_base, false);
            Item it;
            auto tma_load = [&](uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
                if (CG == 2)
                    tma_load_3d_pair(dst, map, bar, c0, c1, c2);
                else
                    tma_load_3d(dst, map, bar, c0, c1, c2);
            };
            while (sch.next(it)) {
                const int batch = it.batch, m_blk = it.m_blk, n_blk = it.n_blk, kb0 = it.kb0, kb1 = it.kb1;
                const int bbatch = p.b_batch_idx ? p.b_batch_idx[batch] : (p.b_batched ? batch : 0);
                const int abatch = p.a_batched ? batch : 0;
                const int n_row0 = n_blk * BN + rank * C::B_ROWS;   // this CTA's rows of the B tile
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty_bar(stage), phase ^ 1u);
                    // CG == 2: both CTAs' bytes are counted on the leader's barrier (the MMA issuer waits there)
                    const uint32_t fb = CG == 2 ? mapa_u32(full_bar(stage), 0) : full_bar(stage);
                    if (p.debug & 4) {
                        if (rank == 0) mbar_arrive(full_bar(stage));
                        if (++stage == C::STAGES) {
                            stage = 0;
                            phase ^= 1u;
                        }
                        continue;
                    }
                    if (rank == 0) mbar_expect_tx(full_bar(stage), (uint32_t)(CG * C::STAGE_BYTES));
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
                            tma_load(sA + t * C::A_BYTES, ma, fb, kb * BK, m_blk * BM, abatch);
                        } else {
#pragma unroll
                            for (int i = 0; i < BM / 64; ++i)
                                tma_load(sA + t * C::A_BYTES + i * (BK * 128), ma, fb, m_blk * BM + i * 64, krow,
                                         p.k_spans_batches ? kslice : abatch);
                        }
                        if (!B_MN) {
                            tma_load(sB + t * C::B_BYTES, mb, fb, kb * BK, n_row0, bbatch);
                        } else {
#pragma unroll
                            for (int i = 0; i < C::B_ROWS / 64; ++i)
                                tma_load(sB + t * C::B_BYTES + i * (BK * 128), mb, fb, n_row0 + i * 64, krow,
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
    } else if (warp == 1 && rank == 0) {
        // ================================ MMA issuer (pair leader only when CG == 2) ==================================
        // The WHOLE warp walks the loops so that every address / descriptor stays warp-uniform (uniform
        // registers feed UTCHMMA directly); one elected lane issues the tcgen05 instructions.  With the loop
        // inside `if (lane == 0)` ptxas rebuilt each descriptor in vector registers and moved it over with
        // R2UR per MMA: ~230 instructions per k-block, slower than the 12 MMAs they issue (ncu, round 1).
        constexpr uint32_t idesc = make_idesc(CG * BM, BN, A_MN, B_MN);
        auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t acc_flag) {
            if (CG == 2)
                tc_mma_bf16_pair(d, da, db, idesc, acc_flag);
            else
                tc_mma_bf16(d, da, db, idesc, acc_flag);
        };
        auto commit = [&](uint32_t bar) {
            if (CG == 2)
                tc_commit_pair(bar);
            else
                tc_commit(bar);
        };
        // K-major SW128: 8-row groups 1024 B apart (SBO), k-step = +32 B inside the swizzle row
        // MN-major SW128: 64-element MN atoms BK*128 B apart (LBO), 8 k-rows = 1024 B (SBO),
        //                 k-step of 16 rows = +2048 B
        constexpr uint32_t A_LBO = A_MN ? BK * 128 : 0, B_LBO = B_MN ? BK * 128 : 0;
        constexpr uint32_t A_KSTEP = A_MN ? UMMA_K * 128 : UMMA_K * 2, B_KSTEP = B_MN ? UMMA_K * 128 : UMMA_K * 2;
        constexpr uint32_t A_TERMS = (NTERMS == 3 ? 2 : 1) * C::A_BYTES;
        // descriptors differ only in the 14-bit start-address field (bytes >> 4): plain 64-bit adds
        const uint64_t dA0 = make_desc(smem_base, A_LBO, 1024);
        const uint64_t dB0 = make_desc(smem_base + A_TERMS, B_LBO, 1024);
        uint64_t dA = dA0, dB = dB0;
        int stage = 0;
        uint32_t phase = 0;
        int local_iter = 0;
        Sched sch(p, CG, rank, bar_base, true);
        Item it;
        for (; sch.next(it); ++local_iter) {
            const int kb0 = it.kb0, kb1 = it.kb1;
            const int acc = local_iter & 1;
            const uint32_t acc_phase = (uint32_t)(local_iter >> 1) & 1u;
            mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
            tc_fence_after();
            const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
            uint32_t accumulate = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(full_bar(stage), phase);
                tc_fence_after();
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < ((p.debug & 2) ? 0 : BK / UMMA_K); ++k) {
                        const uint64_t dAh = dA + (uint64_t)((k * A_KSTEP) >> 4);
                        const uint64_t dBh = dB + (uint64_t)((k * B_KSTEP) >> 4);
                        mma(tmem_d, dAh, dBh, k == 0 ? accumulate : 1u);
                        if (NTERMS == 3) {
                            mma(tmem_d, dAh, dBh + (uint64_t)(C::B_BYTES >> 4), 1);
                            mma(tmem_d, dAh + (uint64_t)(C::A_BYTES >> 4), dBh, 1);
                        }
                    }
                    commit(empty_bar(stage));
                    if (kb == kb1 - 1) commit(tfull_bar(acc));
                }
                __syncwarp();
                accumulate = 1;
                dA += (uint64_t)(C::STAGE_BYTES >> 4);
                dB += (uint64_t)(C::STAGE_BYTES >> 4);
                if (++stage == C::STAGES) {
                    stage = 0;
                    phase ^= 1u;
                    dA = dA0;
                    dB = dB0;
                }
            }
        }
    } else if (warp == 2) {
        // ================================ scheduler (dynamic schedule only) ================================
        // One thread per CTA follows the response ring: it re-arms each slot's barrier after a response has landed and
        // stops at the first failed cancellation; the pair leader's thread also issues the queries (multicast to both
        // CTAs), each as soon as the slot's previous response has been read by every consumer role of the cluster.
        if (p.dynamic && lane == 0) {
            const uint32_t clc_full0 = bar_base + CLC_OFF, clc_empty0 = bar_base + CLC_OFF + 16u;
            const uint32_t resp0 = bar_base + CLC_RESP_OFF;
            if (rank == 0) clc_issue(resp0, clc_full0);
            for (uint32_t use = 0;;) {
                const uint32_t slot = use & 1u;
                mbar_wait(clc_full0 + 8u * slot, (use >> 1) & 1u);
                int ctaid;
                const bool valid = clc_read(resp0 + 16u * slot, ctaid);
                mbar_expect_tx(clc_full0 + 8u * slot, 16);   // arm the slot's next use (issued only after its empty barrier)
                ++use;
                if (!valid) break;
                if (rank == 0) {
                    if (use >= 2) mbar_wait(clc_empty0 + 8u * (use & 1u), ((use - 2) >> 1) & 1u);
                    clc_issue(resp0 + 16u * (use & 1u), clc_full0 + 8u * (use & 1u));
                }
            }
        }
    }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_EPI));
        // ================================ epilogue ====================================
        // 8 warps: TMEM lane quadrant q = warp & 3 (hardware restriction), the two warps of a
        // quadrant take alternate 32-column chunks.  A thread owns ONE output row (its TMEM lane)
        // and 32 consecutive columns of it: all global traffic is 16-byte vectors on that row.
        const int q = warp & 3;
        const int half = (warp - EPI_WARP0) >> 2;
        int local_iter = 0;
        uint32_t F = p.flags;
        constexpr bool EPI_IN = EMODE != 2;
        if (!EPI_IN) F &= ~(F_AUX | F_RES);   // (the host never selects this instantiation with either input set)
        asm volatile("" : "+r"(F));   // pin the stage bits in a register (else ptxas re-reads the constant bank per test)
        Sched sch(p, CG, rank, bar_base, true);
        Item it;
        const uint32_t tempty_remote0 = CG == 2 ? mapa_u32(tempty_bar(0), 0) : 0u;   // the leader's tmem_empty barriers
        float* const epi_buf =
            reinterpret_cast<float*>(smem_gen + C::STAGES * C::STAGE_BYTES) + (warp - EPI_WARP0) * EPI_CHUNK_FLOATS;
        const uint32_t epi_buf_addr = smem_base + C::STAGES * C::STAGE_BYTES + (warp - EPI_WARP0) * (EPI_CHUNK_FLOATS * 4);
        const EpiMaps emaps{&tmC, &tmPre, &tmHi, &tmLo};
        float* const ws_mine = p.sk_ws + (size_t)blockIdx.x * (BM * BN);
        for (; sch.next(it); ++local_iter) {
            const int split = it.split, batch = it.batch, m_blk = it.m_blk, n_blk = it.n_blk;
            const int acc = local_iter & 1;
            const uint32_t acc_phase = (uint32_t)(local_iter >> 1) & 1u;
            const int cbatch = p.c_batch_idx ? p.c_batch_idx[batch] : batch;
            const long long base_off = (long long)cbatch * p.c_batch_stride;
            if ((F & (F_AUX | F_RES)) && it.mode != 1) {
                // the tile's MMAs are still running: pull this warp's aux / residual lines (cold in HBM inside a
                // training step) into L2 now, one 128-byte line per lane (= row) and 32-column chunk
                const int prow = m_blk * BM + q * 32 + lane;
                if (prow < p.m_per_batch) {
                    const long long poff = base_off + (long long)prow * p.ldc + n_blk * BN;
                    for (int c = half; c < BN / 32; c += 2) {
                        if (n_blk * BN + c * 32 >= p.n) break;
                        if (F & F_AUX) prefetch_l2(p.aux + poff + c * 32);
                        if (F & F_RES) prefetch_l2(p.residual + poff + c * 32);
                    }
                }
            }
            // TMA-store epilogue: the aux (else residual) chunk is read one chunk ahead of its use - the first one
            // here, under the tile's MMAs - so that no chunk waits for a memory round trip
            const bool tin_on = EMODE != 2 && (F & F_TMA) && (F & (F_AUX | F_RES)) && it.mode != 1 && !(p.debug & 1);
            const float* const tin_src = (F & F_AUX) ? p.aux : p.residual;
            float4 tin[8], tnext[8];
            if (tin_on && n_blk * BN + half * 32 < p.n)
                chunk_ldg(p, tin_src, base_off, m_blk * BM + q * 32, n_blk * BN + half * 32, lane, tnext);
            mbar_wait(tfull_bar(acc), acc_phase);
            tc_fence_after();
            if (it.mode == 2 && !(p.debug & (8 | 16))) {
                // stream-K owner: the partials of the following CTAs were written at the START of their ranges
                if (lane == 0) {
                    const int want = NUM_EPI_WARPS * it.n_contrib;
                    const long long t0 = clock64();
                    while (ld_acquire_gpu(p.sk_flags + blockIdx.x) < want) {
                        if (clock64() - t0 > (1ll << 32)) {
                            printf("b2t_gemm: stream-K owner %d timed out waiting for %d partials\n", blockIdx.x, it.n_contrib);
                            __trap();
                        }
                    }
                }
                __syncwarp();
            }
            const int bslot = p.b_batch_idx ? p.b_batch_idx[batch] : (p.b_batched ? batch : 0);
            const float* bias_ptr =
                (p.bias != nullptr && split == 0) ? p.bias + (long long)bslot * p.bias_batch_stride : nullptr;
#pragma unroll 1
            for (int c = half; c < BN / 32; c += 2) {
                float v[32];
                if (tin_on) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) tin[i] = tnext[i];
                    if (c + 2 < BN / 32 && n_blk * BN + (c + 2) * 32 < p.n)
                        chunk_ldg(p, tin_src, base_off, m_blk * BM + q * 32, n_blk * BN + (c + 2) * 32, lane, tnext);
                }
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + c * 32);
                tmem_ld_32x32(taddr, v);
                if (c + 2 >= BN / 32) {
                    // this warp's share of the accumulator is in registers: hand the TMEM stage back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (CG == 2)
                            mbar_arrive_cluster(tempty_remote0 + 8u * acc);
                        else
                            mbar_arrive(tempty_bar(acc));
                    }
                }
                // workspace layout [chunk c][quadrant q][j][lane][4]: private to the (contributor, owner) thread pair that
                // shares (warp, lane), so it is chosen for the memory system - every warp instruction moves 512
                // contiguous bytes
                const int ws_off = ((c * 4 + q) * 8 * 32 + lane) * 4;
                if (it.mode == 1) {
                    if (p.debug & (8 | 32)) continue;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        __stcg(reinterpret_cast<float4*>(ws_mine + ws_off + 128 * j),
                               make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
                    continue;
                }
                const int col_base = n_blk * BN + c * 32;
                if (col_base >= p.n || (p.debug & 1)) continue;   // warp-uniform
                if (it.mode == 2 && !(p.debug & (8 | 32))) {
                    for (int j = 1; j <= it.n_contrib; ++j) {
                        const float* w = ws_mine + (size_t)j * CG * (BM * BN) + ws_off;
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            const float4 t = __ldcg(reinterpret_cast<const float4*>(w + 128 * g));
                            v[4 * g] += t.x, v[4 * g + 1] += t.y, v[4 * g + 2] += t.z, v[4 * g + 3] += t.w;
                        }
                    }
                }
                if (F & F_TMA)
                    epi_row32<EMODE>(p, F, epi_buf, epi_buf_addr, v, lane, m_blk * BM + q * 32, cbatch, base_off, col_base, bias_ptr, emaps, tin);
                else if (F & F_VEC)
                    epi_chunk32<true, EMODE != 1>(p, F, epi_buf, v, lane, m_blk * BM + q * 32, base_off, col_base, bias_ptr);
                else
                    epi_chunk32<false, EMODE != 1>(p, F, epi_buf, v, lane, m_blk * BM + q * 32, base_off, col_base, bias_ptr);
            }
            if (p.debug & 8) {
            } else if (it.mode == 1) {
                __threadfence();
                __syncwarp();
                if (lane == 0) red_release_gpu_add(p.sk_flags + it.owner * CG + rank, 1);
            } else if (it.mode == 2) {
                // every epilogue warp is past its wait: clear the flag for the next launch on this stream
                asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
                if (threadIdx.x == EPI_WARP0 * 32) p.sk_flags[blockIdx.x] = 0;
            }
        }
        if (lane == 0) bulk_wait0();   // this warp's bulk stores are complete before the CTA may leave
    }

    tc_fence_before();
    if (CG == 2)
        cluster_sync_all();   // neither CTA leaves (or frees TMEM) while its peer's MMAs / arrives can still touch it
    else
        __syncthreads();
    if (warp == 1) {
        __syncwarp();
        if (CG == 2)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                         : "memory");
        else
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

// A tensor map is a pure function of (pointer, extents, strides, box): a training step re-issues the same few hundred
// GEMMs on the same buffers every iteration (caching allocator), so the encoded maps are kept in a direct-mapped
// table instead of calling cuTensorMapEncodeTiled four times per launch.
struct MapKey {
    const void* ptr;
    int64_t d0, d1, d2, s1, s2;
    int64_t box1;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && s1 == o.s1 && s2 == o.s2 && box1 == o.box1;
    }
};
struct MapSlot {
    MapKey key;
    CUtensorMap map;
    bool used;
};
constexpr int MAP_CACHE_SLOTS = 8192;
MapSlot* map_cache() {
    static thread_local MapSlot* tab = nullptr;   // per host thread: no locking on the launch path
    if (!tab) tab = (MapSlot*)calloc(MAP_CACHE_SLOTS, sizeof(MapSlot));
    return tab;
}
inline uint64_t mix64(uint64_t x) {
    x ^= x >> 33, x *= 0xff51afd7ed558ccdULL, x ^= x >> 33, x *= 0xc4ceb9fe1a85ec53ULL, x ^= x >> 33;
    return x;
}

int encode_map(CUtensorMap* map, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t s1, int64_t s2, int box1,
               const char* what);

// 3-D bf16 map, d0 contiguous, box {64, box1, 1}, 128B swizzle, zero OOB fill
int make_map(CUtensorMap* map, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t s1, int64_t s2, int box1,
             const char* what) {
    MapSlot* tab = map_cache();
    const MapKey key{ptr, d0, d1, d2, s1, s2, box1};
    if (tab) {
        uint64_t h = mix64((uint64_t)(uintptr_t)ptr ^ mix64((uint64_t)d1 * 0x9e3779b97f4a7c15ULL + (uint64_t)s1) ^
                           mix64((uint64_t)d0 + 31 * (uint64_t)d2 + 977 * (uint64_t)s2 + 7919 * (uint64_t)box1));
        MapSlot& sl = tab[h & (MAP_CACHE_SLOTS - 1)];
        if (sl.used && sl.key == key) {
            *map = sl.map;
            return 0;
        }
        const int rc = encode_map(map, ptr, d0, d1, d2, s1, s2, box1, what);
        if (rc == 0) {
            sl.key = key;
            sl.map = *map;
            sl.used = true;
        }
        return rc;
    }
    return encode_map(map, ptr, d0, d1, d2, s1, s2, box1, what);
}

// box1 > 0: bf16 operand map, box {64, box1, 1}, 128-byte swizzle.  box1 == MAP_OUT_F32 / MAP_OUT_BF16: output maps of the
// TMA-store epilogue, box {32, 32, 1} (one warp chunk), fp32 rows = 128-byte swizzle, bf16 rows = 64-byte swizzle.
constexpr int MAP_OUT_F32 = -1, MAP_OUT_BF16 = -2;
int encode_map(CUtensorMap* map, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t s1, int64_t s2, int box1,
               const char* what) {
    EncodeTiledFn fn = get_encode_fn();
    B2T_CHECK_ARG(fn != nullptr, "b2t_gemm: cuTensorMapEncodeTiled entry point not found");
    B2T_CHECK_ARG(((uintptr_t)ptr & 15) == 0, "b2t_gemm: %s not 16-byte aligned", what);
    const bool f32 = box1 == MAP_OUT_F32;
    const int esz = f32 ? 4 : 2;
    B2T_CHECK_ARG((s1 * esz) % 16 == 0 && (s2 * esz) % 16 == 0, "b2t_gemm: %s strides (%lld, %lld) must be multiples of 16 bytes",
                  what, (long long)s1, (long long)s2);
    cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
    cuuint64_t strides[2] = {(cuuint64_t)s1 * esz, (cuuint64_t)s2 * esz};
    cuuint32_t box[3] = {box1 > 0 ? 64u : 32u, box1 > 0 ? (cuuint32_t)box1 : 32u, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims,
                    strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    box1 == MAP_OUT_BF16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                    box1 > 0 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B2T_CHECK_ARG(r == CUDA_SUCCESS,
                  "b2t_gemm: cuTensorMapEncodeTiled(%s) failed with %d (dims %lld %lld %lld strides %lld %lld box1 %d)", what,
                  (int)r, (long long)d0, (long long)d1, (long long)d2, (long long)s1, (long long)s2, box1);
    return 0;
}

// SMs the persistent grid may use.  B2T_GEMM_SMS / b2t_gemm_set_sm_limit leave SMs free for a concurrent
// collective: a persistent grid of one CTA per SM whose tiles are statically strided runs a second wave for
// the CTAs that could not get an SM while NCCL kernels hold some (measured at 2 GPUs: the step grows by about
// the all-reduce time).
}  // namespace
// gemm_tc.cu is compiled twice (build.sh): B2T_GEMM_PART 0 = the API + the bf16x3 (3-term) instantiations, 1 = the plain
// bf16 (1-term) instantiations behind b2t_gemm_part1 - 40 kernel instantiations in one translation unit took 5 minutes.
#ifndef B2T_GEMM_PART
#define B2T_GEMM_PART 0
#endif
#if B2T_GEMM_PART == 0
__attribute__((visibility("hidden"))) int b2t_g_sm_limit = 0;
#else
extern __attribute__((visibility("hidden"))) int b2t_g_sm_limit;
#endif
#if B2T_GEMM_PART == 0
__attribute__((visibility("hidden"))) int b2t_g_dynamic = -1;
#else
extern __attribute__((visibility("hidden"))) int b2t_g_dynamic;
#endif
#define g_sm_limit b2t_g_sm_limit
// internal to the library (hidden: not part of the C ABI of include/b2t_sm100.h)
extern "C" __attribute__((visibility("hidden"))) int b2t_gemm_part1(const b2t_gemm_args* a, const void* kparams, int bn, int cg, void* stream);
namespace {
int num_sms() {
    static int n = 0;
    if (!n) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        const char* e = getenv("B2T_GEMM_SMS");
        if (e && atoi(e) > 0 && g_sm_limit == 0) g_sm_limit = atoi(e);
    }
    return (g_sm_limit > 0 && g_sm_limit < n) ? g_sm_limit : n;
}

template <int NTERMS, bool A_MN, bool B_MN, int BN, int CG, int EMODE>
int launch(const b2t_gemm_args* a, const KParams& kp, cudaStream_t stream) {
    using C = Cfg<NTERMS, BN, CG>;
    CUtensorMap mAh, mAl, mBh, mBl;
    const int a_box1 = A_MN ? BK : BM;
    const int b_box1 = B_MN ? BK : C::B_ROWS;
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
    // output maps of the TMA-store epilogue (aligned outputs only; the kernel ignores them otherwise): columns x rows of
    // one batch x batch slots - rows past m_per_batch and columns past n are clipped by the TMA unit
    CUtensorMap mC = mAh, mPre = mAh, mHi = mAh, mLo = mAh;
    if (kp.flags & F_TMA) {
        const int64_t slots = a->c_batch_idx ? (int64_t)1 << 16 : a->batches;
        const int64_t s2 = slots > 1 ? a->c_batch_stride : (int64_t)a->m_per_batch * a->ldc;
        if (a->out_f32 && (rc = make_map(&mC, a->out_f32, a->n, a->m_per_batch, slots, a->ldc, s2, MAP_OUT_F32, "out_f32"))) return rc;
        if (a->out_pre && (rc = make_map(&mPre, a->out_pre, a->n, a->m_per_batch, slots, a->ldc, s2, MAP_OUT_F32, "out_pre"))) return rc;
        if (a->out_hi && (rc = make_map(&mHi, a->out_hi, a->n, a->m_per_batch, slots, a->ldc, s2, MAP_OUT_BF16, "out_hi"))) return rc;
        if (a->out_lo && (rc = make_map(&mLo, a->out_lo, a->n, a->m_per_batch, slots, a->ldc, s2, MAP_OUT_BF16, "out_lo"))) return rc;
    }
    auto kern = gemm_tc_kernel<NTERMS, A_MN, B_MN, BN, CG, EMODE>;
    {
        // per device: the attribute belongs to the function in ONE device's context
        static bool attr_set[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
    }
    const int units_max = num_sms() / CG;   // CTAs (CG == 1) or CTA pairs (CG == 2) of the persistent grid
    int units = kp.total_items < units_max ? kp.total_items : units_max;
    KParams kq = kp;
    {
        // Off by default: measured on the step it costs more than it gains where it could be measured (round 2, one box:
        // isolated 240-tile GEMMs +3 us of 72, conformer step 52.7 -> 54.1 ms at 1 GPU, 54.7 -> 56.8 ms at 2 GPUs with the
        // exposed all-reduce unchanged at 2.2-2.4 ms).  Its case is an 8-GPU box where NCCL kernels hold SMs for long.
        int& clc = b2t_g_dynamic;
        if (clc < 0) {
            const char* e = getenv("B2T_GEMM_CLC");
            clc = e ? atoi(e) : 0;
        }
        // dynamic schedule: one cluster per work item, handed out by cluster launch control (see Sched); the stream-K
        // ranges and an explicit SM limit keep the fixed persistent grid
        kq.dynamic = clc && !kq.stream_k && !(g_sm_limit > 0) && !kq.debug;
        if (kq.dynamic) units = kp.total_items;
    }
    if (kq.stream_k) {
        // the stream-K line is cut into one range per unit of the FULL grid (the host computed sk_tile0 for it); every
        // CTA is resident (grid <= #SMs, 1 CTA/SM), which the owners' waits rely on
        units = units_max;
        const size_t need = SK_FLAG_BYTES + (size_t)units * CG * BM * BN * sizeof(float);
        if (need > (size_t)a->sk_workspace_bytes || units * CG * (int)sizeof(int) > SK_FLAG_BYTES) {
            kq.stream_k = 0;   // workspace too small for this tile shape: plain persistent schedule
            units = kp.total_items < units_max ? kp.total_items : units_max;
        }
    }
    {
        static int pdl = -1;
        if (pdl < 0) {
            const char* e = getenv("B2T_GEMM_PDL");
            pdl = e ? atoi(e) : 1;
        }
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)(units * CG), 1, 1);
        cfg.blockDim = dim3(NUM_THREADS, 1, 1);
        cfg.dynamicSmemBytes = C::SMEM_BYTES;
        cfg.stream = stream;
        cudaLaunchAttribute at[2];
        int na = 0;
        if (CG > 1) {
            at[na].id = cudaLaunchAttributeClusterDimension;
            at[na].val.clusterDim.x = CG;
            at[na].val.clusterDim.y = 1;
            at[na].val.clusterDim.z = 1;
            ++na;
        }
        if (pdl) {   // the kernel's prologue may overlap the previous kernel's tail (griddepcontrol.wait in the kernel)
            at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        cfg.attrs = at;
        cfg.numAttrs = na;
        B2T_CUDA(cudaLaunchKernelEx(&cfg, kern, mAh, mAl, mBh, mBl, mC, mPre, mHi, mLo, kq));
    }
    B2T_CHECK_LAUNCH("b2t_gemm");
    return 0;
}

template <int NTERMS, int BN, int CG, int EMODE>
int dispatch_major2(const b2t_gemm_args* a, const KParams& kp, cudaStream_t s) {
    if (a->a_mn_major) {
        if (a->b_mn_major) return launch<NTERMS, true, true, BN, CG, EMODE>(a, kp, s);
        return launch<NTERMS, true, false, BN, CG, EMODE>(a, kp, s);
    }
    if (a->b_mn_major) return launch<NTERMS, false, true, BN, CG, EMODE>(a, kp, s);
    return launch<NTERMS, false, false, BN, CG, EMODE>(a, kp, s);
}
template <int NTERMS, int BN, int CG>
int dispatch_major(const b2t_gemm_args* a, const KParams& kp, cudaStream_t s) {
    if constexpr (CG == 2) {   // the specialised epilogues exist for the pair kernel only (40 instantiations in all)
        if (a->aux_mode || a->residual) return dispatch_major2<NTERMS, BN, CG, 1>(a, kp, s);
        return dispatch_major2<NTERMS, BN, CG, 2>(a, kp, s);
    } else {
        return dispatch_major2<NTERMS, BN, CG, 0>(a, kp, s);
    }
}

}  // namespace

#if B2T_GEMM_PART == 0
extern "C" int b2t_gemm_set_sm_limit(int sms) {
    g_sm_limit = sms > 0 ? sms : 0;
    return 0;
}

extern "C" int b2t_gemm_set_dynamic_schedule(int on) {
    b2t_g_dynamic = on ? 1 : 0;
    return 0;
}

extern "C" int64_t b2t_gemm_workspace_bytes(void) {
    return (int64_t)SK_FLAG_BYTES + (int64_t)256 * BM * 256 * (int64_t)sizeof(float);
}

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
    kp.pre_mode = a->pre_mode;
    B2T_CHECK_ARG(a->pre_mode == 0 || (a->pre_mode == 1 && a->out_pre && a->act >= 1 && a->act <= 3 && !a->aux_mode && !a->residual),
                  "b2t_gemm: pre_mode 1 needs out_pre and an activation, and excludes aux / residual");
    B2T_CHECK_ARG(a->aux_mode >= 0 && a->aux_mode <= 4 && a->act >= 0 && a->act <= 3, "b2t_gemm: bad act / aux_mode");
    kp.atomic = a->atomic;
    kp.split_k = a->split_k;
    kp.alpha = a->alpha;
    B2T_CHECK_ARG(a->drop_p >= 0.0f && a->drop_p < 1.0f, "b2t_gemm: drop_p must be in [0, 1)");
    kp.drop_p = a->drop_p;
    kp.drop_scale = 1.0f / (1.0f - a->drop_p);
    kp.drop_thr = drop_threshold(a->drop_p);
    kp.drop_seed = a->drop_seed;
    kp.out_scale = a->out_scale == 0.0f ? 1.0f : a->out_scale;
    {
        auto al16 = [](const void* q) { return q == nullptr || ((uintptr_t)q & 15) == 0; };
        auto al8 = [](const void* q) { return q == nullptr || ((uintptr_t)q & 7) == 0; };
        kp.vec4 = (a->n % 4 == 0) && (a->ldc % 4 == 0) && (a->c_batch_stride % 4 == 0) && al16(a->out_f32) &&
                  al16(a->out_pre) && al16(a->residual) && al16(a->aux) && al16(a->bias) && (a->bias_batch_stride % 4 == 0) &&
                  (a->out_hi == nullptr || (al16(a->out_hi) && al16(a->out_lo) && a->ldc % 8 == 0 && a->c_batch_stride % 8 == 0));
        (void)al8;
    }
    kp.flags = (kp.alpha != 1.0f ? F_ALPHA : 0u) | (kp.bias ? F_BIAS : 0u) | (kp.out_pre ? F_PRE : 0u) | (kp.act ? F_ACT : 0u) |
               (kp.aux_mode ? F_AUX : 0u) | (kp.drop_p > 0.0f ? F_DROP : 0u) | (kp.out_scale != 1.0f ? F_SCALE : 0u) |
               (kp.residual ? F_RES : 0u) | (kp.out_f32 ? F_F32 : 0u) | (kp.atomic ? F_ATOMIC : 0u) | (kp.out_hi ? F_HI : 0u) |
               (kp.out_lo ? F_LO : 0u) | (kp.colsum ? F_COLSUM : 0u) | (kp.vec4 ? F_VEC : 0u) | (kp.pre_mode == 1 ? F_PREFAC : 0u);
    {
        // overlapping-stride outputs (the folded unfold gradient: ldc < n) keep the per-element red.add path
        static int no_tma = -1;
        if (no_tma < 0) {
            const char* e = getenv("B2T_GEMM_NO_TMA_EPILOGUE");
            no_tma = e ? atoi(e) : 0;
        }
        const bool rows_apart = a->ldc >= a->n;
        const bool batches_apart = (a->batches == 1 && !a->c_batch_idx) || a->c_batch_stride >= (int64_t)(a->m_per_batch - 1) * a->ldc + a->n;
        if (kp.vec4 && rows_apart && batches_apart && !no_tma) kp.flags |= F_TMA;
    }

    {
        static int dbg = -1;
        if (dbg < 0) {
            const char* e = getenv("B2T_GEMM_DEBUG");
            dbg = e ? atoi(e) : 0;
        }
        kp.debug = dbg;
    }
    // Tile shape.  cta_group::2 (a CTA pair on 256 x 256 tiles, see Cfg) is what feeds the tensor pipe without hitting the
    // shared-memory port, so it is taken whenever the problem has 256-row blocks to give it; single-CTA tiles
    // (128 x 64/128/256) remain for narrow outputs and short batches (lm_head, day affine, unfold / pos-conv views).
    int bn = a->block_n;
    int cg = a->cta_group;
    B2T_CHECK_ARG(cg == 0 || cg == 1 || cg == 2, "b2t_gemm: cta_group must be 0 (auto), 1 or 2");
    {
        static int force_cg = -1;
        if (force_cg < 0) {
            const char* e = getenv("B2T_GEMM_CG");
            force_cg = e ? atoi(e) : 0;
        }
        if (cg == 0 && (force_cg == 1 || force_cg == 2)) cg = force_cg;
    }
    if (cg == 0) {
        // pairs when there are enough 256 x 256 work items to occupy a good part of the 74 pairs (few, large items lose to
        // many 128 x 128 tiles: measured on the 2048 x 1024 and 1024 x 1024 weight gradients, profiles/gemm_table_r02*)
        const long long items2 = (long long)ceil_div_i(ceil_div_i(a->m_per_batch, BM), 2) * ceil_div_i(a->n, 256) * a->batches * a->split_k;
        cg = (bn == 0 || bn == 256) && a->n >= 256 && a->n % 128 == 0 && a->m_per_batch > BM && num_sms() >= 2 && items2 >= 30 ? 2 : 1;
    }
    if (cg == 2 && (kp.debug & 7)) cg = 1;   // the component-timing switches exist for the single-CTA kernel only
    if (cg == 2) {
        B2T_CHECK_ARG(bn == 0 || bn == 256, "b2t_gemm: cta_group 2 uses block_n 256");
        bn = 256;
    }
    if (bn == 0) {
        bn = a->n <= 64 ? 64 : 128;
        // 128 x 256 tiles move 25 % fewer operand bytes per MMA through shared memory, but halve the tile count:
        // taken only when the wave count (148 persistent CTAs) does not get worse and the tile is long enough
        // for the epilogue of a 128 x 256 tile to hide (measured on the step's shapes, profiles/gemm_bench_r01*)
        if (a->a_lo != nullptr && a->n % 256 == 0 && a->split_k == 1) {
            const long long G = num_sms();
            const long long mt = ceil_div_i(a->m_per_batch, BM) * (long long)a->batches;
            const long long w128 = (mt * (a->n / 128) + G - 1) / G, w256 = (mt * (a->n / 256) + G - 1) / G;
            if (2 * w256 <= w128 && (kp.num_k_blocks >= 32 || w128 >= 4)) bn = 256;
        }
    }
    B2T_CHECK_ARG(bn == 64 || bn == 128 || bn == 256, "b2t_gemm: block_n must be 64, 128 or 256");
    kp.m_tiles = ceil_div_i(ceil_div_i(a->m_per_batch, BM), cg);   // row blocks of one scheduling unit (CTA or CTA pair)
    kp.n_tiles = ceil_div_i(a->n, bn);
    kp.total_items = kp.m_tiles * kp.n_tiles * kp.batches * a->split_k;

    // stream-K (needs a caller-owned workspace).  148 SMs rarely divide the tile count (M = 3776 x N = 4096 is 240 pair
    // tiles on 74 pairs = 3.24 waves): the tiles of the last, partial wave are cut along K into one equal range per unit
    // (partial tiles summed through the workspace by the owner, which runs the full fused epilogue), the other tiles run
    // whole.  stream_k = 1 forces it, 0 = when the partial wave would leave a worthwhile part of the device idle.
    kp.stream_k = 0;
    if (a->sk_workspace && a->stream_k >= 0) {
        const long long tiles = (long long)kp.m_tiles * kp.n_tiles * kp.batches;
        const int G = num_sms() / cg;
        const long long tail = tiles % G;
        // auto: the partial wave is under 3/4 full, its ranges are at least 2 k-blocks long
        const bool worth = tail > 0 && (tail * 4 <= 3ll * G || a->split_k > 1) && tail * kp.num_k_blocks >= 2ll * G;
        const bool want = a->stream_k > 0 ? tail > 0 : worth;
        if (want && kp.num_k_blocks >= 2) {
            B2T_CHECK_ARG(((uintptr_t)a->sk_workspace & 15) == 0, "b2t_gemm: sk_workspace must be 16-byte aligned");
            kp.stream_k = 1;
            kp.split_k = 1;
            kp.total_items = (int)tiles;
            kp.sk_tile0 = (int)(tiles - tail);
            kp.sk_units = tail * kp.num_k_blocks;
            kp.sk_G = kp.sk_units < G ? (int)kp.sk_units : G;
            kp.sk_flags = (int*)a->sk_workspace;
            kp.sk_ws = (float*)((char*)a->sk_workspace + SK_FLAG_BYTES);
        }
    }

    const bool three = a->a_lo != nullptr;
    if (!three) return b2t_gemm_part1(a, &kp, bn, cg, (void*)stream);
    if (cg == 2) return dispatch_major<3, 256, 2>(a, kp, stream);
    if (bn == 64) return dispatch_major<3, 64, 1>(a, kp, stream);
    if (bn == 128) return dispatch_major<3, 128, 1>(a, kp, stream);
    return dispatch_major<3, 256, 1>(a, kp, stream);
}
#else   // B2T_GEMM_PART == 1
extern "C" __attribute__((visibility("hidden"))) int b2t_gemm_part1(const b2t_gemm_args* a, const void* kparams, int bn, int cg, void* stream_) {
    const KParams& kp = *reinterpret_cast<const KParams*>(kparams);
    cudaStream_t stream = (cudaStream_t)stream_;
    if (cg == 2) return dispatch_major<1, 256, 2>(a, kp, stream);
    if (bn == 64) return dispatch_major<1, 64, 1>(a, kp, stream);
    if (bn == 128) return dispatch_major<1, 128, 1>(a, kp, stream);
    return dispatch_major<1, 256, 1>(a, kp, stream);
}
#endif
