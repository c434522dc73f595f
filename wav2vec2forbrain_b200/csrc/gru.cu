// Persistent recurrent GRU kernels (forward and backward), replacing the cuDNN nn.GRU call of
// src/model/brain_feature_extractor.py:39-47,61-65.  The input projections x.W_ih^T + b_ih of
// every layer run beforehand as tcgen05 GEMMs (b2t_gemm; layer 0 through the unfold view); this
// file holds only the strictly sequential part: T' dependent steps of h.W_hh^T + gates.
//
// One thread-block CLUSTER of CS = H/32 CTAs owns one (direction, group of 16 trials) for the
// whole sequence.  CTA `rank` owns hidden units [32*rank, 32*rank+32) and keeps its slice of
// W_hh resident for all T' steps: the bf16 "hi" halves live in REGISTERS as mma fragments, the
// "lo" halves in shared memory (bf16x3 split: hi.hi + hi.lo + lo.hi, fp32 accumulate, so the
// recurrence keeps ~fp32 accuracy over 118 dependent steps).  Per step:
//   forward : gh[16, 96] = h[16, H] . W_slice^T on mma.sync (M=16 trials), gate math for the 32
//             own units, the new h slice is pushed to every CTA of the cluster through
//             distributed shared memory, one barrier.cluster per step.
//   backward: gate derivatives for the own units, partial dh[16, H] = dgh[16, 96] . W_slice on
//             mma.sync, partial sums scattered to the owning CTAs through DSMEM (reduce-scatter),
//             one barrier.cluster per step.  dW_hh / dW_ih / bias gradients are big GEMMs over
//             the stored gate gradients afterwards (b2t_gemm), not per-step work.
// Clusters are independent (no grid-wide barrier), so any number of them may be resident.
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

constexpr int RB = 16;   // trials per cluster == mma M
constexpr int U = 32;    // hidden units per CTA
constexpr int NCOL = 3 * U;

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
    cluster_arrive();
    cluster_wait();
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint4 v) {
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
// remote store that signals `mbar` (an mbarrier of the DESTINATION CTA, cluster address) with its byte
// count: the arrival of the data is the synchronisation, no cluster-wide barrier / membar per step
__device__ __forceinline__ void st_async_v4(uint32_t addr, uint4 v, uint32_t mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(addr),
                 "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void st_async_v2f(uint32_t addr, float a, float b, uint32_t mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(addr), "f"(a),
                 "f"(b), "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
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
        if (spin > (1u << 22)) {   // a lost signal must trap, not hang the box
            printf("b2t_gru: mbarrier timeout block %d thread %d\n", blockIdx.x, threadIdx.x);
            __trap();
        }
    }
}
__device__ __forceinline__ void st_cluster_v2f(uint32_t addr, float a, float b) {
    asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void mma_bf16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ uint32_t pack_bf16(bf16 lo_half, bf16 hi_half) {
    __nv_bfloat162 p = __halves2bfloat162(lo_half, hi_half);
    return *reinterpret_cast<uint32_t*>(&p);
}
__device__ __forceinline__ void split_pair(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    bf16 h0, l0, h1, l1;
    split_bf16(x0, h0, l0);
    split_bf16(x1, h1, l1);
    hi = pack_bf16(h0, h1);
    lo = pack_bf16(l0, l1);
}
__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float keep_scale(float p, unsigned long long seed, unsigned long long off) {
    if (p <= 0.0f) return 1.0f;
    return hash_uniform(seed, off) < p ? 0.0f : 1.0f / (1.0f - p);
}

struct GruFwdArgs {
    const float* gi;      // [M, ndir*3H]  x.W_ih^T + b_ih, gate order r,z,n per direction
    const float* w_hh;    // [ndir][3H][H]
    const float* b_hh;    // [ndir][3H]
    const float* h0;      // [ndir][H] or null (zeros)
    float* out;           // [M, ndir*H]
    bf16* out_hi;         // [M, ndir*H]  (inter-layer dropout applied)
    bf16* out_lo;
    float* save_r;        // [ndir][M][H]   (null in inference)
    float* save_z;
    float* save_n;
    float* save_ghn;
    bf16* hprev_hi;       // [ndir][M][H]   h entering each step (operand of the dW_hh GEMM)
    bf16* hprev_lo;
    int B, T, ndir;
    float drop_p;
    unsigned long long seed;
};

// Warp tiling of the step's matvec gh[16, 96] = h[16, H] . W_slice^T over the 12 warps.  With one warp per 8 gate
// columns and all of K, every warp read the whole h (hi + lo, 32 KB at H = 512): 12 x 32 KB + 96 KB of W lo = 480 KB
// per step against the 128 B/clk shared-memory port = 1.9 us of a 4.8 us step (ncu: the matvec sat on stall_mio).
// KG = 4 splits K four ways instead: warp (gate, kg) owns the gate's 32 columns x a quarter of K, reads a quarter of h
// (96 KB per step in total) and leaves a partial sum per K group that the gate phase adds up.
template <int H>
struct FwdSmem {
    static constexpr int KSTEPS = H / 16;
    static constexpr int KG = KSTEPS % 4 == 0 ? 4 : 1;   // K groups
    static constexpr int NT_W = KG;                        // 8-column tiles per warp (4 x 8 = the gate's 32 columns)
    static constexpr int KS_W = KSTEPS / KG;               // k-steps per warp
    static constexpr int HP = H + 8;   // padded bf16 pitch: conflict-free A-fragment loads
    static constexpr size_t WLO = (size_t)12 * KSTEPS * 32 * 8;
    static constexpr size_t HBUF = (size_t)2 * 2 * RB * HP * 2;
    static constexpr size_t GH = (size_t)KG * RB * (NCOL + 1) * 4;
    static constexpr size_t HOWN = (size_t)RB * U * 4;
    static constexpr size_t STAGE = (size_t)2 * RB * U * 2;
    static constexpr size_t TOTAL = WLO + HBUF + GH + HOWN + STAGE;
};

template <int H>
__global__ void __launch_bounds__(384, 1) gru_fwd_kernel(const GruFwdArgs p) {
    using S = FwdSmem<H>;
    constexpr int KSTEPS = S::KSTEPS, HP = S::HP, CS = H / U;
    extern __shared__ __align__(16) uint8_t smem[];
    uint2* wlo = reinterpret_cast<uint2*>(smem);                       // [12][KSTEPS][32]
    bf16* hbuf = reinterpret_cast<bf16*>(smem + S::WLO);               // [2 buf][2 hi/lo][RB][HP]
    float* gh_s = reinterpret_cast<float*>(smem + S::WLO + S::HBUF);   // [RB][NCOL+1]
    float* hown = reinterpret_cast<float*>(smem + S::WLO + S::HBUF + S::GH);             // [RB][U]
    bf16* stage = reinterpret_cast<bf16*>(smem + S::WLO + S::HBUF + S::GH + S::HOWN);    // [2][RB][U]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tq = lane & 3;
    const uint32_t rank = cluster_ctarank();
    const int cluster_id = blockIdx.x / CS;
    const int dir = cluster_id % p.ndir, group = cluster_id / p.ndir;
    const int row0 = group * RB, u0 = (int)rank * U;
    const long M = (long)p.B * p.T;
    const int T = p.T;

    // ---- resident weights: this warp's gate columns x its K range --------------------------------
    constexpr int KG = S::KG, NT_W = S::NT_W, KS_W = S::KS_W;
    const int gate = warp >> 2, sub = warp & 3;
    const int kg = KG == 4 ? sub : 0;            // K group
    const int nt0 = KG == 4 ? 0 : sub;           // first 8-column tile of the gate's 32 columns
    uint32_t bhi[KS_W][NT_W][2];
    {
#pragma unroll
        for (int nt = 0; nt < NT_W; ++nt) {
            const int col = gate * H + u0 + (nt0 + nt) * 8 + g;   // row of W_hh
            const float* wr = p.w_hh + ((long)dir * 3 * H + col) * H + kg * KS_W * 16;
#pragma unroll
            for (int ks = 0; ks < KS_W; ++ks) {
                uint32_t l0, l1;
                split_pair(wr[ks * 16 + 2 * tq], wr[ks * 16 + 2 * tq + 1], bhi[ks][nt][0], l0);
                split_pair(wr[ks * 16 + 2 * tq + 8], wr[ks * 16 + 2 * tq + 9], bhi[ks][nt][1], l1);
                wlo[((warp * KS_W + ks) * NT_W + nt) * 32 + lane] = make_uint2(l0, l1);
            }
        }
    }
    float bias_c[NT_W][2];
#pragma unroll
    for (int nt = 0; nt < NT_W; ++nt) {
        const long bo = (long)dir * 3 * H + gate * H + u0 + (nt0 + nt) * 8 + 2 * tq;
        bias_c[nt][0] = kg == 0 ? p.b_hh[bo] : 0.0f;       // the bias is added once, by K group 0
        bias_c[nt][1] = kg == 0 ? p.b_hh[bo + 1] : 0.0f;
    }

    // ---- initial state ----------------------------------------------------------------------
    for (int i = tid; i < RB * H; i += blockDim.x) {
        const int r = i / H, k = i - r * H;
        const float v = p.h0 ? p.h0[(long)dir * H + k] : 0.0f;
        bf16 hh, ll;
        split_bf16(v, hh, ll);
        hbuf[(0 * 2 + 0) * RB * HP + r * HP + k] = hh;
        hbuf[(0 * 2 + 1) * RB * HP + r * HP + k] = ll;
    }
    for (int i = tid; i < RB * U; i += blockDim.x) hown[i] = p.h0 ? p.h0[(long)dir * H + u0 + (i % U)] : 0.0f;
    // hbar[b]: "hbuf[b] holds the complete h of the step about to start" — completed by the byte counts
    // of the CS remote slices (st.async), one phase per use
    __shared__ __align__(8) unsigned long long hbar[2];
    const uint32_t hbar0 = (uint32_t)__cvta_generic_to_shared(&hbar[0]);
    constexpr uint32_t STEP_BYTES = (uint32_t)CS * 2 * RB * U * 2;
    if (tid == 0) {
        mbar_init(hbar0, 1);
        mbar_init(hbar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    cluster_sync_all();
    uint32_t hphase[2] = {0u, 0u};

    // gate-phase items: threads 0..255 own unit (tid % 32) for rows tid/32 and tid/32 + 8
    const int gu = tid & 31, gr = tid >> 5;
    const int gdim = p.ndir * 3 * H;

    // input projections of a step, loaded one step ahead so their DRAM latency hides behind the
    // previous step's recurrence (the gate phase never waits on global memory)
    float gi_v[2][3];
    auto load_gi = [&](int s_) {
        const int t_ = dir == 0 ? s_ : T - 1 - s_;
        if (tid < 256) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int b = row0 + gr + 8 * j;
#pragma unroll
                for (int q = 0; q < 3; ++q)
                    gi_v[j][q] = (b < p.B && s_ < T) ? p.gi[((long)b * T + t_) * gdim + dir * 3 * H + q * H + u0 + gu] : 0.0f;
            }
        }
    };
    load_gi(0);

    // push of the new h slice.  Gate warp w (w < 8) computes rows w and w + 8 of the slice, so it also sends them: its 16
    // sixteen-byte pieces (hi / lo x 2 rows x 4) go to the CS CTAs of the cluster, lane -> piece lane & 15, destinations
    // (lane >> 4), + 2, + 4, ...  Staging is warp-private (a __syncwarp instead of a CTA barrier on the critical path),
    // and the cluster addresses are step-invariant (buffer 1 = buffer 0 + a constant): mapa runs once, here.
    constexpr int PER = CS / 2 > 0 ? CS / 2 : 1;
    const int pwhich = (lane & 15) >> 3, prow = warp + 8 * ((lane >> 2) & 1), pch = lane & 3;
    uint32_t rdst[PER], rbar[PER];
    if (warp < 8) {
        const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(hbuf + pwhich * RB * HP + prow * HP + u0 + pch * 8);
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int dest = (lane >> 4) + 2 * i;
            rdst[i] = map_to_cta(dst0, (uint32_t)(dest < CS ? dest : 0));
            rbar[i] = map_to_cta(hbar0, (uint32_t)(dest < CS ? dest : 0));
        }
    }

    for (int s = 0; s < T; ++s) {
        const int t = dir == 0 ? s : T - 1 - s;
        const int cur = s & 1;
        if (s > 0) {   // every slice of h_s has landed in hbuf[cur]
            mbar_wait(hbar0 + 8 * cur, hphase[cur]);
            hphase[cur] ^= 1u;
        }
        if (tid == 0 && s + 1 < T) mbar_expect_tx(hbar0 + 8 * (cur ^ 1), STEP_BYTES);
        // ---- gh = h . W_slice^T (bf16x3) ----
        {
            const bf16* hh = hbuf + (cur * 2 + 0) * RB * HP;
            const bf16* hl = hbuf + (cur * 2 + 1) * RB * HP;
            // lanes 0-15: rows 0-15 at k0, lanes 16-31: the same rows at k0 + 8 -> registers a0..a3 of m16n8k16
            const uint32_t hh_addr = (uint32_t)__cvta_generic_to_shared(hh + (lane & 15) * HP + (lane >> 4) * 8);
            const uint32_t hl_addr = (uint32_t)__cvta_generic_to_shared(hl + (lane & 15) * HP + (lane >> 4) * 8);
            float acc[NT_W][4];
#pragma unroll
            for (int nt = 0; nt < NT_W; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.0f;
#pragma unroll
            for (int ks = 0; ks < KS_W; ++ks) {
                // one ldmatrix.x4 per operand half (registers a0..a3 of m16n8k16)
                uint32_t ah[4], al[4];
                ldsm_x4(ah, hh_addr + (uint32_t)((kg * KS_W + ks) * 32));
                ldsm_x4(al, hl_addr + (uint32_t)((kg * KS_W + ks) * 32));
#pragma unroll
                for (int nt = 0; nt < NT_W; ++nt) {
                    const uint2 bl = wlo[((warp * KS_W + ks) * NT_W + nt) * 32 + lane];
                    mma_bf16_16816(acc[nt], ah, bhi[ks][nt][0], bhi[ks][nt][1]);
                    mma_bf16_16816(acc[nt], ah, bl.x, bl.y);
                    mma_bf16_16816(acc[nt], al, bhi[ks][nt][0], bhi[ks][nt][1]);
                }
            }
            float* ghp = gh_s + kg * RB * (NCOL + 1);
#pragma unroll
            for (int nt = 0; nt < NT_W; ++nt) {
                const int cidx = gate * U + (nt0 + nt) * 8 + 2 * tq;
                ghp[g * (NCOL + 1) + cidx] = acc[nt][0] + bias_c[nt][0];
                ghp[g * (NCOL + 1) + cidx + 1] = acc[nt][1] + bias_c[nt][1];
                ghp[(g + 8) * (NCOL + 1) + cidx] = acc[nt][2] + bias_c[nt][0];
                ghp[(g + 8) * (NCOL + 1) + cidx + 1] = acc[nt][3] + bias_c[nt][1];
            }
        }
        __syncthreads();
        // ---- gates for the own 32 units (brain_feature_extractor.py nn.GRU equations, App. A.2) ----
        float rv[2], zv[2], nv[2], ghnv[2], hpv[2], hnv[2];
        if (tid < 256) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int row = gr + 8 * j;
                float gh_r = 0.0f, gh_z = 0.0f, gh_n = 0.0f;
#pragma unroll
                for (int q = 0; q < KG; ++q) {   // partial sums of the K groups
                    const float* ghq = gh_s + (q * RB + row) * (NCOL + 1);
                    gh_r += ghq[gu];
                    gh_z += ghq[U + gu];
                    gh_n += ghq[2 * U + gu];
                }
                const float r = sigmoid_acc(gi_v[j][0] + gh_r);
                const float z = sigmoid_acc(gi_v[j][1] + gh_z);
                const float n = tanhf(gi_v[j][2] + r * gh_n);
                const float hp = hown[row * U + gu];
                const float hn = (1.0f - z) * n + z * hp;
                hown[row * U + gu] = hn;
                bf16 hh, ll;
                split_bf16(hn, hh, ll);
                stage[row * U + gu] = hh;
                stage[RB * U + row * U + gu] = ll;
                rv[j] = r, zv[j] = z, nv[j] = n, ghnv[j] = gh_n, hpv[j] = hp, hnv[j] = hn;
            }
        }
        // ---- push the new h slice (bf16 hi, lo) to every CTA of the cluster (not needed after the last step) ----
        // No CTA barrier: gh_s is rewritten by the next step's matvec only after hbar[nxt] completed, which needs this
        // CTA's own pushes, i.e. every gate warp is past its reads of gh_s.
        if (warp < 8 && s + 1 < T) {
            __syncwarp();
            const int nxt = cur ^ 1;
            const uint4 v = *reinterpret_cast<const uint4*>(stage + pwhich * RB * U + prow * U + pch * 8);
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                if ((lane >> 4) + 2 * i < CS)
                    st_async_v4(rdst[i] + (uint32_t)(nxt * 2 * RB * HP * 2), v, rbar[i] + (uint32_t)(8 * nxt));
            }
        }
        load_gi(s + 1);
        // ---- outputs and saved-for-backward values (off the recurrence's critical path) ----
        if (tid < 256) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int b = row0 + gr + 8 * j;
                if (b < p.B) {
                    const long m = (long)b * T + t;
                    const long oo = m * (p.ndir * H) + dir * H + u0 + gu;
                    p.out[oo] = hnv[j];
                    if (p.out_hi) {
                        const float hd = hnv[j] * keep_scale(p.drop_p, p.seed, (unsigned long long)oo);
                        bf16 dh_, dl_;
                        split_bf16(hd, dh_, dl_);
                        p.out_hi[oo] = dh_;
                        if (p.out_lo) p.out_lo[oo] = dl_;
                    }
                    if (p.save_r) {
                        const long so = ((long)dir * M + m) * H + u0 + gu;
                        p.save_r[so] = rv[j];
                        p.save_z[so] = zv[j];
                        p.save_n[so] = nv[j];
                        p.save_ghn[so] = ghnv[j];
                        bf16 ph, pl;
                        split_bf16(hpv[j], ph, pl);
                        p.hprev_hi[so] = ph;
                        if (p.hprev_lo) p.hprev_lo[so] = pl;
                    }
                }
            }
        }
    }
    cluster_sync_all();   // no CTA may exit while a peer can still write into its shared memory
}

struct GruBwdArgs {
    const float* dout;     // [M, ndir*H]  gradient w.r.t. the layer output (pre-dropout form is derived here)
    const float* out;      // [M, ndir*H]  forward output (h_t), gives h_{t-1}
    const float* w_hh;     // [ndir][3H][H]
    const float* h0;       // [ndir][H] or null
    const float* save_r;   // [ndir][M][H]
    const float* save_z;
    const float* save_n;
    const float* save_ghn;
    float* dgi;            // [M, ndir*3H]  gradient w.r.t. the input projections
    bf16* dgi_hi;          // [M, ndir*3H]
    bf16* dgi_lo;
    bf16* dgh_hi;          // [ndir][M][3H]  gradient w.r.t. h.W_hh^T + b_hh (operand of the dW_hh GEMM)
    bf16* dgh_lo;
    float* dghn;           // [ndir][M][H]   its n-gate part in fp32 (b_hh gradient)
    float* dh0;            // [ndir][H] += sum over trials (learnable initial state) or null
    int B, T, ndir;
    float drop_p;
    unsigned long long seed;
};

template <int H>
struct BwdSmem {
    static constexpr int CS = H / U;
    static constexpr int NWARP_MMA = CS;                 // warp w produces the partials of CTA w's units
    static constexpr int THREADS = CS * 32 < 256 ? 256 : CS * 32;
    static constexpr int AP = NCOL + 8;                  // padded bf16 pitch of the dgh operand
    static constexpr size_t WLO = (size_t)NWARP_MMA * 24 * 32 * 8;
    static constexpr size_t RECV = (size_t)2 * CS * RB * U * 4;
    static constexpr size_t AOP = (size_t)2 * RB * AP * 2;
    static constexpr size_t DHZ = (size_t)RB * U * 4;
    static constexpr size_t TOTAL = WLO + RECV + AOP + DHZ;
};

template <int H>
__global__ void __launch_bounds__(BwdSmem<H>::THREADS, 1) gru_bwd_kernel(const GruBwdArgs p) {
    using S = BwdSmem<H>;
    constexpr int CS = S::CS, AP = S::AP;
    extern __shared__ __align__(16) uint8_t smem[];
    uint2* wlo = reinterpret_cast<uint2*>(smem);                                   // [CS warps][4 nt][6 ks][32]
    float* recv = reinterpret_cast<float*>(smem + S::WLO);                          // [2][CS src][RB][U]
    bf16* aop = reinterpret_cast<bf16*>(smem + S::WLO + S::RECV);                   // [2 hi/lo][RB][AP]
    float* dhz = reinterpret_cast<float*>(smem + S::WLO + S::RECV + S::AOP);        // [RB][U]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, tq = lane & 3;
    const uint32_t rank = cluster_ctarank();
    const int cluster_id = blockIdx.x / CS;
    const int dir = cluster_id % p.ndir, group = cluster_id / p.ndir;
    const int row0 = group * RB, u0 = (int)rank * U;
    const long M = (long)p.B * p.T;
    const int T = p.T;

    // ---- resident weights: B[k = own gate column kk, n = unit of destination CTA `warp`] = W_hh[col(kk)][unit]
    uint32_t bhi[4][6][2];
    if (warp < CS) {
        const float* wd = p.w_hh + (long)dir * 3 * H * H;
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
            const int unit = warp * U + nt * 8 + g;
#pragma unroll
            for (int ks = 0; ks < 6; ++ks) {
                float w[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int kk = ks * 16 + 2 * tq + (e & 1) + (e >> 1) * 8;
                    const int col = (kk / U) * H + u0 + (kk % U);
                    w[e] = wd[(long)col * H + unit];
                }
                uint32_t l0, l1;
                split_pair(w[0], w[1], bhi[nt][ks][0], l0);
                split_pair(w[2], w[3], bhi[nt][ks][1], l1);
                wlo[((warp * 4 + nt) * 6 + ks) * 32 + lane] = make_uint2(l0, l1);
            }
        }
    }
    for (int i = tid; i < RB * U; i += blockDim.x) dhz[i] = 0.0f;
    for (int i = tid; i < 2 * CS * RB * U; i += blockDim.x) recv[i] = 0.0f;
    // rbar[b]: "recv[b] holds every peer's partial dh for the step about to start" (st.async byte counts)
    __shared__ __align__(8) unsigned long long rbar[2];
    const uint32_t rbar0 = (uint32_t)__cvta_generic_to_shared(&rbar[0]);
    constexpr uint32_t STEP_BYTES = (uint32_t)CS * RB * U * 4;
    if (tid == 0) {
        mbar_init(rbar0, 1);
        mbar_init(rbar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    cluster_sync_all();
    uint32_t rphase[2] = {0u, 0u};

    const int odim = p.ndir * H, gdim = p.ndir * 3 * H;
    // One (trial row, unit) item per thread and pass; its saved forward values are loaded one step
    // ahead (DRAM latency hidden behind the previous step), its global stores are issued after the
    // cluster barrier's arrive (the release does not have to drain them).
    constexpr int ITEMS = (RB * U + S::THREADS - 1) / S::THREADS;
    float f_dout[ITEMS], f_r[ITEMS], f_z[ITEMS], f_n[ITEMS], f_ghn[ITEMS], f_hp[ITEMS];
    auto prefetch = [&](int s_) {
        const int t_ = dir == 0 ? T - 1 - s_ : s_;
        const int tp_ = dir == 0 ? t_ - 1 : t_ + 1;
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int item = tid + it * S::THREADS;
            const int row = item / U, uu = item % U;
            const int b = row0 + row;
            f_dout[it] = f_r[it] = f_z[it] = f_n[it] = f_ghn[it] = f_hp[it] = 0.0f;
            if (item < RB * U && b < p.B && s_ < T) {
                const long m = (long)b * T + t_;
                const long oo = m * odim + dir * H + u0 + uu;
                const long so = ((long)dir * M + m) * H + u0 + uu;
                f_dout[it] = p.dout[oo] * keep_scale(p.drop_p, p.seed, (unsigned long long)oo);
                f_r[it] = p.save_r[so];
                f_z[it] = p.save_z[so];
                f_n[it] = p.save_n[so];
                f_ghn[it] = p.save_ghn[so];
                if (tp_ >= 0 && tp_ < T) f_hp[it] = p.out[((long)b * T + tp_) * odim + dir * H + u0 + uu];
                else f_hp[it] = p.h0 ? p.h0[(long)dir * H + u0 + uu] : 0.0f;
            }
        }
    };
    prefetch(0);
    for (int s = 0; s < T; ++s) {
        const int t = dir == 0 ? T - 1 - s : s;           // reverse of the forward order
        const int cur = s & 1;
        if (s > 0) {   // every peer's partial sums for this step have landed in recv[cur]
            mbar_wait(rbar0 + 8 * cur, rphase[cur]);
            rphase[cur] ^= 1u;
        }
        if (tid == 0) mbar_expect_tx(rbar0 + 8 * (cur ^ 1), STEP_BYTES);
        // ---- gate derivatives for the own units ----
        float o_ar[ITEMS], o_az[ITEMS], o_an[ITEMS], o_ghn[ITEMS];
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int item = tid + it * S::THREADS;
            if (item < RB * U) {
                const int row = item / U, uu = item % U;
                const int b = row0 + row;
                float dar = 0.f, daz = 0.f, dan = 0.f, dghn_v = 0.f, carry = 0.f;
                if (b < p.B) {
                    float rec = dhz[item];
#pragma unroll 4
                    for (int src = 0; src < CS; ++src) rec += recv[((cur * CS + src) * RB + row) * U + uu];
                    const float dh = rec + f_dout[it];
                    const float r = f_r[it], z = f_z[it], n = f_n[it], ghn = f_ghn[it], hp = f_hp[it];
                    const float dn = dh * (1.0f - z);
                    const float dz = dh * (hp - n);
                    carry = dh * z;
                    dan = dn * (1.0f - n * n);
                    dar = dan * ghn * r * (1.0f - r);
                    daz = dz * z * (1.0f - z);
                    dghn_v = dan * r;
                }
                o_ar[it] = dar, o_az[it] = daz, o_an[it] = dan, o_ghn[it] = dghn_v;
                dhz[item] = carry;
                bf16 hh, ll;
                split_bf16(dar, hh, ll);
                aop[row * AP + uu] = hh;
                aop[RB * AP + row * AP + uu] = ll;
                split_bf16(daz, hh, ll);
                aop[row * AP + U + uu] = hh;
                aop[RB * AP + row * AP + U + uu] = ll;
                split_bf16(dghn_v, hh, ll);
                aop[row * AP + 2 * U + uu] = hh;
                aop[RB * AP + row * AP + 2 * U + uu] = ll;
            }
        }
        prefetch(s + 1);
        __syncthreads();
        // ---- partial dh[16, units of CTA `warp`] = dgh[16, 96 own columns] . W_slice, scattered to its owner ----
        if (warp < CS) {
            const uint32_t ah_addr = (uint32_t)__cvta_generic_to_shared(aop + (lane & 15) * AP + (lane >> 4) * 8);
            const uint32_t al_addr = (uint32_t)__cvta_generic_to_shared(aop + RB * AP + (lane & 15) * AP + (lane >> 4) * 8);
            float acc[4][4];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[nt][e] = 0.0f;
#pragma unroll
            for (int ks = 0; ks < 6; ++ks) {
                uint32_t ah[4], al[4];
                ldsm_x4(ah, ah_addr + (uint32_t)(ks * 32));
                ldsm_x4(al, al_addr + (uint32_t)(ks * 32));
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    const uint2 bl = wlo[((warp * 4 + nt) * 6 + ks) * 32 + lane];
                    mma_bf16_16816(acc[nt], ah, bhi[nt][ks][0], bhi[nt][ks][1]);
                    mma_bf16_16816(acc[nt], ah, bl.x, bl.y);
                    mma_bf16_16816(acc[nt], al, bhi[nt][ks][0], bhi[nt][ks][1]);
                }
            }
            const int nxt = cur ^ 1;
            float* slot = recv + ((nxt * CS + (int)rank) * RB) * U;   // [RB][U] slot of source `rank` in the owner
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                const uint32_t a0 = map_to_cta((uint32_t)__cvta_generic_to_shared(slot + g * U + nt * 8 + 2 * tq), (uint32_t)warp);
                const uint32_t a1 =
                    map_to_cta((uint32_t)__cvta_generic_to_shared(slot + (g + 8) * U + nt * 8 + 2 * tq), (uint32_t)warp);
                const uint32_t mb = map_to_cta(rbar0 + 8 * nxt, (uint32_t)warp);
                st_async_v2f(a0, acc[nt][0], acc[nt][1], mb);
                st_async_v2f(a1, acc[nt][2], acc[nt][3], mb);
            }
        }
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int item = tid + it * S::THREADS;
            const int row = item / U, uu = item % U;
            const int b = row0 + row;
            if (item < RB * U && b < p.B) {
                const long m = (long)b * T + t;
                const long so = ((long)dir * M + m) * H + u0 + uu;
                const long go = m * gdim + dir * 3 * H + u0 + uu;
                const long ho = ((long)dir * M + m) * 3 * H + u0 + uu;
                const float dar = o_ar[it], daz = o_az[it], dan = o_an[it], dghn_v = o_ghn[it];
                p.dgi[go] = dar;
                p.dgi[go + H] = daz;
                p.dgi[go + 2 * H] = dan;
                bf16 hh, ll;
                split_bf16(dar, hh, ll);
                p.dgi_hi[go] = hh;
                p.dgh_hi[ho] = hh;
                if (p.dgi_lo) { p.dgi_lo[go] = ll; p.dgh_lo[ho] = ll; }
                split_bf16(daz, hh, ll);
                p.dgi_hi[go + H] = hh;
                p.dgh_hi[ho + H] = hh;
                if (p.dgi_lo) { p.dgi_lo[go + H] = ll; p.dgh_lo[ho + H] = ll; }
                split_bf16(dan, hh, ll);
                p.dgi_hi[go + 2 * H] = hh;
                if (p.dgi_lo) p.dgi_lo[go + 2 * H] = ll;
                split_bf16(dghn_v, hh, ll);
                p.dgh_hi[ho + 2 * H] = hh;
                if (p.dgh_lo) p.dgh_lo[ho + 2 * H] = ll;
                p.dghn[so] = dghn_v;
            }
        }
    }
    {   // the partial sums of the last step (they only feed dh0) must land before anyone exits
        const int cur = T & 1;
        mbar_wait(rbar0 + 8 * cur, rphase[cur]);
    }
    // gradient w.r.t. the initial state (only consumed when hidden_start is learnable)
    if (p.dh0) {
        const int cur = T & 1;
        for (int item = tid; item < RB * U; item += blockDim.x) {
            const int row = item / U, uu = item % U;
            if (row0 + row < p.B) {
                float rec = dhz[item];
                for (int src = 0; src < CS; ++src) rec += recv[((cur * CS + src) * RB + row) * U + uu];
                atomicAdd(p.dh0 + (long)dir * H + u0 + uu, rec);
            }
        }
    }
    cluster_sync_all();   // no CTA may exit while a peer can still write into its shared memory
}

template <int H>
int launch_fwd(const GruFwdArgs& a, cudaStream_t stream) {
    constexpr int CS = H / U;
    const int groups = (a.B + RB - 1) / RB;
    auto kern = gru_fwd_kernel<H>;
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FwdSmem<H>::TOTAL));
        if (CS > 8) B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(CS * groups * a.ndir);
    cfg.blockDim = dim3(384);
    cfg.dynamicSmemBytes = FwdSmem<H>::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    B2T_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
    return 0;
}

template <int H>
int launch_bwd(const GruBwdArgs& a, cudaStream_t stream) {
    constexpr int CS = H / U;
    const int groups = (a.B + RB - 1) / RB;
    auto kern = gru_bwd_kernel<H>;
    static bool attr = false;
    if (!attr) {
        B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BwdSmem<H>::TOTAL));
        if (CS > 8) B2T_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(CS * groups * a.ndir);
    cfg.blockDim = dim3(BwdSmem<H>::THREADS);
    cfg.dynamicSmemBytes = BwdSmem<H>::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    B2T_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
    return 0;
}

}  // namespace

/* diagnostic: how many forward clusters (H/32 CTAs each) the device can keep resident at once */
extern "C" int b2t_gru_max_active_clusters(int H) {
    if (H != 512) return -1;
    constexpr int CS = 512 / U;
    auto kern = gru_fwd_kernel<512>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FwdSmem<512>::TOTAL);
    cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(CS * 16);
    cfg.blockDim = dim3(384);
    cfg.dynamicSmemBytes = FwdSmem<512>::TOTAL;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = -1;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) return -2;
    return n;
}

extern "C" int b2t_gru_supported_hidden(int H) { return (H == 32 || H == 64 || H == 128 || H == 256 || H == 512) ? 1 : 0; }

extern "C" int b2t_gru_fwd(const float* gi, const float* w_hh, const float* b_hh, const float* h0, float* out, void* out_hi,
                           void* out_lo, float* save_r, float* save_z, float* save_n, float* save_ghn, void* hprev_hi,
                           void* hprev_lo, int B, int T, int H, int ndir, float drop_p, uint64_t seed, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(gi && w_hh && b_hh && out, "b2t_gru_fwd: null pointer");
    B2T_CHECK_ARG(B > 0 && T > 0 && (ndir == 1 || ndir == 2), "b2t_gru_fwd: bad sizes");
    B2T_CHECK_ARG(b2t_gru_supported_hidden(H), "b2t_gru_fwd: hidden size %d unsupported (32/64/128/256/512)", H);
    B2T_CHECK_ARG(!save_r || (save_z && save_n && save_ghn && hprev_hi), "b2t_gru_fwd: incomplete save set");
    B2T_CHECK_ARG(drop_p >= 0.0f && drop_p < 1.0f, "b2t_gru_fwd: drop_p");
    GruFwdArgs a;
    a.gi = gi; a.w_hh = w_hh; a.b_hh = b_hh; a.h0 = h0; a.out = out;
    a.out_hi = (bf16*)out_hi; a.out_lo = (bf16*)out_lo;
    a.save_r = save_r; a.save_z = save_z; a.save_n = save_n; a.save_ghn = save_ghn;
    a.hprev_hi = (bf16*)hprev_hi; a.hprev_lo = (bf16*)hprev_lo;
    a.B = B; a.T = T; a.ndir = ndir; a.drop_p = drop_p; a.seed = seed;
    switch (H) {
        case 32: return launch_fwd<32>(a, stream);
        case 64: return launch_fwd<64>(a, stream);
        case 128: return launch_fwd<128>(a, stream);
        case 256: return launch_fwd<256>(a, stream);
        default: return launch_fwd<512>(a, stream);
    }
}

extern "C" int b2t_gru_bwd(const float* dout, const float* out, const float* w_hh, const float* h0, const float* save_r,
                           const float* save_z, const float* save_n, const float* save_ghn, float* dgi, void* dgi_hi,
                           void* dgi_lo, void* dgh_hi, void* dgh_lo, float* dghn, float* dh0, int B, int T, int H, int ndir,
                           float drop_p, uint64_t seed, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(dout && out && w_hh && save_r && save_z && save_n && save_ghn && dgi && dgi_hi && dgh_hi && dghn,
                  "b2t_gru_bwd: null pointer");
    B2T_CHECK_ARG((dgi_lo == nullptr) == (dgh_lo == nullptr), "b2t_gru_bwd: lo outputs must both be set or both NULL");
    B2T_CHECK_ARG(B > 0 && T > 0 && (ndir == 1 || ndir == 2), "b2t_gru_bwd: bad sizes");
    B2T_CHECK_ARG(b2t_gru_supported_hidden(H), "b2t_gru_bwd: hidden size %d unsupported (32/64/128/256/512)", H);
    GruBwdArgs a;
    a.dout = dout; a.out = out; a.w_hh = w_hh; a.h0 = h0;
    a.save_r = save_r; a.save_z = save_z; a.save_n = save_n; a.save_ghn = save_ghn;
    a.dgi = dgi; a.dgi_hi = (bf16*)dgi_hi; a.dgi_lo = (bf16*)dgi_lo;
    a.dgh_hi = (bf16*)dgh_hi; a.dgh_lo = (bf16*)dgh_lo; a.dghn = dghn; a.dh0 = dh0;
    a.B = B; a.T = T; a.ndir = ndir; a.drop_p = drop_p; a.seed = seed;
    switch (H) {
        case 32: return launch_bwd<32>(a, stream);
        case 64: return launch_bwd<64>(a, stream);
        case 128: return launch_bwd<128>(a, stream);
        case 256: return launch_bwd<256>(a, stream);
        default: return launch_bwd<512>(a, stream);
    }
}
