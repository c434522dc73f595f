/* libb2t_sm100.so — C ABI of the B200 (sm_100a) kernels behind the b2p2t_gru+w2v hot path.
 *
 * Drop-in boundary: the reference has no native code; the path it runs is
 * PyTorch library calls inside  W2VBrainEncoderModel.forward
 * (src/model/w2v_custom_feat_extractor.py:65-123) and its callees.  Every entry point below
 * replaces one of those library calls; the reference file:line it replaces is cited per entry.
 *
 * Conventions
 *   - plain pointers + sizes, no torch types.  All pointers are DEVICE pointers unless noted.
 *   - buffers (inputs, outputs, saved-for-backward, workspaces) are owned by the caller.
 *   - every launch goes to `stream` and returns immediately (no host sync, no hidden stream).
 *   - return 0 = ok; <0 = argument error, nothing launched; >0 = cudaError_t of the launch.
 *     b2t_last_error() returns a thread-local message.
 *   - fp32 tensors are `float*`; bf16 tensors are `void*` (uint16 storage).
 *   - "hi/lo" pairs are the bf16x3 operand form:  x ~= hi + lo  (lo may be NULL = plain bf16).
 */
#ifndef B2T_SM100_H
#define B2T_SM100_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef void* b2t_stream;   /* cudaStream_t */

int         b2t_abi_version(void);
const char* b2t_last_error(void);
/* 0 when device `dev` is compute capability 10.x (B200); negative otherwise. */
int         b2t_check_device(int dev);

/* ---------------------------------------------------------------------------------------
 * tcgen05 / TMA GEMM   D = epilogue( alpha * sum_terms A_t . B_t^T )
 * replaces every nn.Linear / einsum / conv-as-GEMM on the path: b2p2t_model.py:155-158 (day
 * affine), brain_feature_extractor.py:61-67 (GRU input projections via cuDNN, FC stack),
 * transformers modeling_wav2vec2.py:326-368,466-573 (pos-conv, q/k/v/out projections, FFN),
 * w2v_custom_feat_extractor.py:149-152 (lm_head) and their autograd backward GEMMs.
 * ------------------------------------------------------------------------------------- */
typedef struct {
    /* operands, bf16.  *_lo == NULL -> 1-term (plain bf16) GEMM, else 3-term bf16x3. */
    const void* a_hi; const void* a_lo;
    const void* b_hi; const void* b_lo;
    /* each operand is described as a 3-D strided view, d0 contiguous; strides in ELEMENTS.
     * K-major  (mn_major=0): d0 = K, d1 = rows (M or N) inside a batch, d2 = batch
     * MN-major (mn_major=1): d0 = M or N, d1 = k rows, d2 = batch (or k-batch, see below) */
    int64_t a_d0, a_d1, a_d2, a_s1, a_s2;
    int64_t b_d0, b_d1, b_d2, b_s1, b_s2;
    int32_t a_mn_major, b_mn_major;
    /* MN-major operands only: contraction runs jointly over (d2, d1) of BOTH operands, every
     * d2 slice padded to a multiple of 64 rows with zeros (implicit-unfold weight gradients). */
    int32_t k_spans_batches;
    int32_t m_per_batch;      /* rows of one A batch (tile rows never cross a batch) */
    int32_t batches;          /* number of output batches (1 for a plain GEMM) */
    int32_t n;                /* output columns */
    int32_t k;                /* contraction length (per d2 slice when k_spans_batches) */
    int32_t a_batched, b_batched;   /* operand d2 index follows the output batch */
    const int32_t* b_batch_idx;     /* optional [batches]: B d2 index per batch (day affine) */
    const int32_t* c_batch_idx;     /* optional [batches]: output batch slot per batch */
    /* outputs, all optional, all addressed  slot*c_batch_stride + row*ldc + col */
    float* out_f32;           /* final value */
    float* out_pre;           /* value before the activation (saved for backward) */
    void*  out_hi; void* out_lo;   /* final value as bf16 hi/lo */
    int64_t ldc, c_batch_stride;
    const float* bias;        /* [n] added before the activation */
    const float* residual;    /* same layout as out, added after the activation */
    const float* aux;         /* same layout as out; value is multiplied by act'(aux) */
    float* colsum;            /* [n] += column sums of the final value (bias gradients) */
    int32_t act;              /* 0 none, 1 gelu(erf), 2 silu, 3 softsign */
    int32_t aux_mode;         /* 0 none, 1 gelu', 2 silu', 3 softsign', 4 value *= aux (see pre_mode) */
    int32_t atomic;           /* out_f32 += value with red.add (split-K, overlapped folds) */
    int32_t split_k;          /* >=1; >1 requires atomic, act==0, no hi/lo/pre */
    float   alpha;
    int32_t block_n;          /* 0 = auto, else 64 / 128 / 256 */
    /* inverted dropout applied after act / act'(aux), before the residual: element `off` is kept
     * iff hash(drop_seed, off) >= drop_p (off = output offset), so the backward GEMM that passes
     * the same seed over the same layout regenerates the mask instead of storing it */
    float    drop_p;
    uint64_t drop_seed;
    /* bias row of batch b = bias + slot(b) * bias_batch_stride, slot = b_batch_idx[b] or b (0: shared) */
    int64_t  bias_batch_stride;
    /* value is multiplied by out_scale after act / act'(aux) / dropout and before the residual
     * (the conformer's half-step feed-forward, modeling_wav2vec2_conformer.py:602,625); 0 reads as 1 */
    float    out_scale;
    /* stream-K schedule.  sk_workspace: caller-owned device scratch of b2t_gemm_workspace_bytes() bytes,
     * ZEROED ONCE at allocation (the kernel leaves its flags at zero), used by one stream at a time;
     * NULL = whole-tile schedule.  stream_k: 0 auto (when the tile count leaves SMs idle in the last wave, or
     * split_k > 1 was asked for: the k-blocks of all tiles are cut into one equal range per SM, partial tiles
     * are summed through the workspace and the owner CTA runs the full fused epilogue, so no red.add and no
     * epilogue restriction), 1 force, -1 off. */
    void*    sk_workspace;
    int64_t  sk_workspace_bytes;
    int32_t  stream_k;
    /* 0 = auto, 1 = single-CTA tiles (128 x block_n), 2 = CTA pair (tcgen05 cta_group::2, 256 x 256 tiles: each CTA
     * stages its own 128 rows of A and half of the B tile, UMMA M = 256) */
    int32_t  cta_group;
    /* 0: out_pre receives the value before the activation (backward multiplies its gradient by act'(aux), aux_mode 1-3,
     *    and regenerates the dropout mask from drop_seed).
     * 1: out_pre receives the whole backward factor of the fused elementwise tail, d(value)/d(pre-activation) =
     *    act'(x) * (dropped ? 0 : 1/(1-drop_p)) * out_scale, computed where act(x) and the keep/drop decision exist
     *    anyway; the backward GEMM then passes it as aux with aux_mode 4 (plain multiply) and no drop_p. */
    int32_t  pre_mode;
    int32_t  reserved_;
} b2t_gemm_args;

int b2t_gemm(const b2t_gemm_args* args, b2t_stream stream);
/* cap the persistent grid at `sms` CTAs (0 = all SMs): leaves SMs to a concurrent collective (also env B2T_GEMM_SMS) */
int b2t_gemm_set_sm_limit(int sms);
/* 1: work items are handed out by the hardware's cluster launch control (one cluster launched per item; resident
 * clusters cancel and take over the items of clusters that are not resident yet) instead of a fixed persistent grid with
 * strided item lists - for GEMMs that share the device with long-running collective kernels.  0 (default): fixed grid.
 * Also env B2T_GEMM_CLC. */
int b2t_gemm_set_dynamic_schedule(int on);
/* bytes of b2t_gemm_args.sk_workspace that cover every tile shape on the current device */
int64_t b2t_gemm_workspace_bytes(void);

/* plain fp32 SIMT GEMM  C[m,n] = sum_k A(m,k) B(n,k) (+bias[n]); element strides; a debugging
 * cross-check for b2t_gemm, never used by the product path. */
int b2t_gemm_simt(const float* a, const float* b, float* c, const float* bias, int m, int n, int k,
                  int64_t sam, int64_t sak, int64_t sbn, int64_t sbk, int64_t ldc, b2t_stream stream);

/* fp32 -> bf16 hi/lo (lo optional).  [rows, cols] row-major in; out row-major, or transposed
 * [cols, rows] when transpose != 0.  `ld_in` = input row stride in elements. */
int b2t_split_bf16(const float* x, void* hi, void* lo, int64_t rows, int64_t cols, int64_t ld_in,
                   int transpose, b2t_stream stream);
/* GRU layer-0 weight re-layout: w[rows, C*KL] with column c*KL+k (nn.Unfold order,
 * b2p2t_model.py:162-167) -> hi/lo [rows, KL*C] with column k*C+c (memory order of the
 * un-materialised unfold view).  inverse=1 maps a gradient in k*C+c order back (fp32 -> fp32). */
int b2t_permute_unfold_cols(const float* w, void* hi, void* lo, float* out_f32, int64_t rows, int C,
                            int KL, int inverse, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * Stage 1: augmentation + Gaussian smoothing   (src/train/train_loop.py:46-62 intended form,
 * src/model/b2p2t_model.py:41-90,150-152 GaussianSmoothing / F.conv1d groups=C, padding="same")
 * x [B,T,C] fp32, taps [C,ntaps]; y = sum_k taps[c][k] * xa[b, t+k-pad_left, c], zeros outside.
 * white_sd / offset_sd = 0 is the parity path.  Outputs: y fp32 and/or hi/lo.
 * ------------------------------------------------------------------------------------- */
int b2t_smooth_fwd(const float* x, const float* taps, float* y, void* y_hi, void* y_lo, int B, int T, int C,
                   int ntaps, int pad_left, float white_sd, float offset_sd, uint64_t seed, b2t_stream stream);
/* backward of softsign (b2p2t_model.py:159): dz = ds/(1+|z|)^2 as bf16 hi/lo, and
 * dbias[day[b]][c] += sum_t dz[b,t,c]  (day_bias gradient, b2p2t_model.py:158) */
int b2t_softsign_bwd(const float* ds, const float* z, const int32_t* day, void* dz_hi, void* dz_lo, float* dbias,
                     int B, int T, int C, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * LayerNorm (transformers modeling_wav2vec2.py:690-693,592-609) and small HBM-bound helpers.
 * fwd: y = LN(x)*gamma+beta [then inverted dropout(drop_p, seed)], saves mean / rstd.
 * bwd: dy may carry that output dropout; writes dx (fp32), the branch copy dx*mask(br_drop) as
 *      bf16 hi/lo (operand of the upstream GEMMs) and accumulates dgamma, dbeta, dcol (+=).
 *      dres (optional, same shape) is added to dx first (fp32 and branch copies): the gradient
 *      that reaches x over the residual connection around a pre-LN branch.
 * ------------------------------------------------------------------------------------- */
int b2t_layernorm_fwd(const float* x, const float* gamma, const float* beta, float eps, float* y, void* y_hi,
                      void* y_lo, float* mean, float* rstd, int rows, int n, float drop_p, uint64_t seed,
                      b2t_stream stream);
int b2t_layernorm_bwd(const float* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                      float* dx, void* dx_hi, void* dx_lo, float* dgamma, float* dbeta, float* dcol, int rows,
                      int n, float out_drop_p, uint64_t out_seed, float br_drop_p, uint64_t br_seed,
                      const float* dres, b2t_stream stream);
/* y = dropout(x) with the library's counter-based mask; outputs fp32 and/or bf16 hi/lo (p=0: split) */
int b2t_dropout(const float* x, float* y, void* hi, void* lo, int64_t n, float p, uint64_t seed, b2t_stream stream);
/* out[c] += sum_r x[r*ld + c] */
int b2t_colsum(const float* x, float* out, int64_t rows, int n, int64_t ld, b2t_stream stream);
/* out = alpha*a + beta*b (b may be NULL) */
int b2t_axpby(const float* a, const float* b, float* out, int64_t n, float alpha, float beta, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * Fused self-attention (transformers modeling_wav2vec2.py:466-549 eager path; conformer
 * modeling_wav2vec2_conformer.py:420-527 after RoPE).  qkv is one [B*T, ld] buffer holding q, k, v
 * at column offsets q_off/k_off/v_off (+ head*d).  probs [B,H,T,T] is saved for backward (may be
 * NULL in inference).  kv_lens (optional) masks keys >= kv_lens[b]; the reference passes no mask.
 * ------------------------------------------------------------------------------------- */
int b2t_attention_fwd(const float* qkv, int64_t ld, int q_off, int k_off, int v_off, const int32_t* kv_lens,
                      float* probs, float* ctx, void* ctx_hi, void* ctx_lo, int64_t ld_ctx, int B, int T, int H,
                      int d, float scale, float drop_p, uint64_t seed, b2t_stream stream);
int b2t_attention_bwd(const float* qkv, int64_t ld, int q_off, int k_off, int v_off, const float* probs,
                      const float* dctx, int64_t ld_ctx, float* ds_ws, float* dqkv, void* dqkv_hi, void* dqkv_lo,
                      int B, int T, int H, int d, float scale, float drop_p, uint64_t seed, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * Persistent GRU recurrence (src/model/brain_feature_extractor.py:39-47,61-65 nn.GRU).
 * gi = x.W_ih^T + b_ih for all frames and both directions, [B*T, ndir*3H] (gate order r,z,n).
 * Hidden sizes 32/64/128/256/512 (a cluster of H/32 CTAs keeps W_hh on chip for the sequence).
 * ------------------------------------------------------------------------------------- */
int b2t_gru_supported_hidden(int H);
/* diagnostic: resident forward clusters (H/32 CTAs each) the current device can hold; < 0 = unknown */
int b2t_gru_max_active_clusters(int H);
int b2t_gru_fwd(const float* gi, const float* w_hh, const float* b_hh, const float* h0, float* out, void* out_hi,
                void* out_lo, float* save_r, float* save_z, float* save_n, float* save_ghn, void* hprev_hi,
                void* hprev_lo, int B, int T, int H, int ndir, float drop_p, uint64_t seed, b2t_stream stream);
int b2t_gru_bwd(const float* dout, const float* out, const float* w_hh, const float* h0, const float* save_r,
                const float* save_z, const float* save_n, const float* save_ghn, float* dgi, void* dgi_hi,
                void* dgi_lo, void* dgh_hi, void* dgh_lo, float* dghn, float* dh0, int B, int T, int H, int ndir,
                float drop_p, uint64_t seed, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * CTC (src/model/w2v_custom_feat_extractor.py:59,81-90: log_softmax + nn.CTCLoss(blank=0,
 * reduction="mean", zero_infinity=True)) and greedy decode (src/train/evaluator.py:108-120).
 * logits[b][t][v] at b*ld_b + t*ld_t + v, V <= 32.  loss (scalar) += mean_b nll_b/max(S_b,1);
 * grad [B,T,V] = d loss / d logits.  alpha_ws: b2t_ctc_workspace_floats(B,T,s_max) floats.
 * ------------------------------------------------------------------------------------- */
int64_t b2t_ctc_workspace_floats(int B, int T, int s_max);
int b2t_ctc_fwd_bwd(const float* logits, int64_t ld_t, int64_t ld_b, const int64_t* targets, int s_max,
                    const int32_t* input_lens, const int64_t* target_lens, float* alpha_ws, float* nll, float* loss,
                    float* grad, int B, int T, int V, int zero_infinity, b2t_stream stream);
/* Allocation-free variant of b2t_ctc_fwd_bwd.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments minus alpha_ws: the lattice scratch is a static device array of b2t_ctc_static_max_floats()
 * floats (2 Mi floats = 8 MiB, held by every GPU that loads the library until exit); a problem whose
 * b2t_ctc_workspace_floats(B,T,s_max) exceeds it is rejected with an argument error.  Allocates nothing (no cudaMalloc /
 * cudaFree / library allocation).  All calls share the array: calls that overlap on the GPU race - issue them on one
 * stream.  The scratch needs no clearing between calls. */
int64_t b2t_ctc_static_max_floats(void);
int b2t_ctc_fwd_bwd_static(const float* logits, int64_t ld_t, int64_t ld_b, const int64_t* targets, int s_max,
                           const int32_t* input_lens, const int64_t* target_lens, float* nll, float* loss, float* grad,
                           int B, int T, int V, int zero_infinity, b2t_stream stream);
/* argmax [B,T]; ids [B,T] = collapsed repeats without blanks (0-padded); lens [B] */
int b2t_greedy_decode(const float* logits, int64_t ld_t, int64_t ld_b, int32_t* argmax, int32_t* ids,
                      int32_t* lens, int B, int T, int V, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * Positional convolution helpers (transformers modeling_wav2vec2.py:326-368).  The grouped conv
 * runs on b2t_gemm per group over an overlapping view of the packed input:
 *   xg [G][B][Tp][64] bf16 hi/lo, xg[g][b][pad_left+t][i] = x[b][t][g*Cg+i]  (zero padded)
 *   wg [D][K*64]      bf16 hi/lo, wg[o][k*64+i] = v[o][i][k]*g[k]/||v[:,:,k]||  (weight_norm, dim=2)
 * ------------------------------------------------------------------------------------- */
int b2t_posconv_pack_input(const float* x, void* hi, void* lo, int B, int T, int D, int G, int pad_left, int Tp,
                           b2t_stream stream);
int b2t_posconv_pack_weight(const float* g, const float* v, float* norm, void* hi, void* lo, int D, int Cg, int K,
                            int G, b2t_stream stream);
/* dx[b][t][g*Cg+i] += dxg[g][b][pad_left+t][i] */
int b2t_posconv_unpack_dx(const float* dxg, float* dx, int B, int T, int D, int G, int pad_left, int Tp,
                          b2t_stream stream);
/* weight-norm backward from dwg [D][K*64] (gradient of wg): dg [K], dv [D][Cg][K] */
int b2t_posconv_weight_bwd(const float* dwg, const float* g, const float* v, const float* norm, float* dg, float* dv,
                           int D, int Cg, int K, b2t_stream stream);
/* out = dy * act'(pre), mode 1 gelu, 2 silu, 3 softsign; fp32 and/or bf16 hi/lo outputs */
int b2t_act_bwd(const float* dy, const float* pre, float* out, void* hi, void* lo, int64_t n, int mode,
                b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * Wav2Vec2-Conformer (RoPE) block pieces: transformers
 * modeling_wav2vec2_conformer.py:125-156 (rotary table), :420-527 (self-attention with the rotary
 * embedding applied to the layer-normed state BEFORE the q/k projections, v from the plain state),
 * :360-417 (convolution module: pointwise-1 -> GLU -> depthwise k=31 -> BatchNorm1d -> SiLU ->
 * pointwise-2; the two pointwise convolutions are b2t_gemm calls).
 * ------------------------------------------------------------------------------------- */
/* cos / sin tables [T, head_dim] of the rotary embedding from inv_freq [head_dim/2] */
int b2t_rope_table(const float* inv_freq, float* cos_t, float* sin_t, int T, int head_dim, b2t_stream stream);
/* y = LN(x)*gamma+beta written twice as bf16 hi/lo: plain (operand of the v projection) and rotated
 * per head (operand of the fused q/k projection); saves mean / rstd.  rows = B*T, n = heads*head_dim */
int b2t_layernorm_rope_fwd(const float* x, const float* gamma, const float* beta, float eps, void* y_hi, void* y_lo,
                           void* r_hi, void* r_lo, float* mean, float* rstd, const float* cos_t, const float* sin_t,
                           int rows, int n, int T, int head_dim, b2t_stream stream);
/* out = rope^T(d_rot) + d_plain: gradient w.r.t. the layer-normed state from the q/k branch (through
 * the transpose of the rotation) and the v branch; ld_* = row strides in elements */
int b2t_rope_bwd_add(const float* d_rot, int64_t ld_rot, const float* d_plain, int64_t ld_plain, float* out,
                     const float* cos_t, const float* sin_t, int64_t rows, int n, int T, int head_dim,
                     b2t_stream stream);
/* g [B*T, 2D] (pointwise-1 output) -> GLU -> depthwise conv over time (w [D,1,K], "same" padding) ->
 * out [B*T, D]; stats [2, D] (double) += per-channel sum and sum of squares (BatchNorm batch statistics) */
int b2t_glu_dwconv_fwd(const float* g, const float* w, float* out, double* stats, int B, int T, int D, int K,
                       b2t_stream stream);
/* BatchNorm1d (training: batch statistics from `stats`, running buffers updated with `momentum`;
 * eval: running buffers) followed by SiLU, written as bf16 hi/lo (operand of pointwise-2) */
int b2t_bn_silu_fwd(const float* x, const double* stats, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                    float* mean_out, float* rstd_out, void* y_hi, void* y_lo, int64_t M, int D, int training,
                    b2t_stream stream);
/* dgamma / dbeta of the BatchNorm (+=) from ds = gradient w.r.t. the SiLU output */
int b2t_bn_silu_bwd_stats(const float* ds, const float* x, const float* mean, const float* rstd, const float* gamma,
                          const float* beta, float* dgamma, float* dbeta, int64_t M, int D, b2t_stream stream);
/* backward of SiLU . BatchNorm . depthwise conv . GLU in one pass: dg (bf16 hi/lo, [B*T, 2D], operand of
 * the pointwise-1 backward GEMMs) and dw [D,1,K] (+=) */
int b2t_glu_dwconv_bwd(const float* ds, const float* x, const float* g, const float* w, const float* mean,
                       const float* rstd, const float* gamma, const float* beta, const float* dgamma,
                       const float* dbeta, void* dg_hi, void* dg_lo, float* dw, int B, int T, int D, int K,
                       int training, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * multi-tensor Adam + global-norm gradient clipping.
 * Replaces torch.optim.Adam(params, lr, weight_decay, eps) — L2 weight-decay form, two parameter
 * groups (src/experiments/experiment.py:25-28, src/experiments/b2t_gru_w2v_experiment.py:109-145) —
 * and torch.nn.utils.clip_grad_norm_ (src/train/train_loop.py:72-78).
 * `tensors` (device): one entry per parameter that has a gradient; `chunks` (device): pairs
 * (tensor index, chunk index), one CTA per B2T_MT_CHUNK elements.  Every call is one launch over
 * all tensors; nothing synchronises with the host (the norm and the clip coefficient stay on the
 * device).
 * ------------------------------------------------------------------------------------- */
#define B2T_MT_CHUNK 65536
#define B2T_ADAM_MAX_GROUPS 32
typedef struct {
    float* p; float* g; float* m; float* v;   /* parameter, gradient, exp_avg, exp_avg_sq (fp32) */
    int64_t n;                                /* elements */
    int32_t group;                            /* index into the groups array */
    int32_t pad_;
} b2t_mt_tensor;
typedef struct {
    /* every derived scalar is computed by the host in double and rounded once, as torch does with its Python
     * floats (1 - 0.999f in fp32 would be 4.7e-5 off 0.001) */
    float beta2, eps, weight_decay;
    float one_minus_beta1;   /* 1 - beta1 */
    float one_minus_beta2;   /* 1 - beta2 */
    float step_size;         /* lr / (1 - beta1^step) */
    float bias_c2_sqrt;      /* sqrt(1 - beta2^step) */
    int32_t pad_;
} b2t_adam_group;
/* p, m, v updated in place exactly as torch.optim.Adam's single-tensor path orders the operations;
 * grad_scale (device scalar, may be NULL) multiplies every gradient first */
int b2t_adam_multi(const b2t_mt_tensor* tensors, const int32_t* chunks, int n_chunks, const b2t_adam_group* groups,
                   int n_groups, const float* grad_scale, b2t_stream stream);
/* *sumsq (device double, caller zeroes it) += sum of squares of every gradient in the table */
int b2t_grad_sumsq_multi(const b2t_mt_tensor* tensors, const int32_t* chunks, int n_chunks, double* sumsq,
                         b2t_stream stream);
/* gradients *= min(1, max_norm / (sqrt(*sumsq) + 1e-6)); *total_norm_out = sqrt(*sumsq) */
int b2t_grad_clip_multi(const b2t_mt_tensor* tensors, const int32_t* chunks, int n_chunks, const double* sumsq,
                        float max_norm, float* total_norm_out, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * WER / CER counts of the greedy CTC decode on the device, replacing the
 * per-batch host path of src/train/evaluator.py:108-120 (argmax -> .cpu() -> tokenizer.batch_decode)
 * and :163-229 (torcheval WordErrorRate, edit_distance.SequenceMatcher character distance).
 * pred_ids [B, ld_pred]: per-frame argmax ids (b2t_greedy_decode's `argmax`); targets [B,S] padded
 * with pad_id.  vocab_chars [V][8]: byte 0 = length (<= 7), bytes 1.. = the token's characters;
 * delim_id is decoded as ' '.  Decode rules are the tokenizer's: group repeats (predictions only),
 * drop pad, join, strip, then " ' " -> "'" when cleanup != 0; eos_id >= 0 additionally cuts the
 * prediction after the first eos string.  counts[4] (int64, +=): character distance, target
 * characters, word distance, target words; per_trial [B,4] (optional) the same per trial.
 * ------------------------------------------------------------------------------------- */
int b2t_ctc_error_counts(const int32_t* pred_ids, int T, int64_t ld_pred, const int64_t* targets, int S,
                         const uint8_t* vocab_chars, int V, int pad_id, int delim_id, int eos_id, int cleanup,
                         int64_t* counts, int32_t* per_trial, int B, b2t_stream stream);

/* ---------------------------------------------------------------------------------------
 * input pipeline over device-resident trials.  frames [sum_T, C] fp32 holds
 * every trial back to back, trial i = rows trial_offsets[i] .. trial_offsets[i+1].
 * b2t_block_zscore: block-wise feature normalisation (src/datasets/preprocessing.py:60-72,100-113):
 *   y = (x - mean_block) / (std_block + eps) per channel, statistics over all frames of the trials
 *   with the same trial_block id (< 0: trial left unnormalised); ws = 3*n_blocks*C doubles.
 * b2t_collate: the collate function of src/datasets/brain2text.py:178-211 for the trials idx[0..B):
 *   out [B, t_max, C] zero padded, targets [B, s_out] from labels [N, s_all] (pad 0), input_lens,
 *   target_lens, day_idxs (all int64).
 * ------------------------------------------------------------------------------------- */
int b2t_block_zscore(const float* x, float* y, const int64_t* trial_offsets, const int32_t* trial_block, int n_trials,
                     int n_blocks, int C, double* ws, double eps, b2t_stream stream);
int b2t_collate(const float* frames, const int64_t* trial_offsets, const int32_t* idx, int B, int C, int t_max, float* out,
                const int64_t* labels, int s_all, const int64_t* label_lens, const int64_t* days, int64_t* out_targets,
                int s_out, int64_t* out_input_lens, int64_t* out_target_lens, int64_t* out_days, b2t_stream stream);

#ifdef __cplusplus
}
#endif
#endif
