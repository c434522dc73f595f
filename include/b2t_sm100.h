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
    int32_t aux_mode;         /* 0 none, 1 gelu', 2 silu', 3 softsign' */
    int32_t atomic;           /* out_f32 += value with red.add (split-K, overlapped folds) */
    int32_t split_k;          /* >=1; >1 requires atomic, act==0, no hi/lo/pre */
    float   alpha;
    int32_t block_n;          /* 0 = auto, else 64 / 128 / 256 */
} b2t_gemm_args;

int b2t_gemm(const b2t_gemm_args* args, b2t_stream stream);

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

#ifdef __cplusplus
}
#endif
#endif
