// conversion / layout kernels around the GEMM (HBM-bound, coalesced, vectorised where aligned)
#include "common.cuh"
#include "../../include/b2t_sm100.h"

namespace {

__global__ void split_kernel(const float* __restrict__ x, bf16* __restrict__ hi, bf16* __restrict__ lo, long rows, long cols,
                             long ld_in) {
    const long n = rows * cols;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long r = i / cols, c = i - r * cols;
        const float v = x[r * ld_in + c];
        const bf16 h = __float2bfloat16_rn(v);
        hi[i] = h;
        if (lo) lo[i] = __float2bfloat16_rn(v - __bfloat162float(h));
    }
}

// contiguous fast path: 4 floats in, 4+4 bf16 out per thread
__global__ void split_vec4_kernel(const float4* __restrict__ x, uint2* __restrict__ hi, uint2* __restrict__ lo, long n4) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
        const float4 v = x[i];
        bf16 h0, h1, h2, h3, l0, l1, l2, l3;
        split_bf16(v.x, h0, l0);
        split_bf16(v.y, h1, l1);
        split_bf16(v.z, h2, l2);
        split_bf16(v.w, h3, l3);
        __nv_bfloat162 a = __halves2bfloat162(h0, h1), b = __halves2bfloat162(h2, h3);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&a);
        o.y = *reinterpret_cast<uint32_t*>(&b);
        hi[i] = o;
        if (lo) {
            a = __halves2bfloat162(l0, l1);
            b = __halves2bfloat162(l2, l3);
            o.x = *reinterpret_cast<uint32_t*>(&a);
            o.y = *reinterpret_cast<uint32_t*>(&b);
            lo[i] = o;
        }
    }
}

// [rows, cols] -> [cols, rows] through a 32x33 smem tile (coalesced on both sides)
__global__ void split_transpose_kernel(const float* __restrict__ x, bf16* __restrict__ hi, bf16* __restrict__ lo, long rows,
                                       long cols, long ld_in) {
    __shared__ float tile[32][33];
    const long c0 = (long)blockIdx.x * 32, r0 = (long)blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const long r = r0 + j, c = c0 + threadIdx.x;
        tile[j][threadIdx.x] = (r < rows && c < cols) ? x[r * ld_in + c] : 0.0f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const long c = c0 + j, r = r0 + threadIdx.x;
        if (c < cols && r < rows) {
            const float v = tile[threadIdx.x][j];
            const bf16 h = __float2bfloat16_rn(v);
            hi[c * rows + r] = h;
            if (lo) lo[c * rows + r] = __float2bfloat16_rn(v - __bfloat162float(h));
        }
    }
}

// w[row, c*KL + k] -> out[row, k*C + c]   (or the inverse for gradients)
__global__ void permute_unfold_kernel(const float* __restrict__ w, bf16* __restrict__ hi, bf16* __restrict__ lo,
                                      float* __restrict__ out_f32, long rows, int C, int KL, int inverse) {
    const long width = (long)C * KL;
    const long n = rows * width;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long r = i / width;
        const int j = (int)(i - r * width);   // output column
        long src;
        if (!inverse) {   // output order k*C + c
            const int k = j / C, c = j - k * C;
            src = r * width + (long)c * KL + k;
        } else {          // output order c*KL + k, input order k*C + c
            const int c = j / KL, k = j - c * KL;
            src = r * width + (long)k * C + c;
        }
        const float v = w[src];
        if (out_f32) out_f32[i] = v;
        if (hi) {
            const bf16 h = __float2bfloat16_rn(v);
            hi[i] = h;
            if (lo) lo[i] = __float2bfloat16_rn(v - __bfloat162float(h));
        }
    }
}

__global__ void gemm_simt_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ c,
                                 const float* __restrict__ bias, int m, int n, int k, long sam, long sak, long sbn, long sbk,
                                 long ldc) {
    __shared__ float ta[32][33], tb[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int row = blockIdx.y * 32 + ty, col = blockIdx.x * 32 + tx;
    float acc = 0.0f;
    for (int k0 = 0; k0 < k; k0 += 32) {
        const int ka = k0 + tx;
        ta[ty][tx] = (row < m && ka < k) ? a[row * sam + ka * sak] : 0.0f;
        const int nb = blockIdx.x * 32 + ty;
        tb[ty][tx] = (nb < n && ka < k) ? b[nb * sbn + ka * sbk] : 0.0f;
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 32; ++kk) acc = fmaf(ta[ty][kk], tb[tx][kk], acc);
        __syncthreads();
    }
    if (row < m && col < n) c[row * ldc + col] = acc + (bias ? bias[col] : 0.0f);
}

inline int grid_for(long n, int block) {
    long g = (n + block - 1) / block;
    const long cap = 148L * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

extern "C" int b2t_split_bf16(const float* x, void* hi, void* lo, int64_t rows, int64_t cols, int64_t ld_in, int transpose,
                              b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(x && hi && rows > 0 && cols > 0 && ld_in >= cols, "b2t_split_bf16: bad arguments");
    if (transpose) {
        dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
        split_transpose_kernel<<<grid, dim3(32, 8), 0, stream>>>(x, (bf16*)hi, (bf16*)lo, rows, cols, ld_in);
    } else if (ld_in == cols && (rows * cols) % 4 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)hi & 7) == 0 &&
               ((uintptr_t)lo & 7) == 0) {
        const long n4 = rows * cols / 4;
        split_vec4_kernel<<<grid_for(n4, 256), 256, 0, stream>>>((const float4*)x, (uint2*)hi, (uint2*)lo, n4);
    } else {
        split_kernel<<<grid_for(rows * cols, 256), 256, 0, stream>>>(x, (bf16*)hi, (bf16*)lo, rows, cols, ld_in);
    }
    B2T_CHECK_LAUNCH("b2t_split_bf16");
    return 0;
}

extern "C" int b2t_permute_unfold_cols(const float* w, void* hi, void* lo, float* out_f32, int64_t rows, int C, int KL,
                                       int inverse, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(w && (hi || out_f32) && rows > 0 && C > 0 && KL > 0, "b2t_permute_unfold_cols: bad arguments");
    permute_unfold_kernel<<<grid_for(rows * C * KL, 256), 256, 0, stream>>>(w, (bf16*)hi, (bf16*)lo, out_f32, rows, C, KL,
                                                                           inverse);
    B2T_CHECK_LAUNCH("b2t_permute_unfold_cols");
    return 0;
}

extern "C" int b2t_gemm_simt(const float* a, const float* b, float* c, const float* bias, int m, int n, int k, int64_t sam,
                             int64_t sak, int64_t sbn, int64_t sbk, int64_t ldc, b2t_stream stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    B2T_CHECK_ARG(a && b && c && m > 0 && n > 0 && k > 0, "b2t_gemm_simt: bad arguments");
    dim3 grid((n + 31) / 32, (m + 31) / 32);
    gemm_simt_kernel<<<grid, dim3(32, 32), 0, stream>>>(a, b, c, bias, m, n, k, sam, sak, sbn, sbk, ldc);
    B2T_CHECK_LAUNCH("b2t_gemm_simt");
    return 0;
}
