// Shared device/host helpers for libb2t_sm100.so (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define B2T_OK 0
#define B2T_EARG (-1)

// thread-local error string, surfaced through b2t_last_error()
void b2t_set_error(const char* fmt, ...);

#define B2T_CHECK_ARG(cond, ...)                 \
    do {                                         \
        if (!(cond)) {                           \
            b2t_set_error(__VA_ARGS__);          \
            return B2T_EARG;                     \
        }                                        \
    } while (0)

#define B2T_CHECK_LAUNCH(name)                                                     \
    do {                                                                           \
        cudaError_t e__ = cudaGetLastError();                                      \
        if (e__ != cudaSuccess) {                                                  \
            b2t_set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
            return (int)e__;                                                       \
        }                                                                          \
    } while (0)

#define B2T_CUDA(call)                                                                  \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            b2t_set_error("%s failed: %s", #call, cudaGetErrorString(e__));             \
            return (int)e__;                                                            \
        }                                                                               \
    } while (0)

static inline int ceil_div_i(long a, long b) { return (int)((a + b - 1) / b); }

#ifdef __CUDACC__
typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// fp32 -> (hi, lo) bf16 pair with hi + lo ~= x to ~2^-17 relative (the bf16x3 GEMM operand form)
__device__ __forceinline__ void split_bf16(float x, bf16& hi, bf16& lo) {
    hi = __float2bfloat16_rn(x);
    lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

// four consecutive values as one 8-byte store per operand half
__device__ __forceinline__ uint2 pack4_bf16(bf16 a, bf16 b, bf16 c, bf16 d) {
    __nv_bfloat162 x = __halves2bfloat162(a, b), y = __halves2bfloat162(c, d);
    uint2 w;
    w.x = *reinterpret_cast<uint32_t*>(&x);
    w.y = *reinterpret_cast<uint32_t*>(&y);
    return w;
}
__device__ __forceinline__ void store_split4(const float (&o)[4], bf16* hi, bf16* lo, long off) {
    bf16 h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) split_bf16(o[e], h[e], l[e]);
    *reinterpret_cast<uint2*>(hi + off) = pack4_bf16(h[0], h[1], h[2], h[3]);
    if (lo) *reinterpret_cast<uint2*>(lo + off) = pack4_bf16(l[0], l[1], l[2], l[3]);
}

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
    const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
    return cdf + x * pdf;
}
// MUFU.RCP (<= 1 ulp) instead of the IEEE division's called slow path: ~1e-7 relative, far inside the 1e-3 gate
__device__ __forceinline__ float rcp_approx(float x) {   // one MUFU.RCP, no denormal slow-path branch
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// exp(-x) as one FMUL + MUFU.EX2 (ftz): __expf carries a denormal-range fix-up (FSETP + 2 predicated FMULs per value)
__device__ __forceinline__ float exp_neg_fast(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * -1.4426950408889634f));
    return r;
}
__device__ __forceinline__ float sigmoidf_(float x) { return rcp_approx(1.0f + exp_neg_fast(x)); }
__device__ __forceinline__ float silu_f(float x) { return x * sigmoidf_(x); }
__device__ __forceinline__ float silu_grad(float x) {
    const float s = sigmoidf_(x);
    return s * (1.0f + x * (1.0f - s));
}

// stateless counter-based uniform in [0,1): same (seed, idx) -> same value in forward and backward.
// 32-bit mixing (murmur3 finaliser over the folded index, seed words xored in before two of the
// multiplies): ~10 integer instructions; the 64-bit splitmix used first cost more than the GEMM epilogue's
// activation math.  tests/abi_emulator.py::hash_uniform is the numpy twin.
__device__ __forceinline__ float hash_uniform(unsigned long long seed, unsigned long long idx) {
    uint32_t h = ((uint32_t)idx + (uint32_t)(idx >> 32) * 0x85EBCA77u) ^ (uint32_t)seed;
    h *= 0x9E3779B1u;
    h ^= (uint32_t)(seed >> 32);
    h ^= h >> 16;
    h *= 0x85EBCA6Bu;
    h ^= h >> 13;
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return (float)(h >> 8) * (1.0f / 16777216.0f);
}
// The same decision `hash_uniform(seed, idx) < p` in integer form for inner loops: (h >> 8) * 2^-24 < p  <=>  (h >> 8) < thr
// with thr = ceil(p * 2^24) (exact: p * 2^24 is computed in double).  hash_fold() is the index-dependent first step.
__host__ __device__ __forceinline__ uint32_t drop_threshold(float p) {
    const double t = (double)p * 16777216.0;
    const uint32_t f = (uint32_t)t;
    return (double)f < t ? f + 1u : f;
}
__device__ __forceinline__ uint32_t hash_fold(unsigned long long idx) { return (uint32_t)idx + (uint32_t)(idx >> 32) * 0x85EBCA77u; }
__device__ __forceinline__ bool hash_drop(uint32_t folded, uint32_t seed_lo, uint32_t seed_hi, uint32_t thr) {
    uint32_t h = folded ^ seed_lo;
    h *= 0x9E3779B1u;
    h ^= seed_hi;
    h ^= h >> 16;
    h *= 0x85EBCA6Bu;
    h ^= h >> 13;
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return (h >> 8) < thr;
}
// Programmatic dependent launch: lets a following kernel that was launched with the programmatic-stream-serialization
// attribute (the tcgen05 GEMM) run its prologue while this grid's last CTAs finish.  Harmless otherwise: a dependent
// still executes griddepcontrol.wait before it touches memory, a normally launched successor ignores the trigger.
#define B2T_PDL_TRIGGER() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
#endif
