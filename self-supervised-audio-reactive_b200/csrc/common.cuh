// Shared helpers for the ssar_b200 CUDA kernels (sm_100a).  No torch types anywhere in csrc/.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ssar_b200.h"

#define SSAR_NUM_SMS 148  // B200: 2 dies x 74 SMs; grids are sized in multiples of this

void ssar_set_error(const char* fmt, ...);

#define SSAR_REQUIRE(cond, ...)            \
    do {                                   \
        if (!(cond)) {                     \
            ssar_set_error(__VA_ARGS__);   \
            return SSAR_ERR_ARG;           \
        }                                  \
    } while (0)

static inline int ssar_launch_status(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        ssar_set_error("%s: %s", what, cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    return SSAR_OK;
}

static inline int ssar_div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

// reflect (edge sample not repeated) index into [0, n): valid while |overshoot| < n
__host__ __device__ __forceinline__ long long ssar_reflect(long long i, long long n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }

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
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// order-preserving float <-> uint32 key (total order incl. negatives; NaN sorts last)
__host__ __device__ __forceinline__ uint32_t f2key(float f) {
    uint32_t u;
#ifdef __CUDA_ARCH__
    u = __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; u = c.u;
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float key2f(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

// atomic max/min on floats via the order-preserving key (buffers hold keys; init with ssar_fill_u32)
// The running extreme is read first (from L2: it only ever moves one way, so a stale value costs one redundant atomic,
// never a wrong result): tens of thousands of warps fold into ONE word per clip, and same-address atomics serialise at
// ~1-2 ns each in L2 — unfiltered they were most of spectral contrast's run time.
__device__ __forceinline__ void atomic_max_key(uint32_t* addr, float v) {
    const uint32_t k = f2key(v);
    if (k > __ldcg(addr)) atomicMax(addr, k);
}
__device__ __forceinline__ void atomic_min_key(uint32_t* addr, float v) {
    const uint32_t k = f2key(v);
    if (k < __ldcg(addr)) atomicMin(addr, k);
}
