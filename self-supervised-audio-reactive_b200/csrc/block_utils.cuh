// Block-level reductions and an in-shared-memory bitonic sort.
#pragma once
#include "common.cuh"

// all threads get the result; `scratch` must hold >= 32 floats; blockDim.x multiple of 32
template <typename Op>
__device__ __forceinline__ float block_reduce(float v, float* scratch, Op op, float identity) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();  // protect scratch from a previous use
    if (l == 0) scratch[w] = v;
    __syncthreads();
    v = (l < nw) ? scratch[l] : identity;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = op(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
struct OpSum { __device__ float operator()(float a, float b) const { return a + b; } };
struct OpMax { __device__ float operator()(float a, float b) const { return fmaxf(a, b); } };
struct OpMin { __device__ float operator()(float a, float b) const { return fminf(a, b); } };

__device__ __forceinline__ float block_sum(float v, float* scratch) { return block_reduce(v, scratch, OpSum(), 0.f); }
__device__ __forceinline__ float block_max(float v, float* scratch) { return block_reduce(v, scratch, OpMax(), -INFINITY); }
__device__ __forceinline__ float block_min(float v, float* scratch) { return block_reduce(v, scratch, OpMin(), INFINITY); }

// ascending bitonic sort of s[0..n) (n a power of two) by the whole block; ends with a barrier
__device__ __forceinline__ void block_bitonic_sort(float* s, int n) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            __syncthreads();
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const float a = s[i], b = s[ixj];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) {
                        s[i] = b;
                        s[ixj] = a;
                    }
                }
            }
        }
    }
    __syncthreads();
}

__host__ __device__ __forceinline__ int next_pow2(int n) {
    int p = 1;
    while (p < n) p <<= 1;
    return p;
}
