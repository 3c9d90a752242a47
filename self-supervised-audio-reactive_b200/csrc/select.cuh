// Exact order statistics by MSB radix select on order-preserving keys (one CTA walks one strided column).
#pragma once
#include "common.cuh"

// subset 0: every non-NaN element; 1: strict interior local maxima (features/processing.py:107-115);
// subset 2: elements whose companion aux[t] is > 0 (rosa/pitch.py:12-16: mag[pitch > 0])
struct SelectCtx {
    const float* xc;   // column base
    const float* aux;  // subset 2 companion (same indexing)
    long long T;
    int C, subset;
    float sub, div;
    bool pre;
};

__device__ __forceinline__ bool sel_value(const SelectCtx& c, long long t, float& v) {
    v = c.xc[(long long)t * c.C];
    if (c.subset == 1) {
        const float plus = c.xc[(t + 1 < c.T ? t + 1 : c.T - 1) * c.C];
        const float minus = c.xc[(t > 0 ? t - 1 : 0) * c.C];
        if (!(v > plus && v > minus)) return false;
    }
    if (c.subset == 2 && !(c.aux[(long long)t * c.C] > 0.f)) return false;
    if (v != v) return false;
    if (c.pre) v = (v - c.sub) / c.div;
    return true;
}

// k-th smallest (0-based) of the selected subset; all threads return the same key
__device__ inline uint32_t radix_select(const SelectCtx& c, long long k, uint32_t* hist, uint32_t* bcast) {
    uint32_t prefix = 0, mask = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        for (long long t = threadIdx.x; t < c.T; t += blockDim.x) {
            float v;
            if (!sel_value(c, t, v)) continue;
            const uint32_t key = f2key(v);
            if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            long long cum = 0;
            int b = 0;
            for (; b < 255; ++b) {
                if (cum + hist[b] > k) break;
                cum += hist[b];
            }
            bcast[0] = (uint32_t)b;
            bcast[1] = (uint32_t)(k - cum);
        }
        __syncthreads();
        prefix |= bcast[0] << shift;
        mask |= 255u << shift;
        k = bcast[1];
        __syncthreads();
    }
    return prefix;
}

