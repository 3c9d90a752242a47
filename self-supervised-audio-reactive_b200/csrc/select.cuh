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

// bucket holding rank k of a 256-bin histogram, found by warp 0 (8 bins per lane + shuffle scan): bcast = (bucket, k - #below)
__device__ __forceinline__ void find_bucket(const uint32_t* hist, long long k, uint32_t* bcast) {
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        uint32_t h[8], sum = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            h[j] = hist[lane * 8 + j];
            sum += h[j];
        }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        long long cum = (long long)(incl - sum);  // elements in the bins of lower lanes
        const bool mine = cum <= k && k < cum + (long long)sum;
        const bool none = __ballot_sync(0xffffffffu, mine) == 0u;  // k beyond the total: last bin, like the serial scan
        if (mine || (none && lane == 31)) {
            int b = 0;
            for (; b < 7; ++b) {
                if (cum + h[b] > k) break;
                cum += h[b];
            }
            if (none) {  // serial-scan semantics: stop at bin 255 with everything below it subtracted
                b = 7;
                cum = (long long)incl - h[7];
            }
            bcast[0] = (uint32_t)(lane * 8 + b);
            bcast[1] = (uint32_t)(k - cum);
        }
    }
}

// k-th smallest (0-based) of the selected subset; all threads return the same key.  keys != NULL: the subset's keys
// staged in shared memory (n_keys of them) — every pass then reads shared memory instead of the strided column.
__device__ inline uint32_t radix_select(const SelectCtx& c, long long k, uint32_t* hist, uint32_t* bcast,
                                        const uint32_t* keys = nullptr, int n_keys = 0) {
    uint32_t prefix = 0, mask = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
        __syncthreads();
        if (keys) {
            for (int i = threadIdx.x; i < n_keys; i += blockDim.x) {
                const uint32_t key = keys[i];
                if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
            }
        } else {
            for (long long t = threadIdx.x; t < c.T; t += blockDim.x) {
                float v;
                if (!sel_value(c, t, v)) continue;
                const uint32_t key = f2key(v);
                if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
            }
        }
        __syncthreads();
        find_bucket(hist, k, bcast);
        __syncthreads();
        prefix |= bcast[0] << shift;
        mask |= 255u << shift;
        k = bcast[1];
        __syncthreads();
    }
    return prefix;
}
