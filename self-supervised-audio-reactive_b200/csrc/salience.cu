// Salience weighting + per-feature normalisation of the `Patch` inputs (ssar/random/mir.py:14-22 and :44):
//     features[k] = normalize(salience_weighted(gaussian_filter(af, sigma=2)))
// with salience_weighted(e) = (g5(e) / g80(e))^2 * e (reflect-mode Gaussians) and normalize over the WHOLE [T, C_k] array of
// each feature.  The eight features of a clip travel as the column groups of one [clip, T, C] matrix (C = 49), so the three
// Gaussians are three ssar_gauss1d launches for the whole batch and the two kernels here finish the job:
//   salience_weight : w = (short / long)^2 * env, and the min / max of every (clip, column group) as order-preserving keys
//   group_normalize : (w - min) / ((max - min) + eps), written FEATURE-BLOCKED: clip b's group g is the contiguous
//                     [T, C_g] block at out + b * T * C + T * group_start[g], the layout `Patch` binds directly
#include "common.cuh"

namespace {

constexpr int SAL_MAX_GROUPS = 32;
constexpr int SAL_THREADS = 256;

__global__ void __launch_bounds__(SAL_THREADS) salience_weight_kernel(const float* __restrict__ env,
                                                                       const float* __restrict__ sshort,
                                                                       const float* __restrict__ slong, long long n_per_clip,
                                                                       int C, const int* __restrict__ col_group, int G,
                                                                       float* __restrict__ out, uint32_t* __restrict__ min_key,
                                                                       uint32_t* __restrict__ max_key,
                                                                       const int* __restrict__ clip_frames) {
    __shared__ uint32_t smin[SAL_MAX_GROUPS], smax[SAL_MAX_GROUPS];
    const int clip = blockIdx.y;
    // ragged batches: only the clip's own rows feed the minima / maxima (the rest is written, as zeros)
    const long long n_valid = clip_frames ? min(n_per_clip, (long long)__ldg(clip_frames + clip) * C) : n_per_clip;
    if (threadIdx.x < G) {
        smin[threadIdx.x] = 0xFFFFFFFFu;
        smax[threadIdx.x] = 0u;
    }
    __syncthreads();
    const long long base = (long long)clip * n_per_clip;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x) {
        if (i >= n_valid) {
            out[base + i] = 0.f;
            continue;
        }
        const float e = env[base + i];
        const float r = sshort[base + i] / slong[base + i];  // mir.py:19: (short / long) ** 2 * envelope
        const float w = (r * r) * e;
        out[base + i] = w;
        const int g = __ldg(col_group + (int)(i % C));
        const uint32_t k = f2key(w);
        if (k < smin[g]) atomicMin(&smin[g], k);
        if (k > smax[g]) atomicMax(&smax[g], k);
    }
    __syncthreads();
    if (threadIdx.x < G) {
        const uint32_t lo = smin[threadIdx.x], hi = smax[threadIdx.x];
        if (lo <= hi) {  // the CTA saw at least one element of this group
            if (lo < __ldcg(min_key + clip * G + threadIdx.x)) atomicMin(min_key + clip * G + threadIdx.x, lo);
            if (hi > __ldcg(max_key + clip * G + threadIdx.x)) atomicMax(max_key + clip * G + threadIdx.x, hi);
        }
    }
}

__global__ void __launch_bounds__(SAL_THREADS) group_normalize_kernel(const float* __restrict__ w, int T, int C,
                                                                       const int* __restrict__ col_group,
                                                                       const int* __restrict__ group_start, int G,
                                                                       const uint32_t* __restrict__ min_key,
                                                                       const uint32_t* __restrict__ max_key, float eps,
                                                                       float* __restrict__ out) {
    __shared__ float smn[SAL_MAX_GROUPS], sden[SAL_MAX_GROUPS];
    __shared__ int sstart[SAL_MAX_GROUPS + 1];
    const int clip = blockIdx.y;
    if (threadIdx.x < G) {
        const float mn = key2f(min_key[clip * G + threadIdx.x]), mx = key2f(max_key[clip * G + threadIdx.x]);
        smn[threadIdx.x] = mn;
        sden[threadIdx.x] = (mx - mn) + eps;  // processing.py:52-55: (array - min) / (max(array - min) + 1e-8)
        sstart[threadIdx.x] = group_start[threadIdx.x];
    }
    if (threadIdx.x == 0) sstart[G] = C;
    __syncthreads();
    const long long n = (long long)T * C, base = (long long)clip * n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i / C), j = (int)(i - (long long)t * C);
        const int g = __ldg(col_group + j);
        const int cg = sstart[g + 1] - sstart[g];
        out[base + (long long)T * sstart[g] + (long long)t * cg + (j - sstart[g])] = (w[base + i] - smn[g]) / sden[g];
    }
}

}  // namespace

extern "C" int ssar_salience_weight(const float* env, const float* smooth_short, const float* smooth_long, int n_clips,
                                    int n_frames, int n_cols, const int* col_group, int n_groups, float* out,
                                    uint32_t* min_key, uint32_t* max_key, const int* clip_frames, void* stream) {
    if (n_clips <= 0 || n_frames <= 0 || n_cols <= 0) return SSAR_OK;
    SSAR_REQUIRE(env && smooth_short && smooth_long && col_group && out && min_key && max_key, "salience_weight: null pointer");
    SSAR_REQUIRE(n_groups >= 1 && n_groups <= SAL_MAX_GROUPS, "salience_weight: 1..%d column groups", SAL_MAX_GROUPS);
    const long long n = (long long)n_frames * n_cols;
    int gx = ssar_div_up(n, SAL_THREADS * 4);
    if (gx > SSAR_NUM_SMS * 4) gx = SSAR_NUM_SMS * 4;
    salience_weight_kernel<<<dim3(gx, n_clips), SAL_THREADS, 0, (cudaStream_t)stream>>>(env, smooth_short, smooth_long, n, n_cols,
                                                                                     col_group, n_groups, out, min_key, max_key,
                                                                                     clip_frames);
    return ssar_launch_status("salience_weight");
}

extern "C" int ssar_group_normalize(const float* w, int n_clips, int n_frames, int n_cols, const int* col_group,
                                    const int* group_start, int n_groups, const uint32_t* min_key, const uint32_t* max_key,
                                    float eps, float* out, void* stream) {
    if (n_clips <= 0 || n_frames <= 0 || n_cols <= 0) return SSAR_OK;
    SSAR_REQUIRE(w && col_group && group_start && min_key && max_key && out && w != out, "group_normalize: null pointer or aliasing");
    SSAR_REQUIRE(n_groups >= 1 && n_groups <= SAL_MAX_GROUPS, "group_normalize: 1..%d column groups", SAL_MAX_GROUPS);
    const long long n = (long long)n_frames * n_cols;
    int gx = ssar_div_up(n, SAL_THREADS * 4);
    if (gx > SSAR_NUM_SMS * 4) gx = SSAR_NUM_SMS * 4;
    group_normalize_kernel<<<dim3(gx, n_clips), SAL_THREADS, 0, (cudaStream_t)stream>>>(w, n_frames, n_cols, col_group, group_start,
                                                                                     n_groups, min_key, max_key, eps, out);
    return ssar_launch_status("group_normalize");
}
