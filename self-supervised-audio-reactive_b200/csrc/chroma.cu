// Chroma from a constant-Q transform, CENS post-processing and tonnetz.
//   chroma_quant_kernel : |CQT| -> 0/1 fold onto n_chroma classes (rosa/convert.py:69-117 cq_to_chroma, rosa/spectral.py
//                         :286-325 chroma_cqt with norm=False) -> L1 normalisation -> soft quantiser = cubic-spline map +
//                         sigmoid staircase (rosa/spectral.py:164-219 spline_quantize / step_function).
//   cens_smooth_kernel  : depthwise correlation along time with the normalised smoothing window, zero "same" padding,
//                         then L2 normalisation per frame (rosa/spectral.py:273-280).
//   tonnetz_kernel      : phi[6, 12] @ (chroma / ||chroma||_1) (features/audio.py:48-57).
#include "common.cuh"

namespace {

constexpr int CQ_WARPS = 8;
constexpr int MAX_CHROMA = 32;
constexpr int CQ_LD = 8;  // row loads in flight per lane

struct SplineTable {
    const float* knots;  // [n_knots]
    const float* a;      // [n_knots - 1]
    const float* b;
    const float* c;
    const float* d;
    int n_knots;
};

// rosa/spectral.py:192-202 spline_eval: index = clamp(bucketize(t, knots) - 1, 0, n_knots - 2), cubic in (t - knot)
__device__ __forceinline__ float spline_eval(const SplineTable& sp, float t) {
    // torch.bucketize(right=False): first i with knots[i] >= t
    int lo = 0, hi = sp.n_knots;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(sp.knots + mid) < t) lo = mid + 1;
        else hi = mid;
    }
    int idx = lo - 1;
    idx = min(max(idx, 0), sp.n_knots - 2);
    const float f = t - __ldg(sp.knots + idx);
    float inner = __ldg(sp.c + idx) + __ldg(sp.d + idx) * f;
    inner = __ldg(sp.b + idx) + inner * f;
    return __ldg(sp.a + idx) + inner * f;
}

// rosa/spectral.py:205-214: h * (floor(w - .5) + inv2m / (1 + exp(-2 alpha r(w)))), r(w) = (w - .5) - floor(w - .5) - .5
__device__ __forceinline__ float step_function(float w, float h, float two_alpha, float inv2m) {
    const float u = w - 0.5f;
    const float fl = floorf(u);
    const float r = u - fl - 0.5f;
    return h * (fl + inv2m * 1.f / (1.f + expf(-two_alpha * r)));
}

__global__ void __launch_bounds__(CQ_WARPS * 32) chroma_quant_kernel(
    const float2* __restrict__ cqt, long long ldc, long long n_rows, int n_bins, const int* __restrict__ chroma_of_bin,
    int n_chroma, SplineTable sp, float h, float two_alpha, float inv2m, int quantize, float* __restrict__ out) {
    extern __shared__ float sm[];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    float* mags = sm + (size_t)w * n_bins;
    // bins grouped by chroma class (ascending inside a class: the summation order of the fold matrix product), built
    // once per CTA: a lane then walks its own ~21 bins instead of testing all 252
    int* by_class = reinterpret_cast<int*>(sm + (size_t)CQ_WARPS * n_bins);  // [n_bins]
    __shared__ int class_start[MAX_CHROMA + 1];
    if (threadIdx.x < n_chroma) {
        int n = 0;
        for (int k = 0; k < n_bins; ++k) n += __ldg(chroma_of_bin + k) == (int)threadIdx.x;
        class_start[threadIdx.x + 1] = n;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        class_start[0] = 0;
        for (int c = 0; c < n_chroma; ++c) class_start[c + 1] += class_start[c];
    }
    __syncthreads();
    if (threadIdx.x < n_chroma) {
        int p = class_start[threadIdx.x];
        for (int k = 0; k < n_bins; ++k)
            if (__ldg(chroma_of_bin + k) == (int)threadIdx.x) by_class[p++] = k;
    }
    __syncthreads();
    for (long long row = (long long)blockIdx.x * CQ_WARPS + w; row < n_rows; row += (long long)gridDim.x * CQ_WARPS) {
        const float2* c = cqt + row * ldc;
        if (n_bins <= 32 * CQ_LD) {  // (252 bins) the lane's loads of a row are issued together
            float2 zz[CQ_LD];
#pragma unroll
            for (int u = 0; u < CQ_LD; ++u) {
                const int k = l + 32 * u;
                zz[u] = k < n_bins ? __ldg(c + k) : make_float2(0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < CQ_LD; ++u) {
                const int k = l + 32 * u;
                if (k < n_bins) mags[k] = sqrtf(fmaf(zz[u].x, zz[u].x, zz[u].y * zz[u].y));
            }
        } else {
            for (int k = l; k < n_bins; k += 32) {
                const float2 z = __ldg(c + k);
                mags[k] = sqrtf(fmaf(z.x, z.x, z.y * z.y));
            }
        }
        __syncwarp();
        float v = 0.f;
        if (l < n_chroma)
            for (int q = class_start[l]; q < class_start[l + 1]; ++q) v += mags[by_class[q]];
        // chroma_cqt threshold 0.0 (:317-318) is a no-op on sums of magnitudes
        float norm = warp_sum(l < n_chroma ? fabsf(v) : 0.f);
        if (l < n_chroma) {
            if (quantize) {
                v = v / norm;  // rosa/spectral.py:264 (0/0 = NaN on silent frames, like the reference)
                v = step_function(spline_eval(sp, v), h, two_alpha, inv2m);
            }
            out[row * n_chroma + l] = v;
        }
        __syncwarp();
    }
}

// block = (32 lanes over chroma, 8 frames)
__global__ void cens_smooth_kernel(const float* __restrict__ q, int cap_frames, int n_chroma, const float* __restrict__ win,
                                   int win_len, int do_l2, float* __restrict__ out, const int* __restrict__ clip_frames) {
    const int clip = blockIdx.y;
    const int t = blockIdx.x * blockDim.y + threadIdx.y;
    const int l = threadIdx.x;
    const float* qc = q + (long long)clip * cap_frames * n_chroma;
    // ragged batches: the zero "same" padding starts at the clip's own last frame; rows past it are written as zeros
    const int n_frames = clip_frames ? __ldg(clip_frames + clip) : cap_frames;
    if (t >= n_frames) {
        if (t < cap_frames && l < n_chroma) out[((long long)clip * cap_frames + t) * n_chroma + l] = 0.f;
        return;
    }
    float v = 0.f;
    if (t < n_frames && l < n_chroma) {
        const int left = (win_len - 1) / 2;  // conv1d padding="same": total pad = win_len - 1, left = total / 2
        for (int k = 0; k < win_len; ++k) {
            const int tt = t + k - left;
            if (tt >= 0 && tt < n_frames) v = fmaf(__ldg(win + k), __ldg(qc + (long long)tt * n_chroma + l), v);
        }
    }
    const float ss = warp_sum((t < n_frames && l < n_chroma) ? v * v : 0.f);
    if (t < n_frames && l < n_chroma)
        out[((long long)clip * cap_frames + t) * n_chroma + l] = do_l2 ? v / sqrtf(ss) : v;
}

__global__ void tonnetz_kernel(const float* __restrict__ chroma, long long n_rows, int n_chroma, const float* __restrict__ phi,
                               int n_out, float* __restrict__ out) {
    extern __shared__ float sphi[];
    for (int i = threadIdx.x; i < n_out * n_chroma; i += blockDim.x) sphi[i] = phi[i];
    __syncthreads();
    for (long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x; row < n_rows; row += (long long)gridDim.x * blockDim.x) {
        float c[MAX_CHROMA];
        float l1 = 0.f;
#pragma unroll 4
        for (int j = 0; j < n_chroma; ++j) {
            c[j] = chroma[row * n_chroma + j];
            l1 += fabsf(c[j]);
        }
        for (int o = 0; o < n_out; ++o) {
            float acc = 0.f;
            for (int j = 0; j < n_chroma; ++j) acc = fmaf(sphi[o * n_chroma + j], c[j] / l1, acc);
            out[row * n_out + o] = acc;
        }
    }
}

}  // namespace

extern "C" int ssar_chroma_from_cqt(const float* cqt, long long ldc, int n_clips, int n_frames, int n_bins,
                                    const int* chroma_of_bin, int n_chroma, int quantize, const float* knots,
                                    const float* sa, const float* sb, const float* sc, const float* sd, int n_knots,
                                    float step_h, float step_alpha, float* out, void* stream) {
    const long long rows = (long long)n_clips * n_frames;
    if (rows == 0) return SSAR_OK;
    SSAR_REQUIRE(cqt && chroma_of_bin && out, "chroma_from_cqt: null pointer");
    SSAR_REQUIRE(n_chroma > 0 && n_chroma <= 32 && n_bins > 0 && n_bins <= 2048, "chroma_from_cqt: bad sizes");
    SSAR_REQUIRE(!quantize || (knots && sa && sb && sc && sd && n_knots >= 2), "chroma_from_cqt: spline table missing");
    SplineTable sp = {knots, sa, sb, sc, sd, n_knots};
    // m(alpha) = 1/(1+exp(-alpha)) - 0.5 evaluated in double like the reference's numpy scalar (rosa/spectral.py:209-210)
    const double m = 1.0 / (1.0 + exp(-(double)step_alpha)) - 0.5;
    long long grid = (rows + CQ_WARPS - 1) / CQ_WARPS;
    if (grid > (long long)SSAR_NUM_SMS * 8) grid = (long long)SSAR_NUM_SMS * 8;
    if (grid > (long long)SSAR_NUM_SMS * 4) grid = (long long)SSAR_NUM_SMS * 4;  // persistent: the class lists are built per CTA
    chroma_quant_kernel<<<(unsigned)grid, CQ_WARPS * 32, (size_t)(CQ_WARPS + 1) * n_bins * sizeof(float), (cudaStream_t)stream>>>(
        reinterpret_cast<const float2*>(cqt), ldc, rows, n_bins, chroma_of_bin, n_chroma, sp, step_h, 2.f * step_alpha,
        (float)(1.0 / (2.0 * m)), quantize, out);
    return ssar_launch_status("chroma_from_cqt");
}

extern "C" int ssar_cens_smooth(const float* q, int n_clips, int n_frames, int n_chroma, const float* window, int win_len,
                                int l2_normalize, float* out, const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(q && window && out && q != out, "cens_smooth: null pointer or aliasing");
    SSAR_REQUIRE(n_chroma > 0 && n_chroma <= 32 && win_len > 0, "cens_smooth: bad sizes");
    dim3 block(32, 8);
    dim3 grid(ssar_div_up(n_frames, 8), n_clips);
    cens_smooth_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(q, n_frames, n_chroma, window, win_len, l2_normalize, out,
                                                                 clip_frames);
    return ssar_launch_status("cens_smooth");
}

extern "C" int ssar_tonnetz(const float* chroma, long long n_rows, int n_chroma, const float* phi, int n_out, float* out,
                            void* stream) {
    if (n_rows == 0) return SSAR_OK;
    SSAR_REQUIRE(chroma && phi && out, "tonnetz: null pointer");
    SSAR_REQUIRE(n_chroma > 0 && n_chroma <= MAX_CHROMA && n_out > 0 && n_out <= 64, "tonnetz: bad sizes");
    int grid = ssar_div_up(n_rows, 128);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    tonnetz_kernel<<<grid, 128, (size_t)n_out * n_chroma * sizeof(float), (cudaStream_t)stream>>>(chroma, n_rows, n_chroma, phi,
                                                                                                  n_out, out);
    return ssar_launch_status("tonnetz");
}
