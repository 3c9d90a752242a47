// Harmonic / percussive separation masks on a complex STFT (frame-major [clip, frame, bin]).
//
// Replaces rosa/spectral.py:145-161 (hpss) = magphase (:113-117) + two median_filter2d calls
// (features/processing.py:74-84: 31-tap median along time -> "harm", along frequency -> "perc", reflect
// padding, torch.median = exact 16th smallest of 31) + two softmask calls (:120-142, power 2, margin m).
//
// One CTA computes a 32-frame x 64-bin output tile from a (32+30) x (64+30) magnitude tile staged in shared
// memory (magnitudes are computed on the fly from the complex input, so |D| never round-trips HBM).  Each
// output needs two 31-element selections; both run as a register-resident Batcher odd-even merge network
// (data-independent; the generated include keeps only the 152 compare-exchanges that can reach the median).
#include "common.cuh"

namespace {

constexpr int KS = 31, HALF = 15;
constexpr int TT = 32, TF = 64;
constexpr int ROWS = TT + 2 * HALF, COLS = TF + 2 * HALF;
constexpr int LDS = COLS + 1;

__device__ __forceinline__ void cswap(float& a, float& b) {
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo;
    b = hi;
}

// 16th smallest of v[0..30] -> returned; pruned Batcher odd-even merge network (median31_net.inc, generated)
__device__ __forceinline__ float median31(float (&v)[31]) {
#define CS(a, b) cswap(v[a], v[b]);
#define MV(a, b) { float t_ = v[a]; v[a] = v[b]; v[b] = t_; }
#include "median31_net.inc"
#undef CS
#undef MV
    return v[15];
}

__device__ __forceinline__ float softmask2(float X, float Xref, float split) {
    // rosa/spectral.py:120-142 with power = 2; split = 0.5 when split_zeros (margin == 1) else 0
    float Z = fmaxf(X, Xref);
    if (Z < 1.17549435e-38f) return split;
    const float a = X / Z, b = Xref / Z;
    const float m = a * a, r = b * b;
    return m / (m + r);
}

__global__ void __launch_bounds__(256) hpss_kernel(const float2* __restrict__ D, int n_frames, int n_bins, long long ld,
                                                   float margin_h, float margin_p, float2* __restrict__ out_h,
                                                   float2* __restrict__ out_p) {
    __shared__ float mag[ROWS * LDS];
    const int clip = blockIdx.z;
    const int t0 = blockIdx.y * TT, f0 = blockIdx.x * TF;
    const float2* Dc = D + (long long)clip * n_frames * ld;

    for (int i = threadIdx.x; i < ROWS * COLS; i += blockDim.x) {
        const int r = i / COLS, c = i - r * COLS;
        int t = t0 + r - HALF, f = f0 + c - HALF;
        // reflect; tiles hanging over the far edge read clamped (unused) positions
        t = (int)ssar_reflect(t, n_frames);
        f = (int)ssar_reflect(f, n_bins);
        t = min(max(t, 0), n_frames - 1);
        f = min(max(f, 0), n_bins - 1);
        const float2 z = __ldg(&Dc[(long long)t * ld + f]);
        mag[r * LDS + c] = sqrtf(fmaf(z.x, z.x, z.y * z.y));
    }
    __syncthreads();

    for (int o = threadIdx.x; o < TT * TF; o += blockDim.x) {
        const int r = o / TF, c = o - r * TF;
        const int t = t0 + r, f = f0 + c;
        if (t >= n_frames || f >= n_bins) continue;
        float v[31];
        // median along time (window rows r .. r+30 at column c+15) -> harmonic-enhanced
#pragma unroll
        for (int k = 0; k < KS; ++k) v[k] = mag[(r + k) * LDS + c + HALF];
        const float harm = median31(v);
        // median along frequency (row r+15, columns c .. c+30) -> percussive-enhanced
#pragma unroll
        for (int k = 0; k < KS; ++k) v[k] = mag[(r + HALF) * LDS + c + k];
        const float perc = median31(v);

        const long long idx = ((long long)clip * n_frames + t) * ld + f;
        const float2 z = __ldg(&D[idx]);
        const float S = mag[(r + HALF) * LDS + c + HALF];
        const float split = (margin_h == 1.f && margin_p == 1.f) ? 0.5f : 0.f;  // rosa/spectral.py:158
        // (S * mask) * exp(i angle(D)): phase = D / |D| (1 when |D| = 0)
        float2 ph = make_float2(1.f, 0.f);
        if (S > 0.f) ph = make_float2(z.x / S, z.y / S);
        if (out_h) {
            const float g = S * softmask2(harm, perc * margin_h, split);
            out_h[idx] = make_float2(g * ph.x, g * ph.y);
        }
        if (out_p) {
            const float g = S * softmask2(perc, harm * margin_p, split);
            out_p[idx] = make_float2(g * ph.x, g * ph.y);
        }
    }
}

}  // namespace

extern "C" int ssar_hpss(const float* stft, int n_clips, int n_frames, int n_bins, long long ld, float margin_harm,
                         float margin_perc, float* out_harm, float* out_perc, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(stft && (out_harm || out_perc), "hpss: null pointer");
    SSAR_REQUIRE(n_frames > HALF && n_bins > HALF, "hpss: reflect padding of 15 needs more than 15 frames and bins");
    SSAR_REQUIRE(n_clips <= 65535, "hpss: too many clips per launch");
    dim3 grid(ssar_div_up(n_bins, TF), ssar_div_up(n_frames, TT), n_clips);
    hpss_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(stft), n_frames, n_bins, ld,
                                                       margin_harm, margin_perc, reinterpret_cast<float2*>(out_harm),
                                                       reinterpret_cast<float2*>(out_perc));
    return ssar_launch_status("hpss");
}
