// Harmonic / percussive separation masks on a complex STFT (frame-major [clip, frame, bin]).
//
// Replaces rosa/spectral.py:145-161 (hpss) = magphase (:113-117) + two median_filter2d calls
// (features/processing.py:74-84: 31-tap median along time -> "harm", along frequency -> "perc", reflect
// padding, torch.median = exact 16th smallest of 31) + two softmask calls (:120-142, power 2, margin m).
//
// One CTA (512 threads) computes a 32-frame x 64-bin output tile from a (32+30) x (64+30) magnitude tile staged in
// shared memory (magnitudes are computed on the fly from the complex input, so |D| never round-trips HBM).
//
// Selection with a SHARED CORE: four neighbouring outputs along the filter axis have windows [j, j+30], j = 0..3, of one
// 34-element span; the 28 elements [3, 30] are common to all four.  One pruned Batcher network (select28_net.inc, 138
// compare-exchanges, generated + self-checked by scripts/gen_select_net.py) leaves the core's sorted ranks 12..15 in
// place; the median of window j is then the 4th smallest of those four and the window's 3 private elements (everything
// below rank 12 of the core is below it, everything above rank 15 is above it): a 3-element sort and
// min(max(a0,b2), max(a1,b1), max(a2,b0), a3).  41 compare-exchange equivalents per output instead of 152, and
// 8.5 shared-memory loads instead of 31.
//   phase 1: medians along time   — thread = (bin column, 4 frames), lanes along bins   -> Hs tile
//   phase 2: medians along frequency — thread = (frame row, 4 bins), lanes along frames (odd row stride: conflict-free)
//            + both soft masks -> gain tiles
//   phase 3: coalesced pass: out = gain * D / |D|
#include "common.cuh"

namespace {

constexpr int HALF = 15;
constexpr int TT = 32, TF = 64;
constexpr int ROWS = TT + 2 * HALF, COLS = TF + 2 * HALF;
constexpr int LDS = COLS + 1;  // 95: odd
constexpr int LDG = TF + 1;    // 65: odd
constexpr int HP_THREADS = 512;

__device__ __forceinline__ void cswap(float& a, float& b) {
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo;
    b = hi;
}

// 4th smallest of {a0 <= a1 <= a2 <= a3} U {x, y, z}
__device__ __forceinline__ float fourth_of_4_plus_3(float a0, float a1, float a2, float a3, float x, float y, float z) {
    cswap(x, y);
    cswap(y, z);
    cswap(x, y);
    return fminf(fminf(fmaxf(a0, z), fmaxf(a1, y)), fminf(fmaxf(a2, x), a3));
}

// medians of the four 31-element windows [j, j+30] of the 34-element span s[0], s[stride], ..., s[33 * stride]
__device__ __forceinline__ void median31_x4(const float* __restrict__ s, int stride, float (&out)[4]) {
    float v[28];
#pragma unroll
    for (int k = 0; k < 28; ++k) v[k] = s[(k + 3) * stride];
    const float x0 = s[0], x1 = s[stride], x2 = s[2 * stride];
    const float y0 = s[31 * stride], y1 = s[32 * stride], y2 = s[33 * stride];
#define CS(a, b) cswap(v[a], v[b]);
#define MV(a, b) { float t_ = v[a]; v[a] = v[b]; v[b] = t_; }
#include "select28_net.inc"
#undef CS
#undef MV
    out[0] = fourth_of_4_plus_3(v[12], v[13], v[14], v[15], x0, x1, x2);
    out[1] = fourth_of_4_plus_3(v[12], v[13], v[14], v[15], x1, x2, y0);
    out[2] = fourth_of_4_plus_3(v[12], v[13], v[14], v[15], x2, y0, y1);
    out[3] = fourth_of_4_plus_3(v[12], v[13], v[14], v[15], y0, y1, y2);
}

__device__ __forceinline__ float softmask2(float X, float Xref, float split) {
    // rosa/spectral.py:120-142 with power = 2; split = 0.5 when split_zeros (margin == 1) else 0
    float Z = fmaxf(X, Xref);
    if (Z < 1.17549435e-38f) return split;
    // approximate divisions (2 ulp): the ratios are in [0, 1] and the result is held to 1e-5 against the reference;
    // the six IEEE divisions per bin were ~10 % of the kernel's instructions
    const float iz = __fdividef(1.f, Z);
    const float a = X * iz, b = Xref * iz;
    const float m = a * a, r = b * b;
    return __fdividef(m, m + r);
}

__global__ void __launch_bounds__(HP_THREADS, 2) hpss_kernel(const float2* __restrict__ D, int cap_frames, int n_bins,
                                                             long long ld, float margin_h, float margin_p,
                                                             float2* __restrict__ out_h, float2* __restrict__ out_p,
                                                             const int* __restrict__ clip_frames) {
    __shared__ float mag[ROWS * LDS];
    __shared__ float gh[TT * LDG];  // harmonic medians, then harmonic gains
    __shared__ float gp[TT * LDG];  // percussive gains
    const int clip = blockIdx.z;
    const int t0 = blockIdx.y * TT, f0 = blockIdx.x * TF;
    const float2* Dc = D + (long long)clip * cap_frames * ld;
    // ragged batches: the clip's own frame count bounds the time axis (reflection at ITS last frame); rows of the
    // capacity beyond it are written as zeros
    const int n_frames = clip_frames ? __ldg(clip_frames + clip) : cap_frames;
    if (t0 >= n_frames) {
        for (int o = threadIdx.x; o < TT * TF; o += HP_THREADS) {
            const int r = o / TF, c = o - r * TF;
            const int t = t0 + r, f = f0 + c;
            if (t >= cap_frames || f >= n_bins) continue;
            const long long idx = ((long long)clip * cap_frames + t) * ld + f;
            if (out_h) out_h[idx] = make_float2(0.f, 0.f);
            if (out_p) out_p[idx] = make_float2(0.f, 0.f);
        }
        return;
    }

    const bool interior = t0 >= HALF && t0 + TT + HALF <= n_frames && f0 >= HALF && f0 + TF + HALF <= n_bins;
    if (interior) {  // no reflection / clamping anywhere in the tile: plain strided rows
        // every load of the thread's ~12 tile elements is issued before the first magnitude is formed (ncu: the kernel's
        // largest stall was the long scoreboard — the compiler kept load -> sqrt -> store per element)
        const float2* base = Dc + (long long)(t0 - HALF) * ld + (f0 - HALF);
        constexpr int NLD = (ROWS * COLS + HP_THREADS - 1) / HP_THREADS;
        float2 zz[NLD];
#pragma unroll
        for (int u = 0; u < NLD; ++u) {
            const int i = threadIdx.x + u * HP_THREADS;
            const int r = i / COLS, c = i - r * COLS;
            zz[u] = i < ROWS * COLS ? __ldg(base + (long long)r * ld + c) : make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < NLD; ++u) {
            const int i = threadIdx.x + u * HP_THREADS;
            const int r = i / COLS, c = i - r * COLS;
            if (i < ROWS * COLS) mag[r * LDS + c] = sqrtf(fmaf(zz[u].x, zz[u].x, zz[u].y * zz[u].y));
        }
    } else {  // edge tiles (3 of the 17 tile columns of a 1025-bin spectrum): same batching, reflected indices
        constexpr int NLD = (ROWS * COLS + HP_THREADS - 1) / HP_THREADS;
        float2 zz[NLD];
#pragma unroll
        for (int u = 0; u < NLD; ++u) {
            const int i = threadIdx.x + u * HP_THREADS;
            const int r = i / COLS, c = i - r * COLS;
            int t = t0 + r - HALF, f = f0 + c - HALF;
            // reflect; tiles hanging over the far edge read clamped (unused) positions
            t = (int)ssar_reflect(t, n_frames);
            f = (int)ssar_reflect(f, n_bins);
            t = min(max(t, 0), n_frames - 1);
            f = min(max(f, 0), n_bins - 1);
            zz[u] = i < ROWS * COLS ? __ldg(&Dc[(long long)t * ld + f]) : make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < NLD; ++u) {
            const int i = threadIdx.x + u * HP_THREADS;
            const int r = i / COLS, c = i - r * COLS;
            if (i < ROWS * COLS) mag[r * LDS + c] = sqrtf(fmaf(zz[u].x, zz[u].x, zz[u].y * zz[u].y));
        }
    }
    __syncthreads();

    {   // phase 1: median along time (window rows r .. r+30 at column c+15) -> harmonic-enhanced
        const int c = threadIdx.x % TF, r = (threadIdx.x / TF) * 4;
        if (f0 + c < n_bins && t0 + r < n_frames) {  // 1025 bins = 16 tiles + 1 column: the last tile is nearly empty
            float m[4];
            median31_x4(mag + r * LDS + c + HALF, LDS, m);
#pragma unroll
            for (int j = 0; j < 4; ++j) gh[(r + j) * LDG + c] = m[j];
        }
    }
    __syncthreads();
    {   // phase 2: median along frequency (row r+15, columns c .. c+30) -> percussive-enhanced; soft masks -> gains
        const int r = threadIdx.x % TT, c = (threadIdx.x / TT) * 4;
        if (f0 + c < n_bins && t0 + r < n_frames) {
            float m[4];
            median31_x4(mag + (r + HALF) * LDS + c, 1, m);
            const float split = (margin_h == 1.f && margin_p == 1.f) ? 0.5f : 0.f;  // rosa/spectral.py:158
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float harm = gh[r * LDG + c + j], perc = m[j];
                const float S = mag[(r + HALF) * LDS + c + j + HALF];
                gh[r * LDG + c + j] = S * softmask2(harm, perc * margin_h, split);
                gp[r * LDG + c + j] = S * softmask2(perc, harm * margin_p, split);
            }
        }
    }
    __syncthreads();
    // phase 3: (S * mask) * exp(i angle(D)): phase = D / |D| (1 when |D| = 0); coalesced along bins
    constexpr int NOUT = TT * TF / HP_THREADS;
    static_assert(NOUT * HP_THREADS == TT * TF, "output tile / threads");
    float2 zo[NOUT];
#pragma unroll
    for (int u = 0; u < NOUT; ++u) {  // the tile's own D values again (L2): all loads first
        const int o = threadIdx.x + u * HP_THREADS;
        const int r = o / TF, c = o - r * TF;
        const int t = t0 + r, f = f0 + c;
        zo[u] = (t < n_frames && f < n_bins) ? __ldg(&D[((long long)clip * cap_frames + t) * ld + f]) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < NOUT; ++u) {
        const int o = threadIdx.x + u * HP_THREADS;
        const int r = o / TF, c = o - r * TF;
        const int t = t0 + r, f = f0 + c;
        if (t >= cap_frames || f >= n_bins) continue;
        const long long idx = ((long long)clip * cap_frames + t) * ld + f;
        if (t >= n_frames) {  // (ragged) past this clip's end
            if (out_h) out_h[idx] = make_float2(0.f, 0.f);
            if (out_p) out_p[idx] = make_float2(0.f, 0.f);
            continue;
        }
        const float2 z = zo[u];
        const float S = mag[(r + HALF) * LDS + c + HALF];
        float2 ph = make_float2(1.f, 0.f);
        if (S > 0.f) {
            const float is = __fdividef(1.f, S);
            ph = make_float2(z.x * is, z.y * is);
        }
        if (out_h) {
            const float g = gh[r * LDG + c];
            out_h[idx] = make_float2(g * ph.x, g * ph.y);
        }
        if (out_p) {
            const float g = gp[r * LDG + c];
            out_p[idx] = make_float2(g * ph.x, g * ph.y);
        }
    }
}

}  // namespace

extern "C" int ssar_hpss(const float* stft, int n_clips, int n_frames, int n_bins, long long ld, float margin_harm,
                         float margin_perc, float* out_harm, float* out_perc, const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(stft && (out_harm || out_perc), "hpss: null pointer");
    SSAR_REQUIRE(n_frames > HALF && n_bins > HALF, "hpss: reflect padding of 15 needs more than 15 frames and bins");
    SSAR_REQUIRE(n_clips <= 65535, "hpss: too many clips per launch");
    dim3 grid(ssar_div_up(n_bins, TF), ssar_div_up(n_frames, TT), n_clips);
    hpss_kernel<<<grid, HP_THREADS, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(stft), n_frames, n_bins, ld,
                                                       margin_harm, margin_perc, reinterpret_cast<float2*>(out_harm),
                                                       reinterpret_cast<float2*>(out_perc), clip_frames);
    return ssar_launch_status("hpss");
}
