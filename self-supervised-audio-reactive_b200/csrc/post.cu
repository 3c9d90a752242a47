// Post-processing of per-frame feature matrices [clip, T, C] (time-major, C contiguous):
// Gaussian temporal smoothing, min/max normalisation, per-column order statistics (radix select), percentile
// clamps, emphasize, spectral flux, and the per-column (count, sum, sum of squares) statistics that feed the
// cross-GPU mean/std all-reduce.
// Replaces ssar/features/processing.py:10-48 (gaussian_filter), :52-55 (normalize), :58-61 (standardize),
// :88-98 (spectral_flux, onset_envelope), :101-129 (percentile clamps), :132-138 (emphasize) and the dataset
// statistics of ssar/supervised/data.py:317-319.
#include "block_utils.cuh"
#include "select.cuh"

namespace {

// ---------------------------------------------------------------------------------------------------------------
// gaussian_filter: out[b,t,j] = sum_k w[k] * x[b, idx(t + k - r), j]
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int pad_index(int v, int T, int radius, int mode) {
    if (radius > T) v = min(max(v, -T), 2 * T - 1);  // outer replicate pad of the short-sequence branch (:30-35)
    if (mode == 0) {  // circular
        v %= T;
        if (v < 0) v += T;
    } else if (mode == 1) {  // reflect (edge not repeated)
        if (v < 0) v = -v;
        if (v >= T) v = 2 * (T - 1) - v;
        v = min(max(v, 0), T - 1);
    } else {  // replicate
        v = min(max(v, 0), T - 1);
    }
    return v;
}

// TPT consecutive time steps per thread.  The taps travel through a circular register window (slot u holds tap i for
// i = u mod TPT), so an input row costs one global load, one shared-memory tap read and TPT FMAs; every output still
// accumulates its taps in ascending order, so the result does not depend on TPT.  TPT = 4 for the short kernels of the
// [T, 59] smoothing, 16 for the long ones (sigma 80 of random/mir.py:14-22: 641 taps).
// Ragged batches: clip_frames[c] <= T rows of clip c are real; the padding (wrap / reflection) happens at ITS length and the
// rows beyond are written as zeros.
template <int TPT>
__global__ void gauss1d_kernel(const float* __restrict__ x, int cap_T, long long inner, const float* __restrict__ w,
                               int radius, int mode, float* __restrict__ out, const int* __restrict__ clip_frames) {
    extern __shared__ float ws[];
    const int taps = 2 * radius + 1;
    for (int i = threadIdx.x; i < taps; i += blockDim.x) ws[i] = w[i];
    __syncthreads();
    const int clip = blockIdx.z;
    const float* xc = x + (long long)clip * cap_T * inner;
    float* oc = out + (long long)clip * cap_T * inner;
    const int T = clip_frames ? __ldg(clip_frames + clip) : cap_T;
    // (time block, column) pairs are packed into the CTA, columns fastest: narrow matrices ([T, 4] envelopes) still
    // fill their warps
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long tb = gid / inner;
    const long long j = gid - tb * inner;
    if (tb * TPT >= cap_T) return;
    const int t0 = (int)tb * TPT;
    if (t0 >= T) {  // (ragged) a block of rows past the clip's end
#pragma unroll
        for (int o = 0; o < TPT; ++o)
            if (t0 + o < cap_T) oc[(long long)(t0 + o) * inner + j] = 0.f;
        return;
    }
    float acc[TPT], wr[TPT];
#pragma unroll
    for (int o = 0; o < TPT; ++o) acc[o] = 0.f, wr[o] = 0.f;
    // input rows t0 - r .. t0 + TPT - 1 + r; row i contributes to output o with tap k = i - o (relative indices)
    const int n_in = taps + TPT - 1;
    if (t0 - radius >= 0 && t0 + TPT - 1 + radius < T) {
        // interior: no padding arithmetic, the row pointer just advances
        const float* __restrict__ px = xc + (long long)(t0 - radius) * inner + j;
        for (int i0 = 0; i0 < n_in; i0 += TPT) {
#pragma unroll
            for (int u = 0; u < TPT; ++u) {
                const int i = i0 + u;
                if (i < n_in) {  // (uniform across the CTA)
                    const float v = __ldg(px + (long long)i * inner);
                    wr[u] = i < taps ? ws[i] : 0.f;  // tap i enters the window; taps beyond the kernel are zeros
#pragma unroll
                    for (int o = 0; o < TPT; ++o) acc[o] = fmaf(wr[(u - o + TPT) % TPT], v, acc[o]);
                }
            }
        }
    } else {
        for (int i0 = 0; i0 < n_in; i0 += TPT) {
#pragma unroll
            for (int u = 0; u < TPT; ++u) {
                const int i = i0 + u;
                if (i < n_in) {
                    const float v = __ldg(xc + (long long)pad_index(t0 + i - radius, T, radius, mode) * inner + j);
                    wr[u] = i < taps ? ws[i] : 0.f;
#pragma unroll
                    for (int o = 0; o < TPT; ++o) acc[o] = fmaf(wr[(u - o + TPT) % TPT], v, acc[o]);
                }
            }
        }
    }
#pragma unroll
    for (int o = 0; o < TPT; ++o)
        if (t0 + o < cap_T) oc[(long long)(t0 + o) * inner + j] = t0 + o < T ? acc[o] : 0.f;
}

// Long kernels (sigma 80 of random/mir.py:14-22: 641 taps): a CTA stages the (tile_rows + 2 radius) x inner input rows of
// its output tile in shared memory ONCE (padding resolved per row while loading), then every thread makes 16 consecutive
// outputs of one column from the tile with the same register tap window and the same ascending-tap fmaf chain as
// gauss1d_kernel — identical results; two conflict-free shared-memory reads per 16 FMAs instead of a global load and the
// padding arithmetic per tap.
constexpr int GT_TPT = 16;
__host__ __device__ inline int gt_round16(int v) { return (v + GT_TPT - 1) / GT_TPT * GT_TPT; }
// shared memory: taps zero-padded to a multiple of 16 entries | tile of (16 * groups + padded taps) rows, zero rows past
// the staged ones: the inner loop then runs in whole blocks of 16 input rows with no bound checks at all
__global__ void __launch_bounds__(512) gauss1d_tiled_kernel(const float* __restrict__ x, int cap_T, int inner,
                                                            const float* __restrict__ w, int radius, int mode,
                                                            float* __restrict__ out, const int* __restrict__ clip_frames,
                                                            int tile_rows) {
    extern __shared__ float gsm[];
    const int taps = 2 * radius + 1;
    const int n_pad = gt_round16(taps + GT_TPT - 1);  // input rows walked per group of 16 outputs
    float* ws = gsm;
    float* tile = gsm + n_pad;
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * cap_T * inner;
    float* oc = out + (long long)clip * cap_T * inner;
    const int T = clip_frames ? __ldg(clip_frames + clip) : cap_T;
    const int t0 = blockIdx.x * tile_rows;
    const int rows_out = min(tile_rows, cap_T - t0);
    if (t0 >= T) {  // (ragged) a tile past the clip's end
        for (int i = threadIdx.x; i < rows_out * inner; i += blockDim.x) oc[(long long)t0 * inner + i] = 0.f;
        return;
    }
    for (int i = threadIdx.x; i < n_pad; i += blockDim.x) ws[i] = i < taps ? w[i] : 0.f;
    const int groups = (rows_out + GT_TPT - 1) / GT_TPT;
    const int rows_in = rows_out + 2 * radius;
    const int rows_all = (groups - 1) * GT_TPT + n_pad;  // >= rows_in
    for (int i = threadIdx.x; i < rows_all * inner; i += blockDim.x) {
        const int r = i / inner, j = i - r * inner;
        tile[i] = r < rows_in ? __ldg(xc + (long long)pad_index(t0 - radius + r, T, radius, mode) * inner + j) : 0.f;
    }
    __syncthreads();
    for (int item = threadIdx.x; item < groups * inner; item += blockDim.x) {
        const int g = item / inner, j = item - g * inner;
        const int r0 = g * GT_TPT;
        float acc[GT_TPT], wr[GT_TPT];
#pragma unroll
        for (int o = 0; o < GT_TPT; ++o) acc[o] = 0.f, wr[o] = 0.f;
        const float* col = tile + r0 * inner + j;
        const float* wp = ws;
        for (int i0 = 0; i0 < n_pad; i0 += GT_TPT) {
#pragma unroll
            for (int u = 0; u < GT_TPT; ++u) {
                const float v = *col;
                col += inner;
                wr[u] = wp[u];  // tap i0 + u enters the window (zeros beyond the kernel)
#pragma unroll
                for (int o = 0; o < GT_TPT; ++o) acc[o] = fmaf(wr[(u - o + GT_TPT) % GT_TPT], v, acc[o]);
            }
            wp += GT_TPT;
        }
#pragma unroll
        for (int o = 0; o < GT_TPT; ++o) {
            const int t = t0 + r0 + o;
            if (r0 + o < rows_out) oc[(long long)t * inner + j] = t < T ? acc[o] : 0.f;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// per-(clip, column) min / max of [clip, T, C]
// ---------------------------------------------------------------------------------------------------------------
__global__ void col_minmax_kernel(const float* __restrict__ x, int cap_T, int C, float* __restrict__ mn_out,
                                  float* __restrict__ mx_out, const int* __restrict__ clip_frames) {
    __shared__ float scratch[32];
    const int col = blockIdx.x, clip = blockIdx.y;
    const float* xc = x + (long long)clip * cap_T * C + col;
    const int T = clip_frames ? __ldg(clip_frames + clip) : cap_T;  // ragged batches: the clip's own rows only
    float mn = INFINITY, mx = -INFINITY;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const float v = xc[(long long)t * C];
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = block_min(mn, scratch);
    mx = block_max(mx, scratch);
    if (threadIdx.x == 0) {
        if (mn_out) mn_out[clip * C + col] = mn;
        if (mx_out) mx_out[clip * C + col] = mx;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// per-(clip, column) quantile by MSB radix select on order-preserving keys.  One CTA per (column, clip).
//   subset 0: every non-NaN element;  subset 1: strict interior local maxima (processing.py:107-115)
//   interp 0: torch.quantile "linear" (rank = float(q) * (n - 1) in fp32, torch.lerp in fp32)
//   interp 3: efficient_quantile midpoint (rank = double(float(q)) * (n - 1), trunc / ceil neighbours, mean in fp64)
//   pre-transform: v -> (v - sub[col]) / div[col] when sub != NULL (emphasize's rescale, processing.py:133-136)
// ---------------------------------------------------------------------------------------------------------------
__global__ void col_quantile_kernel(const float* __restrict__ x, int cap_T, int C, int subset, double q, int interp,
                                    const float* __restrict__ pre_sub, const float* __restrict__ pre_div,
                                    float* __restrict__ out, int* __restrict__ count_out, int staged_keys,
                                    const int* __restrict__ clip_frames) {
    __shared__ uint32_t hist[256];
    __shared__ uint32_t bcast[2];
    __shared__ int s_count;
    const int col = blockIdx.x, clip = blockIdx.y;
    const int T = clip_frames ? __ldg(clip_frames + clip) : cap_T;  // ragged batches: the clip's own rows only
    SelectCtx c;
    c.xc = x + (long long)clip * cap_T * C + col;
    c.T = T;
    c.C = C;
    c.aux = nullptr;
    c.subset = subset;
    c.pre = pre_sub != nullptr;
    c.sub = c.pre ? pre_sub[clip * C + col] : 0.f;
    c.div = c.pre ? pre_div[clip * C + col] : 1.f;
    if (threadIdx.x == 0) s_count = 0;
    __syncthreads();
    extern __shared__ uint32_t q_keys[];  // staged keys of the selected subset (when the launch provides room for T)
    const bool staged = staged_keys != 0;
    int local = 0;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        float v;
        if (!sel_value(c, t, v)) continue;
        if (staged) q_keys[atomicAdd(&s_count, 1)] = f2key(v);
        else ++local;
    }
    if (local) atomicAdd(&s_count, local);
    __syncthreads();
    const int n = s_count;
    if (count_out && threadIdx.x == 0) count_out[clip * C + col] = n;
    if (n == 0) {  // torch.quantile raises on an empty input; efficient_quantile returns NaN. The host mirrors that.
        if (threadIdx.x == 0) out[clip * C + col] = __int_as_float(0x7fc00000);
        return;
    }
    long long lo, hi;
    float w32 = 0.f;
    if (interp == 0) {
        const float rank = (float)q * (float)(n - 1);
        const float below = floorf(rank);
        lo = (long long)below;
        hi = (long long)ceilf(rank);
        w32 = rank - below;
    } else {
        const double rank = (double)(float)q * (double)(n - 1);
        lo = (long long)rank;
        hi = (long long)ceil(rank);
    }
    const uint32_t* kp = staged ? q_keys : nullptr;
    const float vlo = key2f(radix_select(c, lo, hist, bcast, kp, n));
    const float vhi = hi == lo ? vlo : key2f(radix_select(c, hi, hist, bcast, kp, n));
    if (threadIdx.x == 0) {
        float r;
        if (interp == 0) {  // ATen lerp: w < 0.5 ? a + w (b - a) : b - (b - a)(1 - w)
            const float d = vhi - vlo;
            r = w32 < 0.5f ? vlo + w32 * d : vhi - d * (1.f - w32);
        } else {
            const double delta = hi > lo ? 0.5 : 0.0;
            const double a = (double)vlo, b = (double)vhi;
            r = (float)(delta < 0.5 ? a + delta * (b - a) : b - (b - a) * (1.0 - delta));
        }
        out[clip * C + col] = r;
    }
}

// clamp [clip, T, C] by per-(clip, column) bounds (either may be NULL)
__global__ void clamp_cols_kernel(const float* __restrict__ x, long long n_per_clip, int C, const float* __restrict__ lo,
                                  const float* __restrict__ hi, float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * n_per_clip;
    float* oc = out + (long long)clip * n_per_clip;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x) {
        const int col = (int)(i % C);
        float v = xc[i];
        // torch.clamp(min, max): min applied first, then max; NaN propagates
        if (lo) v = (v != v) ? v : fmaxf(v, lo[clip * C + col]);
        if (hi) v = (v != v) ? v : fminf(v, hi[clip * C + col]);
        oc[i] = v;
    }
}

// normalize (processing.py:52-55): (x - min) / ((max - min) + 1e-8) over the whole clip; min/max as keys
__global__ void normalize_kernel(const float* __restrict__ x, long long n_per_clip, const uint32_t* __restrict__ min_key,
                                 const uint32_t* __restrict__ max_key, float eps, float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float mn = key2f(min_key[clip]);
    const float den = (key2f(max_key[clip]) - mn) + eps;
    const float* xc = x + (long long)clip * n_per_clip;
    float* oc = out + (long long)clip * n_per_clip;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x)
        oc[i] = (xc[i] - mn) / den;
}

// normalize for small per-clip arrays (<= 256 k elements, L2 resident): ONE CTA per clip finds min / max and rescales —
// one launch instead of two key fills + minmax + normalize.  Same min / max (exact) and the same expression per element.
__global__ void __launch_bounds__(1024) normalize_fused_kernel(const float* __restrict__ x, long long cap_per_clip, float eps,
                                                                float* __restrict__ out, const int* __restrict__ clip_rows,
                                                                long long row_elems) {
    __shared__ float scratch[32];
    __shared__ float s_mn, s_mx;
    const int clip = blockIdx.x;
    const float* xc = x + (long long)clip * cap_per_clip;
    float* oc = out + (long long)clip * cap_per_clip;
    // ragged batches: the first clip_rows[c] * row_elems elements of the clip are real; the rest is written as zeros
    const long long n_per_clip = clip_rows ? min(cap_per_clip, (long long)__ldg(clip_rows + clip) * row_elems) : cap_per_clip;
    for (long long i = n_per_clip + threadIdx.x; i < cap_per_clip; i += blockDim.x) oc[i] = 0.f;
    float mn = INFINITY, mx = -INFINITY;
    for (long long i = threadIdx.x; i < n_per_clip; i += blockDim.x) {
        const float v = xc[i];
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = block_min(mn, scratch);
    __syncthreads();
    mx = block_max(mx, scratch);
    if (threadIdx.x == 0) {
        s_mn = mn;
        s_mx = mx;
    }
    __syncthreads();
    const float lo = s_mn, den = (s_mx - s_mn) + eps;
    for (long long i = threadIdx.x; i < n_per_clip; i += blockDim.x) oc[i] = (xc[i] - lo) / den;
}

// emphasize (processing.py:132-138): xn = (x - min)/max' ; xn * (1 + tanh(s (xn - qv))) * max' + min
__global__ void emphasize_kernel(const float* __restrict__ x, long long n_per_clip, int C, const float* __restrict__ mn,
                                 const float* __restrict__ rng, const float* __restrict__ qv, float strength,
                                 float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * n_per_clip;
    float* oc = out + (long long)clip * n_per_clip;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x) {
        const int col = clip * C + (int)(i % C);
        const float m = mn[col], r = rng[col];
        float v = (xc[i] - m) / r;
        v = v * (1.f + tanhf(strength * (v - qv[col])));
        oc[i] = v * r + m;
    }
}

// per-column range = max(x - min) (== max - min in fp32) from separate min / max arrays
__global__ void range_kernel(const float* __restrict__ mn, const float* __restrict__ mx, int n, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = mx[i] - mn[i];
}

// spectral_flux (processing.py:88-89): out[t] = x[t+1] - x[t], last row = 0 - x[T-1]
__global__ void flux_kernel(const float* __restrict__ x, int T, long long inner, float* __restrict__ out) {
    const int clip = blockIdx.y;
    const long long n = (long long)T * inner;
    const float* xc = x + clip * n;
    float* oc = out + clip * n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float nxt = (i + inner < n) ? xc[i + inner] : 0.f;
        oc[i] = nxt - xc[i];
    }
}

// half-wave rectified row sums (processing.py:92): u[t] = sum_j 0.5 (f + |f|)
__global__ void hwr_rowsum_kernel(const float* __restrict__ x, long long rows, int inner, float* __restrict__ out) {
    const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, l = threadIdx.x & 31;
    for (long long r = (long long)blockIdx.x * warps + w; r < rows; r += (long long)gridDim.x * warps) {
        float acc = 0.f;
        for (int j = l; j < inner; j += 32) {
            const float f = x[r * inner + j];
            acc += 0.5f * (f + fabsf(f));
        }
        acc = warp_sum(acc);
        if (l == 0) out[r] = acc;
    }
}

// (count, sum, sum of squares) per column in fp64, atomically accumulated into stats[3, C] (zeroed by the caller).
// rows = n_clips * T flattened (the statistic is over all windows and frames, supervised/data.py:317-319).
__global__ void feature_stats_kernel(const float* __restrict__ x, long long rows, int C, double* __restrict__ stats,
                                     int cap_T, const int* __restrict__ clip_frames) {
    // blockDim = (32, 8): x = column lane, y = row lane
    for (int col = blockIdx.x * 32 + threadIdx.x; col < C; col += gridDim.x * 32) {
        double s = 0.0, ss = 0.0;
        long long cnt = 0;
        for (long long r = (long long)blockIdx.y * blockDim.y + threadIdx.y; r < rows; r += (long long)gridDim.y * blockDim.y) {
            if (clip_frames && (int)(r % cap_T) >= __ldg(clip_frames + r / cap_T)) continue;  // (ragged) past the clip's end
            const double v = (double)x[r * C + col];
            s += v;
            ss += v * v;
            ++cnt;
        }
        atomicAdd(&stats[col], (double)cnt);
        atomicAdd(&stats[C + col], s);
        atomicAdd(&stats[2 * C + col], ss);
    }
}

// overlapping_slices (supervised/data.py:255-258): window k = rows [offsets[k], offsets[k] + L) of x [T, inner] ->
// out [n, L, inner].  A window is L * inner contiguous floats on both sides: 128-bit copies when aligned.
__global__ void window_slices_kernel(const float* __restrict__ x, long long inner, const long long* __restrict__ offsets,
                                     int L, float* __restrict__ out) {
    const long long per = (long long)L * inner;
    const float* src = x + __ldg(offsets + blockIdx.y) * inner;
    float* dst = out + (long long)blockIdx.y * per;
    if ((per & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (per >> 2); i += (long long)gridDim.x * blockDim.x)
            d4[i] = __ldg(s4 + i);
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < per; i += (long long)gridDim.x * blockDim.x)
            dst[i] = __ldg(src + i);
    }
}

}  // namespace

#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" int ssar_window_slices(const float* x, long long n_rows, long long inner, const long long* offsets, int n_windows,
                                  int window, float* out, void* stream) {
    if (n_windows == 0 || window == 0 || inner == 0) return SSAR_OK;
    SSAR_REQUIRE(x && offsets && out && n_rows >= window && n_windows <= 65535 * 64, "window_slices: null pointer or bad sizes");
    const long long per = (long long)window * inner;
    int gx = ssar_div_up(per, 256 * 8);
    if (gx > 64) gx = 64;
    for (int w0 = 0; w0 < n_windows; w0 += 65535) {
        const int nw = n_windows - w0 < 65535 ? n_windows - w0 : 65535;
        window_slices_kernel<<<dim3(gx, nw), 256, 0, ST(stream)>>>(x, inner, offsets + w0, window, out + (long long)w0 * per);
    }
    return ssar_launch_status("window_slices");
}

extern "C" int ssar_gauss1d(const float* x, int n_clips, int n_frames, long long inner, const float* weights, int radius,
                            int mode, float* out, const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_frames == 0 || inner == 0) return SSAR_OK;
    SSAR_REQUIRE(x && weights && out, "gauss1d: null pointer");
    SSAR_REQUIRE(x != out, "gauss1d: in-place filtering is not supported");
    SSAR_REQUIRE(radius >= 0 && radius <= 3 * n_frames, "gauss1d: radius must be <= 3 * n_frames");
    SSAR_REQUIRE(mode >= 0 && mode <= 2, "gauss1d: mode must be 0 (circular), 1 (reflect) or 2 (replicate)");
    SSAR_REQUIRE(mode != 1 || radius < n_frames, "gauss1d: reflect padding needs radius < n_frames");
    SSAR_REQUIRE((2 * radius + 1) * sizeof(float) <= 200 * 1024, "gauss1d: kernel too long");
    SSAR_REQUIRE(n_clips <= 65535, "gauss1d: too many clips per launch");
    if (radius >= 32 && inner <= 4096) {  // (shorter kernels: the tile staging costs more than it saves, measured)
        // tiled variant: tile rows chosen so that the launch is a whole number of waves of one CTA per SM (when it fits)
        const int n_pad = gt_round16(2 * radius + 1 + GT_TPT - 1);
        const long long smem_cap = 200 * 1024 - (long long)n_pad * 4;
        const long long max_rows = smem_cap / (4 * inner) - n_pad;  // rows_all = 16 (groups - 1) + n_pad <= tile_rows + n_pad
        if (max_rows >= 4 * GT_TPT) {
            int tile_rows = (int)(max_rows / GT_TPT * GT_TPT);
            // fewest tiles per clip that fit, then the smallest tile height giving that many tiles
            const int tiles = (n_frames + tile_rows - 1) / tile_rows;
            int waves_tiles = tiles;
            // n_clips * tiles CTAs: round the tile count up until the grid fills its last wave to >= 85 %
            for (int tt = tiles; tt <= tiles + 8; ++tt) {
                const long long ctas = (long long)n_clips * tt;
                const long long waves = (ctas + SSAR_NUM_SMS - 1) / SSAR_NUM_SMS;
                if ((double)ctas / (double)(waves * SSAR_NUM_SMS) >= 0.85) {
                    waves_tiles = tt;
                    break;
                }
            }
            tile_rows = ((n_frames + waves_tiles - 1) / waves_tiles + GT_TPT - 1) / GT_TPT * GT_TPT;
            const int n_tiles = (n_frames + tile_rows - 1) / tile_rows;
            const size_t smem = ((size_t)n_pad + (size_t)(tile_rows + n_pad) * inner) * sizeof(float);
            cudaFuncSetAttribute(gauss1d_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            gauss1d_tiled_kernel<<<dim3(n_tiles, n_clips), 512, smem, ST(stream)>>>(x, n_frames, (int)inner, weights, radius, mode, out,
                                                                                  clip_frames, tile_rows);
            return ssar_launch_status("gauss1d");
        }
    }
    const int threads = 128;
    // long kernels on enough rows: 16 outputs per thread (16 FMAs per loaded row instead of 4)
    const int tpt = (radius >= 32 && (long long)n_clips * inner * ((n_frames + 15) / 16) >= (long long)SSAR_NUM_SMS * threads) ? 16 : 4;
    const long long items = inner * ((n_frames + tpt - 1) / tpt);
    SSAR_REQUIRE((items + threads - 1) / threads <= 0x7fffffffLL, "gauss1d: too many elements per launch");
    dim3 grid((unsigned)((items + threads - 1) / threads), 1, n_clips);
    const size_t smem = (2 * radius + 1) * sizeof(float);
    auto kern = tpt == 16 ? gauss1d_kernel<16> : gauss1d_kernel<4>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    kern<<<grid, threads, smem, ST(stream)>>>(x, n_frames, inner, weights, radius, mode, out, clip_frames);
    return ssar_launch_status("gauss1d");
}

extern "C" int ssar_col_minmax(const float* x, int n_clips, int n_frames, int n_cols, float* min_out, float* max_out,
                               const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_cols == 0) return SSAR_OK;
    SSAR_REQUIRE(x && n_frames > 0, "col_minmax: null pointer or empty columns");
    col_minmax_kernel<<<dim3(n_cols, n_clips), 256, 0, ST(stream)>>>(x, n_frames, n_cols, min_out, max_out, clip_frames);
    return ssar_launch_status("col_minmax");
}

extern "C" int ssar_col_quantile(const float* x, int n_clips, int n_frames, int n_cols, int subset, double q, int interp,
                                 const float* pre_sub, const float* pre_div, float* out, int* count_out,
                                 const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_cols == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out, "col_quantile: null pointer");
    SSAR_REQUIRE(q >= 0.0 && q <= 1.0, "col_quantile: q must be in [0, 1]");
    SSAR_REQUIRE(subset == 0 || subset == 1, "col_quantile: subset must be 0 (all) or 1 (strict interior peaks)");
    SSAR_REQUIRE(interp == 0 || interp == 3, "col_quantile: interp must be 0 (linear) or 3 (midpoint)");
    SSAR_REQUIRE((pre_sub == nullptr) == (pre_div == nullptr), "col_quantile: pre_sub and pre_div go together");
    SSAR_REQUIRE(n_clips <= 65535, "col_quantile: too many clips per launch");
    const int threads = n_frames >= 4096 ? 512 : 256;
    // columns of up to 11 776 frames are staged (as order-preserving keys) in shared memory once; the ~9 radix passes
    // then never touch the strided column again
    const int staged = n_frames <= 11776 ? 1 : 0;  // 46 KB + the static arrays stay under the 48 KB default
    col_quantile_kernel<<<dim3(n_cols, n_clips), threads, staged ? (size_t)n_frames * sizeof(uint32_t) : 0, ST(stream)>>>(
        x, n_frames, n_cols, subset, q, interp, pre_sub, pre_div, out, count_out, staged, clip_frames);
    return ssar_launch_status("col_quantile");
}

static inline dim3 flat_grid(long long n_per_clip, int n_clips) {
    int gx = ssar_div_up(n_per_clip, 256 * 4);
    if (gx > SSAR_NUM_SMS * 8) gx = SSAR_NUM_SMS * 8;
    if (gx < 1) gx = 1;
    return dim3(gx, n_clips);
}

extern "C" int ssar_clamp_cols(const float* x, int n_clips, int n_frames, int n_cols, const float* lo, const float* hi,
                               float* out, void* stream) {
    if (n_clips == 0 || n_frames == 0 || n_cols == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out, "clamp_cols: null pointer");
    const long long n = (long long)n_frames * n_cols;
    clamp_cols_kernel<<<flat_grid(n, n_clips), 256, 0, ST(stream)>>>(x, n, n_cols, lo, hi, out);
    return ssar_launch_status("clamp_cols");
}

extern "C" int ssar_normalize(const float* x, int n_clips, long long n_per_clip, const uint32_t* min_key,
                              const uint32_t* max_key, float eps, float* out, void* stream) {
    if (n_clips == 0 || n_per_clip == 0) return SSAR_OK;
    SSAR_REQUIRE(x && min_key && max_key && out, "normalize: null pointer");
    normalize_kernel<<<flat_grid(n_per_clip, n_clips), 256, 0, ST(stream)>>>(x, n_per_clip, min_key, max_key, eps, out);
    return ssar_launch_status("normalize");
}

extern "C" int ssar_normalize_fused(const float* x, int n_clips, long long n_per_clip, float eps, float* out,
                                    const int* clip_rows, long long row_elems, void* stream) {
    if (n_clips == 0 || n_per_clip == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out, "normalize_fused: null pointer");
    SSAR_REQUIRE(n_per_clip <= (1LL << 18), "normalize_fused: at most 262 144 elements per clip (use minmax_keys + normalize)");
    SSAR_REQUIRE(!clip_rows || row_elems > 0, "normalize_fused: row_elems must be positive with clip_rows");
    normalize_fused_kernel<<<n_clips, 1024, 0, ST(stream)>>>(x, n_per_clip, eps, out, clip_rows, row_elems);
    return ssar_launch_status("normalize_fused");
}

extern "C" int ssar_emphasize(const float* x, int n_clips, int n_frames, int n_cols, const float* col_min,
                              const float* col_max, float strength, double q, float* scratch, float* out,
                              const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_frames == 0 || n_cols == 0) return SSAR_OK;
    SSAR_REQUIRE(x && col_min && col_max && scratch && out, "emphasize: null pointer");
    const int nc = n_clips * n_cols;
    float* rng = scratch;        // [n_clips * n_cols]
    float* qv = scratch + nc;    // [n_clips * n_cols]
    range_kernel<<<ssar_div_up(nc, 256), 256, 0, ST(stream)>>>(col_min, col_max, nc, rng);
    int rc = ssar_col_quantile(x, n_clips, n_frames, n_cols, 0, q, 0, col_min, rng, qv, nullptr, clip_frames, stream);
    if (rc != SSAR_OK) return rc;
    const long long n = (long long)n_frames * n_cols;
    emphasize_kernel<<<flat_grid(n, n_clips), 256, 0, ST(stream)>>>(x, n, n_cols, col_min, rng, qv, strength, out);
    return ssar_launch_status("emphasize");
}

extern "C" int ssar_spectral_flux(const float* x, int n_clips, int n_frames, long long inner, float* out, void* stream) {
    if (n_clips == 0 || n_frames == 0 || inner == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out && x != out, "spectral_flux: null pointer or aliasing");
    flux_kernel<<<flat_grid((long long)n_frames * inner, n_clips), 256, 0, ST(stream)>>>(x, n_frames, inner, out);
    return ssar_launch_status("spectral_flux");
}

extern "C" int ssar_hwr_rowsum(const float* x, long long n_rows, int inner, float* out, void* stream) {
    if (n_rows == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out && inner > 0, "hwr_rowsum: null pointer");
    int grid = ssar_div_up(n_rows, 8);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    hwr_rowsum_kernel<<<grid, 256, 0, ST(stream)>>>(x, n_rows, inner, out);
    return ssar_launch_status("hwr_rowsum");
}

extern "C" int ssar_feature_stats(const float* x, long long n_rows, int n_cols, double* stats, int frames_per_clip,
                                  const int* clip_frames, void* stream) {
    if (n_rows == 0 || n_cols == 0) return SSAR_OK;
    SSAR_REQUIRE(x && stats, "feature_stats: null pointer");
    int gy = ssar_div_up(n_rows, 8 * 64);
    if (gy > SSAR_NUM_SMS * 2) gy = SSAR_NUM_SMS * 2;
    if (gy < 1) gy = 1;
    SSAR_REQUIRE(!clip_frames || (frames_per_clip > 0 && n_rows % frames_per_clip == 0),
                 "feature_stats: ragged batches need rows = n_clips * frames_per_clip");
    feature_stats_kernel<<<dim3(ssar_div_up(n_cols, 32), gy), dim3(32, 8), 0, ST(stream)>>>(x, n_rows, n_cols, stats,
                                                                                       frames_per_clip, clip_frames);
    return ssar_launch_status("feature_stats");
}
