// Tuning estimation for the constant-Q basis (rosa/pitch.py:9-120: estimate_tuning = piptrack + pitch_tuning).
//
//   piptrack_kernel : per STFT frame (magnitude spectrum, n_fft 2048, hop n_fft/4): column maximum, threshold,
//                     local maxima inside [fmin, fmax), parabolic peak interpolation -> (pitch, mag) pairs appended to a
//                     per-clip list (warp-aggregated atomic append; the statistics that follow are order-independent,
//                     so the result is deterministic).  Capacity per clip: n_frames * ((R + 1) / 2) pairs, R = bins in
//                     the feasible range (two local maxima cannot be adjacent).
//   tuning_kernel   : one CTA per clip: lower median of the peak magnitudes (radix select over the slots with
//                     pitch > 0), then the 100-bin histogram of the 1/36-octave residuals of the pitches whose
//                     magnitude reaches the median, arg-max -> tuning index in [0, 100] (value = edges[index]).
#include "block_utils.cuh"
#include "select.cuh"

namespace {

constexpr int PIP_WARPS = 8;

__global__ void __launch_bounds__(PIP_WARPS * 32) piptrack_kernel(const float* __restrict__ S, long long ld, long long n_rows,
                                                                int n_frames, int n_bins, int k_lo, int k_hi,
                                                                float threshold, float sr_f, float n_fft_f,
                                                                long long capacity, float* __restrict__ pitch,
                                                                float* __restrict__ mag, int* __restrict__ count) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * PIP_WARPS + w; row < n_rows; row += (long long)gridDim.x * PIP_WARPS) {
        const float* s = S + row * ld;
        const int clip = (int)(row / n_frames);
        float mx = -INFINITY;
        for (int k = l; k < n_bins; k += 32) mx = fmaxf(mx, __ldg(s + k));
        mx = warp_max(mx);
        const float ref = threshold * mx;  // rosa/pitch.py:69
        for (int kb = k_lo; kb < k_hi; kb += 32) {
            const int k = kb + l;
            float p = 0.f, m = 0.f;
            bool is_peak = false;
            if (k < k_hi) {
                const float c = __ldg(s + k);
                const float prev = k > 0 ? __ldg(s + k - 1) : 0.f;
                const float next = k + 1 < n_bins ? __ldg(s + k + 1) : 0.f;
                // localmax of S * (S > ref) along frequency with zero padding (rosa/pitch.py:71, 86-95)
                const float cm = c > ref ? c : 0.f;
                const float pm = (k > 0 && prev > ref) ? prev : 0.f;
                const float nm = (k + 1 < n_bins && next > ref) ? next : 0.f;
                if (cm > pm && cm >= nm) {
                    float avg = 0.f, shift = 0.f;
                    if (k > 0 && k + 1 < n_bins) {  // the parabola terms are zero-padded at the first / last bin (:57-58)
                        avg = 0.5f * (next - prev);
                        const float den = 2.f * c - next - prev;
                        shift = avg / (den + (fabsf(den) < 1.17549435e-38f ? 1.f : 0.f));
                    }
                    p = ((float)k + shift) * sr_f / n_fft_f;
                    m = c + 0.5f * avg * shift;
                    is_peak = p > 0.f;  // estimate_tuning only ever looks at pitch > 0 (rosa/pitch.py:12-21)
                }
            }
            const unsigned ballot = __ballot_sync(0xffffffffu, is_peak);
            if (ballot) {
                int base = 0;
                if (l == 0) base = atomicAdd(count + clip, __popc(ballot));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (is_peak) {
                    const long long slot = base + __popc(ballot & ((1u << l) - 1u));
                    if (slot < capacity) {
                        pitch[clip * capacity + slot] = p;
                        mag[clip * capacity + slot] = m;
                    }
                }
            }
        }
    }
}

__global__ void __launch_bounds__(1024) tuning_kernel(const float* __restrict__ pitch, const float* __restrict__ mag,
                                                      long long capacity, const int* __restrict__ count,
                                                      const float* __restrict__ edges, int n_hist,
                                                      float bins_per_octave, int* __restrict__ tuning_idx,
                                                      float* __restrict__ tuning_val) {
    __shared__ uint32_t hist[256];
    __shared__ uint32_t bcast[2];
    __shared__ unsigned long long s_count;
    __shared__ int s_any;
    const int clip = blockIdx.x;
    const long long n_per_clip = min((long long)count[clip], capacity);
    SelectCtx c;
    c.xc = mag + (long long)clip * capacity;
    c.aux = pitch + (long long)clip * capacity;
    c.T = n_per_clip;
    c.C = 1;
    c.subset = 2;
    c.pre = false;
    c.sub = 0.f;
    c.div = 1.f;
    if (threadIdx.x == 0) {
        s_count = 0ull;
        s_any = 0;
    }
    __syncthreads();
    unsigned long long local = 0;
    for (long long t = threadIdx.x; t < n_per_clip; t += blockDim.x) {
        float v;
        local += sel_value(c, t, v) ? 1 : 0;
    }
    if (local) atomicAdd(&s_count, local);
    __syncthreads();
    const long long n = (long long)s_count;
    // threshold = torch.median(mag[pitch > 0]) (lower median) or 0 when nothing was tracked (rosa/pitch.py:12-17)
    float thr = 0.f;
    if (n > 0) thr = key2f(radix_select(c, (n - 1) / 2, hist, bcast));
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const float lo = edges[0], hi = edges[n_hist];
    for (long long t = threadIdx.x; t < n_per_clip; t += blockDim.x) {
        const float f = c.aux[t];
        if (!(f > 0.f) || !(c.xc[t] >= thr)) continue;
        // rosa/pitch.py:108-113: residual of bins_per_octave * log2(f / (440 / 16)) modulo 1, folded to [-0.5, 0.5)
        float r = bins_per_octave * log2f(f / 27.5f);
        r = r - floorf(r);
        if (r >= 0.5f) r -= 1.f;
        if (r < lo || r > hi) continue;
        // torch.histc (ATen HistogramKernel, LINEAR_INTERPOLATION without the local search torch.histogram adds):
        // pos = (int)((r - lo) * n / (hi - lo)) in fp32, the right edge folded into the last bin.  Checked against
        // torch.histc on 2e5 values within a quarter ulp of the edges (identical counts; the local search moved 8 %).
        int pos = (int)((r - lo) * (float)n_hist / (hi - lo));
        pos = min(max(pos, 0), n_hist - 1);
        atomicAdd(&hist[pos], 1u);
        s_any = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int best = 0;
        if (s_any) {
            for (int i = 1; i < n_hist; ++i)
                if (hist[i] > hist[best]) best = i;  // first maximum wins (torch.argmax)
            tuning_idx[clip] = best;
            tuning_val[clip] = edges[best];
        } else {  // pitch_tuning returns 0.0 on an empty set (rosa/pitch.py:104-105): index of the 0.0 edge
            tuning_idx[clip] = n_hist / 2;
            tuning_val[clip] = 0.f;
        }
    }
}

}  // namespace

extern "C" long long ssar_piptrack_capacity(int n_frames, int k_lo, int k_hi) {
    return (long long)n_frames * ((k_hi - k_lo + 1) / 2);
}

extern "C" int ssar_piptrack(const float* mag_spec, long long ld, int n_clips, int n_frames, int n_bins, int k_lo,
                             int k_hi, float threshold, float sr, int n_fft, long long capacity, float* pitch, float* mag,
                             int* count, void* stream) {
    const long long rows = (long long)n_clips * n_frames;
    if (rows == 0) return SSAR_OK;
    SSAR_REQUIRE(mag_spec && pitch && mag && count, "piptrack: null pointer");
    SSAR_REQUIRE(0 <= k_lo && k_lo < k_hi && k_hi <= n_bins, "piptrack: bad bin range");
    SSAR_REQUIRE(capacity >= ssar_piptrack_capacity(n_frames, k_lo, k_hi), "piptrack: capacity below ssar_piptrack_capacity");
    cudaError_t e = cudaMemsetAsync(count, 0, sizeof(int) * n_clips, (cudaStream_t)stream);
    if (e != cudaSuccess) {
        ssar_set_error("piptrack: memset: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    long long grid = (rows + PIP_WARPS - 1) / PIP_WARPS;
    if (grid > (long long)SSAR_NUM_SMS * 8) grid = (long long)SSAR_NUM_SMS * 8;
    piptrack_kernel<<<(unsigned)grid, PIP_WARPS * 32, 0, (cudaStream_t)stream>>>(
        mag_spec, ld, rows, n_frames, n_bins, k_lo, k_hi, threshold, sr, (float)n_fft, capacity, pitch, mag, count);
    return ssar_launch_status("piptrack");
}

extern "C" int ssar_tuning_from_peaks(const float* pitch, const float* mag, int n_clips, long long capacity,
                                      const int* count, const float* edges, int n_hist, float bins_per_octave,
                                      int* tuning_idx, float* tuning_val, void* stream) {
    if (n_clips == 0) return SSAR_OK;
    SSAR_REQUIRE(pitch && mag && count && edges && tuning_idx && tuning_val, "tuning_from_peaks: null pointer");
    SSAR_REQUIRE(n_hist > 0 && n_hist <= 256, "tuning_from_peaks: histogram size must be in [1, 256]");
    tuning_kernel<<<n_clips, 1024, 0, (cudaStream_t)stream>>>(pitch, mag, capacity, count, edges, n_hist, bins_per_octave,
                                                             tuning_idx, tuning_val);
    return ssar_launch_status("tuning_from_peaks");
}
