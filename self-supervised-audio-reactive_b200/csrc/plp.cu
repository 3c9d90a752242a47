// Predominant local pulse, spectral step (rosa/beat.py:54-66): on the Fourier tempogram [clip, frame, bin] (complex,
// frame-major) keep, per frame, only the bin(s) inside the tempo range whose log1p(1e6 |X|) equals the frame maximum,
// and scale them to unit magnitude: X / (sqrt(tiny) + max |X|).  In place.  One warp per frame.
#include "common.cuh"

namespace {

constexpr int PLP_WARPS = 8;

__global__ void __launch_bounds__(PLP_WARPS * 32) plp_peaks_kernel(float2* __restrict__ X, long long ld, long long n_rows,
                                                                 int n_bins, int k_lo, int k_hi) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    for (long long row = (long long)blockIdx.x * PLP_WARPS + w; row < n_rows; row += (long long)gridDim.x * PLP_WARPS) {
        float2* x = X + row * ld;
        // ftmag peak over ALL bins of the masked tempogram (bins outside the range are zero -> log1p(0) = 0)
        float peak = 0.f;
        for (int k = k_lo + l; k < k_hi; k += 32) {
            const float2 z = x[k];
            peak = fmaxf(peak, log1pf(1e6f * sqrtf(fmaf(z.x, z.x, z.y * z.y))));
        }
        peak = warp_max(peak);
        float amax = 0.f;
        for (int k = k_lo + l; k < k_hi; k += 32) {
            const float2 z = x[k];
            const float a = sqrtf(fmaf(z.x, z.x, z.y * z.y));
            if (!(log1pf(1e6f * a) < peak)) amax = fmaxf(amax, a);
        }
        amax = warp_max(amax);
        const float den = 1.0842021724855044e-19f + amax;  // finfo(float32).tiny ** 0.5
        for (int k = l; k < n_bins; k += 32) {
            float2 z = make_float2(0.f, 0.f);
            if (k >= k_lo && k < k_hi) {
                const float2 v = x[k];
                const float a = sqrtf(fmaf(v.x, v.x, v.y * v.y));
                if (!(log1pf(1e6f * a) < peak)) z = make_float2(v.x / den, v.y / den);
            }
            x[k] = z;
        }
    }
}

}  // namespace

extern "C" int ssar_plp_peaks(float* tempogram, long long ld, int n_clips, int n_frames, int n_bins, int k_lo, int k_hi,
                              void* stream) {
    const long long rows = (long long)n_clips * n_frames;
    if (rows == 0) return SSAR_OK;
    SSAR_REQUIRE(tempogram, "plp_peaks: null pointer");
    SSAR_REQUIRE(0 <= k_lo && k_lo <= k_hi && k_hi <= n_bins, "plp_peaks: bad tempo bin range");
    long long grid = (rows + PLP_WARPS - 1) / PLP_WARPS;
    if (grid > (long long)SSAR_NUM_SMS * 8) grid = (long long)SSAR_NUM_SMS * 8;
    plp_peaks_kernel<<<(unsigned)grid, PLP_WARPS * 32, 0, (cudaStream_t)stream>>>(reinterpret_cast<float2*>(tempogram), ld,
                                                                                rows, n_bins, k_lo, k_hi);
    return ssar_launch_status("plp_peaks");
}
