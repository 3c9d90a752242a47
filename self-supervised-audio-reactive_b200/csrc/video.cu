// Video-side feature twins (ssar/features/video.py:11-31, 58-75): per-frame colour histograms, visual variance and the
// frame-to-frame absolute difference.  All three are single passes (two for the histogram: its range is the frame's own
// min / max) over [T, C, H, W] video tensors: HBM-bound, 4 B per pixel read.
//   video_hist_kernel     : torch.histc(frame, bins) with the default range (the frame's min / max; a constant frame
//                           widens it by 1 on both sides) -> counts / max count          (redogram / greenogram / blueogram)
//   video_variance_kernel : std over (C, H, W) (unbiased, two-pass) squared                (visual_variance)
//   video_absdiff_kernel  : sum |x[t + 1] - x[t]| over (C, H, W); the last frame repeats its predecessor (absdiff)
#include "block_utils.cuh"
#include "common.cuh"

namespace {

constexpr int VH_MAX_BINS = 256;

__global__ void __launch_bounds__(512) video_hist_kernel(const float* __restrict__ video, long long frame_stride,
                                                         long long channel_offset, int n_pix, int bins,
                                                         float* __restrict__ out, long long out_ld, int out_col0) {
    __shared__ float scratch[32];
    __shared__ unsigned int hist[VH_MAX_BINS];
    __shared__ float s_lo, s_hi;
    __shared__ unsigned int s_max;
    const float* __restrict__ f = video + (long long)blockIdx.x * frame_stride + channel_offset;
    float mn = INFINITY, mx = -INFINITY;
    for (int i = threadIdx.x; i < n_pix; i += blockDim.x) {
        const float v = __ldg(f + i);
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = block_min(mn, scratch);
    __syncthreads();
    mx = block_max(mx, scratch);
    for (int b = threadIdx.x; b < bins; b += blockDim.x) hist[b] = 0u;
    if (threadIdx.x == 0) {
        if (mn == mx) {  // ATen histc: a degenerate range is widened by one on both sides
            mn -= 1.f;
            mx += 1.f;
        }
        s_lo = mn;
        s_hi = mx;
        s_max = 0u;
    }
    __syncthreads();
    const float lo = s_lo, hi = s_hi;
    for (int i = threadIdx.x; i < n_pix; i += blockDim.x) {
        const float v = __ldg(f + i);
        if (v < lo || v > hi) continue;  // (NaN)
        int pos = (int)((v - lo) * (float)bins / (hi - lo));  // ATen's linear binning; the right edge joins the last bin
        if (pos >= bins) pos = bins - 1;
        atomicAdd(&hist[pos], 1u);
    }
    __syncthreads();
    unsigned int m = 0u;
    for (int b = threadIdx.x; b < bins; b += blockDim.x) m = max(m, hist[b]);
    if (m) atomicMax(&s_max, m);
    __syncthreads();
    const float den = (float)s_max;
    for (int b = threadIdx.x; b < bins; b += blockDim.x)
        out[(long long)blockIdx.x * out_ld + out_col0 + b] = (float)hist[b] / den;
}

__global__ void __launch_bounds__(512) video_variance_kernel(const float* __restrict__ video, long long n_per_frame,
                                                             float* __restrict__ out) {
    __shared__ float scratch[32];
    __shared__ float s_b;
    const float* __restrict__ f = video + (long long)blockIdx.x * n_per_frame;
    float s = 0.f;
    for (long long i = threadIdx.x; i < n_per_frame; i += blockDim.x) s += __ldg(f + i);
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) s_b = s / (float)n_per_frame;
    __syncthreads();
    const float mean = s_b;
    float ss = 0.f;
    for (long long i = threadIdx.x; i < n_per_frame; i += blockDim.x) {
        const float d = __ldg(f + i) - mean;
        ss = fmaf(d, d, ss);
    }
    __syncthreads();
    ss = block_sum(ss, scratch);
    if (threadIdx.x == 0) {
        const float sd = sqrtf(ss / (float)(n_per_frame - 1));  // video.std((1, 2, 3)): unbiased
        out[blockIdx.x] = sd * sd;                                // ... ** 2
    }
}

__global__ void __launch_bounds__(512) video_absdiff_kernel(const float* __restrict__ video, long long n_per_frame, int T,
                                                            float* __restrict__ out) {
    __shared__ float scratch[32];
    const int t = blockIdx.x;  // t in [0, T - 1)
    const float* __restrict__ a = video + (long long)t * n_per_frame;
    const float* __restrict__ b = a + n_per_frame;
    float s = 0.f;
    for (long long i = threadIdx.x; i < n_per_frame; i += blockDim.x) s += fabsf(__ldg(b + i) - __ldg(a + i));
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) {
        out[t] = s;
        if (t == T - 2) out[T - 1] = s;  // video.py:71: the last value is repeated
    }
}

}  // namespace

extern "C" int ssar_video_hist(const float* video, int n_frames, int n_channels, int height, int width, int channel,
                               int bins, float* out, long long out_ld, int out_col0, void* stream) {
    if (n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(video && out, "video_hist: null pointer");
    SSAR_REQUIRE(channel >= 0 && channel < n_channels && bins >= 1 && bins <= VH_MAX_BINS && height > 0 && width > 0,
                 "video_hist: bad channel / bins (<= 256) / size");
    SSAR_REQUIRE((long long)height * width < (1LL << 31), "video_hist: frame too large");
    const long long n_pix = (long long)height * width;
    video_hist_kernel<<<n_frames, 512, 0, (cudaStream_t)stream>>>(video, n_pix * n_channels, n_pix * channel, (int)n_pix, bins, out,
                                                                 out_ld, out_col0);
    return ssar_launch_status("video_hist");
}

extern "C" int ssar_video_variance(const float* video, int n_frames, long long n_per_frame, float* out, void* stream) {
    if (n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(video && out && n_per_frame > 1, "video_variance: null pointer or single-element frames");
    video_variance_kernel<<<n_frames, 512, 0, (cudaStream_t)stream>>>(video, n_per_frame, out);
    return ssar_launch_status("video_variance");
}

extern "C" int ssar_video_absdiff(const float* video, int n_frames, long long n_per_frame, float* out, void* stream) {
    SSAR_REQUIRE(video && out && n_frames >= 2 && n_per_frame > 0, "video_absdiff: needs at least two frames");
    video_absdiff_kernel<<<n_frames - 1, 512, 0, (cudaStream_t)stream>>>(video, n_per_frame, n_frames, out);
    return ssar_launch_status("video_absdiff");
}
