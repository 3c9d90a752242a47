// Framed DFT / inverse for ARBITRARY n_fft (not a power of two), restricted to a band of bins.
//
// The radix kernels (frames_fft.cu, istft.cu) serve n_fft 1024 / 2048, which is everything the path uses EXCEPT the
// Fourier tempogram of a clip shorter than 1024 frames: rosa/beat.py:46-49 takes win_length = min(len(onset), 1024), so a
// 16 s clip (the reference generators' own length, features/generate.py:51-54) transforms T + 1 frames of n_fft = T = 384
// at hop 1, and only the 60-180 BPM bins [T/24, T/8] survive the mask (beat.py:54-57).  That is (T+1) x T/12 x T
// complex multiply-adds (< 0.1 G for T = 1008): a direct DFT over the band, with an exact twiddle index (k m) mod n_fft
// into a shared-memory table, is both simpler and more accurate than Bluestein.
//
//   ssar_frames_dft : torch.stft(center=True, reflect) semantics, bins [k_lo, k_hi) -> complex [clip, frame, k_hi - k_lo]
//   ssar_istft_band : torch.istft(center=True, length) semantics from the same band (all other bins zero)
#include "common.cuh"

namespace {

constexpr int DFT_THREADS = 256;
constexpr int DFT_MAX_NFFT = 4096;
constexpr int DFT_SPAN = 8192;  // floats of (reflect-padded) signal staged per CTA

// tw[r] = exp(-2 pi i r / n), r in [0, n): evaluated in double so the table is correctly rounded
__device__ __forceinline__ void fill_twiddle_table(float2* tw, int n) {
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        double s, c;
        sincospi(2.0 * (double)r / (double)n, &s, &c);
        tw[r] = make_float2((float)c, (float)-s);
    }
}

__global__ void __launch_bounds__(DFT_THREADS) frames_dft_kernel(const float* __restrict__ x, long long n_samples,
                                                                  long long clip_stride, int n_fft, int hop, int n_frames,
                                                                  const float* __restrict__ window, int k_lo, int k_hi,
                                                                  int frames_per_cta, float2* __restrict__ out) {
    extern __shared__ float4 dft_smem_raw[];
    float2* tw = reinterpret_cast<float2*>(dft_smem_raw);  // [n_fft]
    float* sig = reinterpret_cast<float*>(tw + n_fft);     // [(frames_per_cta - 1) * hop + n_fft], window NOT applied
    float* win = sig + DFT_SPAN;                           // [n_fft]
    const int clip = blockIdx.y;
    const int t0 = blockIdx.x * frames_per_cta;
    const int nf = min(frames_per_cta, n_frames - t0);
    const int nb = k_hi - k_lo;
    const float* __restrict__ xc = x + (long long)clip * clip_stride;

    fill_twiddle_table(tw, n_fft);
    for (int i = threadIdx.x; i < n_fft; i += blockDim.x) win[i] = window ? __ldg(window + i) : 1.f;
    const int span = (nf - 1) * hop + n_fft;
    const long long first = (long long)t0 * hop - n_fft / 2;  // sample index of sig[0] (reflect outside [0, n_samples))
    for (int i = threadIdx.x; i < span; i += blockDim.x) sig[i] = __ldg(xc + ssar_reflect(first + i, n_samples));
    __syncthreads();

    for (int item = threadIdx.x; item < nf * nb; item += blockDim.x) {
        const int f = item / nb, k = k_lo + (item - f * nb);
        const float* s = sig + f * hop;
        float ar = 0.f, ai = 0.f, br = 0.f, bi = 0.f;  // two accumulator pairs: even / odd m
        int r = 0;                                      // (k * m) mod n_fft
        const int kk = k % n_fft;
        int m = 0;
        for (; m + 1 < n_fft; m += 2) {
            const float v0 = s[m] * win[m];
            const float2 w0 = tw[r];
            r += kk;
            if (r >= n_fft) r -= n_fft;
            const float v1 = s[m + 1] * win[m + 1];
            const float2 w1 = tw[r];
            r += kk;
            if (r >= n_fft) r -= n_fft;
            ar = fmaf(v0, w0.x, ar);
            ai = fmaf(v0, w0.y, ai);
            br = fmaf(v1, w1.x, br);
            bi = fmaf(v1, w1.y, bi);
        }
        if (m < n_fft) {
            const float v0 = s[m] * win[m];
            const float2 w0 = tw[r];
            ar = fmaf(v0, w0.x, ar);
            ai = fmaf(v0, w0.y, ai);
        }
        out[((long long)clip * n_frames + t0 + f) * nb + (k - k_lo)] = make_float2(ar + br, ai + bi);
    }
}

// One thread per output sample: sum over the frames that cover it of window[m] * irfft(frame)[m], divided by the summed
// squared window (torch.istft).  irfft of a one-sided spectrum whose only non-zero bins are [k_lo, k_hi):
//   y[m] = (1 / n) sum_k c_k (Re X_k cos(2 pi k m / n) - Im X_k sin(2 pi k m / n)),  c_k = 1 for k = 0 and k = n / 2 (n even:
//   the imaginary parts of those two bins are ignored by a C2R transform), else 2.
__global__ void __launch_bounds__(DFT_THREADS) istft_band_kernel(const float2* __restrict__ spec, int n_frames, int n_fft,
                                                                  int hop, const float* __restrict__ window, int k_lo,
                                                                  int k_hi, long long length, float* __restrict__ out) {
    extern __shared__ float4 dft_smem_raw[];
    float2* tw = reinterpret_cast<float2*>(dft_smem_raw);  // [n_fft]
    float* win = reinterpret_cast<float*>(tw + n_fft);     // [n_fft]
    fill_twiddle_table(tw, n_fft);
    for (int i = threadIdx.x; i < n_fft; i += blockDim.x) win[i] = window ? __ldg(window + i) : 1.f;
    __syncthreads();

    const int clip = blockIdx.y, nb = k_hi - k_lo;
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= length) return;
    const long long p = j + n_fft / 2;  // position in the un-trimmed overlap-add buffer
    long long t_lo = (p - n_fft + hop) / hop;  // ceil((p - n_fft + 1) / hop)
    if (p - n_fft + 1 <= 0) t_lo = 0;
    long long t_hi = p / hop;
    if (t_hi > n_frames - 1) t_hi = n_frames - 1;
    const float2* __restrict__ sc = spec + (long long)clip * n_frames * nb;
    const bool even = (n_fft & 1) == 0;
    float acc = 0.f, env = 0.f;
    for (long long t = t_lo; t <= t_hi; ++t) {
        const int m = (int)(p - t * hop);
        const float wv = win[m];
        env = fmaf(wv, wv, env);
        const float2* __restrict__ row = sc + t * nb;
        int r = (int)(((long long)k_lo * m) % n_fft);
        float y = 0.f;
        for (int k = k_lo; k < k_hi; ++k) {
            const float2 X = __ldg(row + (k - k_lo));
            if (X.x != 0.f || X.y != 0.f) {
                const float2 w = tw[r];  // (cos, -sin)
                if (k == 0 || (even && 2 * k == n_fft))
                    y = fmaf(X.x, w.x, y);
                else
                    y += 2.f * fmaf(X.x, w.x, X.y * w.y);
            }
            r += m;
            if (r >= n_fft) r -= n_fft;
        }
        acc = fmaf(wv, y / (float)n_fft, acc);
    }
    out[(long long)clip * length + j] = (t_hi >= t_lo && env > 1e-11f) ? acc / env : 0.f;
}

}  // namespace

extern "C" int ssar_frames_dft(const float* x, long long n_samples, long long clip_stride, int n_clips, int n_fft, int hop,
                               int n_frames, const float* window, int k_lo, int k_hi, float* out_complex, void* stream) {
    if (n_clips <= 0 || n_frames <= 0 || k_hi <= k_lo) return SSAR_OK;
    SSAR_REQUIRE(x && out_complex, "frames_dft: null pointer");
    SSAR_REQUIRE(n_fft >= 2 && n_fft <= DFT_MAX_NFFT, "frames_dft: n_fft=%d outside [2, %d]", n_fft, DFT_MAX_NFFT);
    SSAR_REQUIRE(hop >= 1 && hop <= DFT_SPAN - DFT_MAX_NFFT, "frames_dft: hop=%d out of range", hop);
    SSAR_REQUIRE(0 <= k_lo && k_hi <= n_fft / 2 + 1, "frames_dft: bins [%d, %d) outside the one-sided spectrum", k_lo, k_hi);
    SSAR_REQUIRE(n_samples > n_fft / 2, "frames_dft: clip shorter than n_fft/2+1 samples cannot be reflect-padded");
    SSAR_REQUIRE((long long)(n_frames - 1) * hop <= n_samples + 2 * (long long)(n_fft / 2) - n_fft,
                 "frames_dft: n_frames=%d exceeds the padded clip", n_frames);
    int fpc = (DFT_SPAN - n_fft) / hop + 1;
    if (fpc > 32) fpc = 32;
    if (fpc > n_frames) fpc = n_frames;
    const size_t smem = (size_t)n_fft * sizeof(float2) + (size_t)DFT_SPAN * sizeof(float) + (size_t)n_fft * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(frames_dft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("frames_dft: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    dim3 grid((unsigned)ssar_div_up(n_frames, fpc), (unsigned)n_clips);
    frames_dft_kernel<<<grid, DFT_THREADS, smem, (cudaStream_t)stream>>>(x, n_samples, clip_stride, n_fft, hop, n_frames, window,
                                                                         k_lo, k_hi, fpc, reinterpret_cast<float2*>(out_complex));
    return ssar_launch_status("frames_dft");
}

extern "C" int ssar_istft_band(const float* spec, int n_clips, int n_frames, int n_fft, int hop, const float* window, int k_lo,
                               int k_hi, long long length, float* out, void* stream) {
    if (n_clips <= 0 || length <= 0) return SSAR_OK;
    SSAR_REQUIRE(spec && out, "istft_band: null pointer");
    SSAR_REQUIRE(n_fft >= 2 && n_fft <= DFT_MAX_NFFT, "istft_band: n_fft=%d outside [2, %d]", n_fft, DFT_MAX_NFFT);
    SSAR_REQUIRE(hop >= 1 && n_frames >= 1, "istft_band: bad hop / n_frames");
    SSAR_REQUIRE(0 <= k_lo && k_lo <= k_hi && k_hi <= n_fft / 2 + 1, "istft_band: bins [%d, %d) outside the one-sided spectrum",
                 k_lo, k_hi);
    const size_t smem = (size_t)n_fft * (sizeof(float2) + sizeof(float));
    cudaError_t e = cudaFuncSetAttribute(istft_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("istft_band: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    dim3 grid((unsigned)ssar_div_up(length, DFT_THREADS), (unsigned)n_clips);
    istft_band_kernel<<<grid, DFT_THREADS, smem, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(spec), n_frames, n_fft,
                                                                         hop, window, k_lo, k_hi, length, out);
    return ssar_launch_status("istft_band");
}
