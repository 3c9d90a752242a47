// Inverse STFT = per-frame inverse real FFT * window  ->  overlap-add / window-envelope normalisation.
// Replaces torch.istft behind rosa/spectral.py:24-32 (used by features/audio.py:13-24 harmonic/percussive with
// hop = n_fft/2 and by rosa/beat.py:69 plp with hop = 1).
//
// Kernel 1 (irfft_frames): same shared-memory Stockham core as the forward kernel.  The one-sided spectrum is
//   re-packed into an M = n_fft/2 point complex transform (Z = E + iO), run through the FORWARD passes on conj(Z)
//   (inverse = conj(FFT(conj(.)))/M), windowed and stored as [clip, frame, n_fft].
// Kernel 2 (overlap_add): gather formulation, one thread per output sample: sum the <= ceil(n_fft/hop) frames
//   that cover it, divide by the summed squared window (torch.istft's envelope), center-trim by n_fft/2.
#include "fft_core.cuh"

namespace {

template <int M>
__global__ void __launch_bounds__(256, 2) irfft_frames_kernel(const float2* __restrict__ X, long long ldx,
                                                              const float* __restrict__ win,
                                                              const float2* __restrict__ tw, long long total_frames,
                                                              float* __restrict__ out) {
    using S = FftShape<M>;
    constexpr int TPF = S::TPF, FPB = 256 / TPF, N = 2 * M;
    constexpr int ZLEN = fft_padded_len(M);
    extern __shared__ float4 smem_raw[];
    const int g = threadIdx.x / TPF, lt = threadIdx.x % TPF;
    float2* s = reinterpret_cast<float2*>(smem_raw) + g * ZLEN;
    const float inv_m = 1.0f / (float)M;

    for (long long base = (long long)blockIdx.x * FPB; base < total_frames; base += (long long)gridDim.x * FPB) {
        const long long f = base + g;
        const bool active = f < total_frames;
        const float2* Xf = X + (active ? f : 0) * ldx;
        // ---- inverse untangle: conj(Z[k]) with Z[k] = E + iO, E = (X[k] + conj(X[M-k]))/2,
        //      O = conj(w^k) (X[k] - conj(X[M-k]))/2 ; imaginary parts of DC and Nyquist are ignored (C2R semantics)
        for (int k = lt; k <= M / 2; k += TPF) {
            float2 xk = active ? __ldg(&Xf[k]) : make_float2(0.f, 0.f);
            float2 xm = active ? __ldg(&Xf[M - k]) : make_float2(0.f, 0.f);
            if (k == 0) {
                xk.y = 0.f;
                xm.y = 0.f;
            }
            const float2 w = __ldg(&tw[k]);
            const float2 e = make_float2(0.5f * (xk.x + xm.x), 0.5f * (xk.y - xm.y));
            const float2 d = make_float2(0.5f * (xk.x - xm.x), 0.5f * (xk.y + xm.y));
            const float2 o = cmul(cconj(w), d);
            // Z[k] = e + i o = (e.x - o.y, e.y + o.x);  Z[M-k] = conj(e) + i conj(o) = (e.x + o.y, -e.y + o.x)
            s[PADI(k)] = make_float2(e.x - o.y, -(e.y + o.x));           // conj(Z[k])
            if (k > 0) s[PADI(M - k)] = make_float2(e.x + o.y, e.y - o.x);  // conj(Z[M-k])
        }
        __syncthreads();
        stockham_pass<M, S::R1, TPF>(s, tw, lt, 1);
        fft_tail_passes<M>(s, tw, lt);
        // ---- x[2n] = Re(conj(Y[n]))/M, x[2n+1] = Im(conj(Y[n]))/M, times the synthesis window
        if (active) {
            float2* of = reinterpret_cast<float2*>(out + f * N);
            for (int n = lt; n < M; n += TPF) {
                const float2 y = s[PADI(n)];
                float a = y.x * inv_m, b = -y.y * inv_m;
                if (win) {
                    const float2 w = __ldg(reinterpret_cast<const float2*>(win) + n);
                    a *= w.x;
                    b *= w.y;
                }
                of[n] = make_float2(a, b);
            }
        }
        __syncthreads();
    }
}

__global__ void overlap_add_kernel(const float* __restrict__ frames, const float* __restrict__ win, int n_fft, int hop,
                                   int n_frames, long long length, float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float* fc = frames + (long long)clip * n_frames * n_fft;
    float* oc = out + (long long)clip * length;
    const long long total = (long long)n_fft + (long long)hop * (n_frames - 1);
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < length; n += (long long)gridDim.x * blockDim.x) {
        const long long p = n + n_fft / 2;
        float acc = 0.f, env = 0.f;
        if (p < total) {
            long long t_hi = p / hop;
            if (t_hi > n_frames - 1) t_hi = n_frames - 1;
            long long t_lo = (p - n_fft + hop) / hop;  // ceil((p - n_fft + 1)/hop) for p - n_fft + 1 > 0
            if (p - n_fft + 1 <= 0) t_lo = 0;
            for (long long t = t_lo; t <= t_hi; ++t) {
                const int i = (int)(p - t * hop);
                acc += __ldg(fc + t * n_fft + i);
                const float w = win ? __ldg(win + i) : 1.f;
                env = fmaf(w, w, env);
            }
        }
        oc[n] = env > 1e-11f ? acc / env : 0.f;
    }
}

template <int M>
int launch_irfft(const float* spec, long long ld, const float* win, const float* tw, long long total, float* frames,
                 cudaStream_t st) {
    constexpr int FPB = 256 / FftShape<M>::TPF;
    const size_t smem = (size_t)FPB * fft_padded_len(M) * sizeof(float2);
    cudaError_t e = cudaFuncSetAttribute(irfft_frames_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("irfft_frames: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    long long grid = (total + FPB - 1) / FPB;
    if (grid > (long long)SSAR_NUM_SMS * 4) grid = (long long)SSAR_NUM_SMS * 4;
    irfft_frames_kernel<M><<<(unsigned)grid, 256, smem, st>>>(reinterpret_cast<const float2*>(spec), ld, win,
                                                             reinterpret_cast<const float2*>(tw), total, frames);
    return ssar_launch_status("irfft_frames");
}

}  // namespace

extern "C" int ssar_irfft_frames(const float* spec, long long ld, int n_clips, int n_frames, int n_fft,
                                 const float* window, const float* twiddle, float* frames, void* stream) {
    const long long total = (long long)n_clips * n_frames;
    if (total == 0) return SSAR_OK;
    SSAR_REQUIRE(spec && twiddle && frames, "irfft_frames: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (n_fft == 2048) return launch_irfft<1024>(spec, ld, window, twiddle, total, frames, st);
    if (n_fft == 1024) return launch_irfft<512>(spec, ld, window, twiddle, total, frames, st);
    ssar_set_error("irfft_frames: n_fft=%d unsupported by the fast path (1024 or 2048)", n_fft);
    return SSAR_ERR_UNSUPPORTED;
}

extern "C" int ssar_overlap_add(const float* frames, const float* window, int n_clips, int n_frames, int n_fft, int hop,
                                long long length, float* out, void* stream) {
    if (n_clips == 0 || length == 0) return SSAR_OK;
    SSAR_REQUIRE(frames && out, "overlap_add: null pointer");
    SSAR_REQUIRE(hop > 0 && n_fft > 0 && n_frames > 0, "overlap_add: bad sizes");
    int gx = ssar_div_up(length, 256);
    if (gx > SSAR_NUM_SMS * 16) gx = SSAR_NUM_SMS * 16;
    overlap_add_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(frames, window, n_fft, hop, n_frames, length, out);
    return ssar_launch_status("overlap_add");
}
