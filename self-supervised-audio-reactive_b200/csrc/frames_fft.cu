// Framed, windowed real FFT with fused epilogues (power/magnitude spectrum, complex spectrum, banded real
// filterbanks such as mel, banded complex projections such as the CQT basis, RMS, spectral flatness).
//
// One launch covers every clip of a batch.  A CTA of 256 threads owns FPB = 256/TPF frames at a time and walks
// the (clip, frame) list persistently; consecutive frames of one clip land in the same / neighbouring CTAs, so
// the 4x (hop 512) or 2x (hop 1024) overlapped waveform reads are served by L1/L2 and HBM sees each sample once.
// The spectrum of a frame lives only in shared memory: |X|^2 -> filterbanks happen before anything is written,
// so the [frames, 1025] spectrogram never round-trips HBM unless the caller asks for it as an output.
#include "fft_core.cuh"

namespace {

template <int M>
struct FramesSmem {
    static constexpr int TPF = FftShape<M>::TPF;
    static constexpr int FPB = 256 / TPF;
    static constexpr int ZLEN = fft_padded_len(M);        // float2 per frame group
    static constexpr int PLEN = ((M + 1 + 3) / 4) * 4;    // floats per frame group
    static constexpr size_t bytes = (size_t)FPB * ZLEN * sizeof(float2) + (size_t)FPB * PLEN * sizeof(float) +
                                    (size_t)FPB * 8 * sizeof(float);
};

template <int M>
__global__ void __launch_bounds__(256, 2) frames_fft_kernel(const ssar_frames_args a, const long long total_frames) {
    using L = FramesSmem<M>;
    constexpr int TPF = L::TPF, FPB = L::FPB, N = 2 * M, NBINS = M + 1;
    constexpr int WPF = TPF / 32;  // warps per frame group
    extern __shared__ float4 smem_raw[];
    float2* zall = reinterpret_cast<float2*>(smem_raw);
    float* pall = reinterpret_cast<float*>(zall + FPB * L::ZLEN);
    float* redall = pall + FPB * L::PLEN;

    const int g = threadIdx.x / TPF, lt = threadIdx.x % TPF;
    const int wig = lt >> 5;  // warp index inside the frame group
    float2* s = zall + g * L::ZLEN;
    float* pw = pall + g * L::PLEN;
    float* red = redall + g * 8;
    const float2* __restrict__ tw = reinterpret_cast<const float2*>(a.twiddle);
    const float* __restrict__ win = a.window;
    const bool need_pw = a.out_spec || a.out_flatness || a.fb[0].n_bands > 0 || a.fb[1].n_bands > 0;

    for (long long base = (long long)blockIdx.x * FPB; base < total_frames; base += (long long)gridDim.x * FPB) {
        const long long f = base + g;
        const bool active = f < total_frames;
        const int clip = active ? (int)(f / a.n_frames) : 0;
        const int t = active ? (int)(f - (long long)clip * a.n_frames) : 0;

        // ---- pass 1 (radix 16, Ns = 1): the 16 inputs of butterfly `lt` come straight from global memory
        float2 v[16];
        float ss = 0.f;
        if (active) {
            const float* __restrict__ xc = a.x + (long long)clip * a.clip_stride;
            const long long start = (long long)t * a.hop - M;
            const bool interior = start >= 0 && start + N <= a.n_samples;
            const bool vec = interior && ((((long long)clip * a.clip_stride + start) & 1LL) == 0) &&
                             ((reinterpret_cast<uintptr_t>(a.x) & 7) == 0);
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int n = lt + r * (M / 16);
                float x0, x1;
                if (vec) {
                    const float2 p = __ldg(reinterpret_cast<const float2*>(xc + start) + n);
                    x0 = p.x;
                    x1 = p.y;
                } else if (interior) {
                    x0 = __ldg(xc + start + 2 * n);
                    x1 = __ldg(xc + start + 2 * n + 1);
                } else {
                    x0 = __ldg(xc + ssar_reflect(start + 2 * n, a.n_samples));
                    x1 = __ldg(xc + ssar_reflect(start + 2 * n + 1, a.n_samples));
                }
                ss = fmaf(x0, x0, ss);
                ss = fmaf(x1, x1, ss);
                if (win) {
                    const float2 w = __ldg(reinterpret_cast<const float2*>(win) + n);
                    x0 *= w.x;
                    x1 *= w.y;
                }
                v[r] = make_float2(x0, x1);
            }
        } else {
#pragma unroll
            for (int r = 0; r < 16; ++r) v[r] = make_float2(0.f, 0.f);
        }
        stockham_finish<M, 16>(v, s, tw, lt, 1);
        __syncthreads();
        fft_tail_passes<M>(s, tw, lt);

        // ---- untangle the packed real transform: X[k] = E + w^k O, X[M-k] = conj(E - w^k O)
        for (int k = lt; k <= M / 2; k += TPF) {
            const float2 zk = s[PADI(k)];
            const float2 zm = s[PADI(k == 0 ? 0 : M - k)];
            const float2 w = __ldg(&tw[k]);
            const float2 e = make_float2(0.5f * (zk.x + zm.x), 0.5f * (zk.y - zm.y));
            const float2 d = make_float2(zk.x - zm.x, zk.y + zm.y);
            const float2 o = make_float2(0.5f * d.y, -0.5f * d.x);
            const float2 wo = cmul(w, o);
            s[PADI(k)] = cadd(e, wo);
            s[PADI(M - k)] = cconj(csub(e, wo));
        }
        __syncthreads();

        // ---- epilogue 1: complex / power / magnitude outputs, power tile for the filterbanks, flatness sums
        float slog = 0.f, slin = 0.f;
        {
            float2* oc = (active && a.out_complex)
                             ? reinterpret_cast<float2*>(a.out_complex) + ((long long)clip * a.n_frames + t) * a.ldc
                             : nullptr;
            float* os = (active && a.out_spec) ? a.out_spec + ((long long)clip * a.n_frames + t) * a.lds : nullptr;
            for (int k = lt; k < NBINS; k += TPF) {
                const float2 X = s[PADI(k)];
                if (oc) oc[k] = X;
                if (need_pw) {
                    const float p = fmaf(X.x, X.x, X.y * X.y);
                    pw[k] = p;
                    if (os) os[k] = (a.spec_power == 2) ? p : sqrtf(p);
                    if (a.out_flatness) {
                        const float pt = fmaxf(1e-10f, p);
                        slog += logf(pt);
                        slin += pt;
                    }
                }
            }
        }
        // group reductions (rms, flatness) through a tiny smem scratch
        ss = warp_sum(ss);
        slog = warp_sum(slog);
        slin = warp_sum(slin);
        if ((lt & 31) == 0) {
            red[wig] = ss;
            red[2 + wig] = slog;
            red[4 + wig] = slin;
        }
        __syncthreads();
        if (active && lt == 0) {
            float r0 = red[0], r1 = red[2], r2 = red[4];
            if (WPF == 2) {
                r0 += red[1];
                r1 += red[3];
                r2 += red[5];
            }
            const long long o = (long long)clip * a.n_frames + t;
            if (a.out_rms) a.out_rms[o] = sqrtf(r0 / (float)N);
            if (a.out_flatness) a.out_flatness[o] = expf(r1 / (float)NBINS) / (r2 / (float)NBINS);
        }

        // ---- epilogue 2: banded real filterbanks on the shared power tile
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const ssar_filterbank& fb = a.fb[q];
            if (fb.n_bands <= 0) continue;
            float bmax = -INFINITY;
            for (int b = lt; b < fb.n_bands; b += TPF) {
                const int st = __ldg(fb.start + b), cn = __ldg(fb.count + b);
                const float* __restrict__ w = fb.weights + __ldg(fb.offset + b);
                float acc = 0.f;
                if (fb.on_power) {
                    for (int k = 0; k < cn; ++k) acc = fmaf(__ldg(w + k), pw[st + k], acc);
                } else {
                    for (int k = 0; k < cn; ++k) acc = fmaf(__ldg(w + k), sqrtf(pw[st + k]), acc);
                }
                if (active) fb.out[((long long)clip * a.n_frames + t) * fb.ld + b] = acc;
                bmax = fmaxf(bmax, acc);
            }
            if (fb.max_key) {
                bmax = warp_max(bmax);
                if (active && (lt & 31) == 0) atomic_max_key(fb.max_key + clip, bmax);
            }
        }

        // ---- epilogue 3: banded complex projection of the complex spectrum (CQT basis)
        if (a.proj.n_rows > 0) {
            const ssar_projection& pj = a.proj;
            const int slot = (pj.clip_slot && active) ? __ldg(pj.clip_slot + clip) : 0;
            const int rb = slot * pj.n_rows;
            for (int r = lt; r < pj.n_rows; r += TPF) {
                const int st = __ldg(pj.start + rb + r), cn = __ldg(pj.count + rb + r);
                const float2* __restrict__ w = reinterpret_cast<const float2*>(pj.weights) + __ldg(pj.offset + rb + r);
                float2 acc = make_float2(0.f, 0.f);
                for (int k = 0; k < cn; ++k) {
                    const float2 wk = __ldg(w + k);
                    const float2 X = s[PADI(st + k)];
                    acc.x = fmaf(wk.x, X.x, acc.x);
                    acc.x = fmaf(-wk.y, X.y, acc.x);
                    acc.y = fmaf(wk.x, X.y, acc.y);
                    acc.y = fmaf(wk.y, X.x, acc.y);
                }
                const float sc = pj.row_scale ? pj.scale * __ldg(pj.row_scale + slot * pj.row_scale_ld + r) : pj.scale;
                if (active)
                    reinterpret_cast<float2*>(pj.out)[((long long)clip * a.n_frames + t) * pj.ld + r] =
                        make_float2(acc.x * sc, acc.y * sc);
            }
        }
        __syncthreads();  // the next iteration overwrites s / pw / red
    }
}

template <int M>
int launch_frames(const ssar_frames_args& a, cudaStream_t st) {
    using L = FramesSmem<M>;
    const long long total = (long long)a.n_clips * a.n_frames;
    if (total == 0) return SSAR_OK;
    cudaError_t e = cudaFuncSetAttribute(frames_fft_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::bytes);
    if (e != cudaSuccess) {
        ssar_set_error("frames_fft: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    const long long groups = (total + L::FPB - 1) / L::FPB;
    // persistent grid: a multiple of the SM count, 4 CTAs per SM resident when shared memory allows
    long long grid = (long long)SSAR_NUM_SMS * 4;
    if (groups < grid) grid = groups;
    frames_fft_kernel<M><<<(unsigned)grid, 256, L::bytes, st>>>(a, total);
    return ssar_launch_status("frames_fft");
}

__global__ void fill_u32_kernel(uint32_t* dst, long long n, uint32_t v) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dst[i] = v;
}

}  // namespace

extern "C" int ssar_frames_fft(const ssar_frames_args* args, void* stream) {
    SSAR_REQUIRE(args != nullptr, "frames_fft: null args");
    const ssar_frames_args& a = *args;
    SSAR_REQUIRE(a.x && a.twiddle, "frames_fft: x and twiddle are required");
    SSAR_REQUIRE(a.n_clips >= 0 && a.n_frames >= 0 && a.hop > 0, "frames_fft: bad sizes");
    SSAR_REQUIRE(a.n_samples > a.n_fft / 2, "frames_fft: clip shorter than n_fft/2+1 samples cannot be reflect-padded");
    SSAR_REQUIRE(a.spec_power == 1 || a.spec_power == 2 || !a.out_spec, "frames_fft: spec_power must be 1 or 2");
    SSAR_REQUIRE(a.n_frames == 0 || (long long)(a.n_frames - 1) * a.hop <= a.n_samples,
                 "frames_fft: n_frames=%d exceeds 1 + n_samples/hop", a.n_frames);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (a.n_fft == 2048) return launch_frames<1024>(a, st);
    if (a.n_fft == 1024) return launch_frames<512>(a, st);
    ssar_set_error("frames_fft: n_fft=%d unsupported by the fast path (1024 or 2048); use ssar_frames_dft", a.n_fft);
    return SSAR_ERR_UNSUPPORTED;
}

extern "C" int ssar_fill_u32(uint32_t* dst, long long n, uint32_t value, void* stream) {
    if (n <= 0) return SSAR_OK;
    SSAR_REQUIRE(dst != nullptr, "fill_u32: null dst");
    int grid = ssar_div_up(n, 256);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    fill_u32_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(dst, n, value);
    return ssar_launch_status("fill_u32");
}
