// Inverse STFT = per-frame inverse real FFT * window  ->  overlap-add / window-envelope normalisation.
// Replaces torch.istft behind rosa/spectral.py:24-32 (used by features/audio.py:13-24 harmonic/percussive with
// hop = n_fft/2 and by rosa/beat.py:69 plp with hop = 1).
//
// irfft_frames: warp-per-frame transform (fft_warp.cuh).  The one-sided spectrum is re-packed into an M = n_fft/2
//   point complex transform (Z = E + iO), run through the FORWARD passes on conj(Z) (inverse = conj(FFT(conj(.)))/M),
//   windowed and stored as [clip, frame, n_fft].
// overlap_add: gather formulation: sum the <= ceil(n_fft/hop) frames that cover an output sample, divide by the summed
//   squared window (torch.istft's envelope), center-trim by n_fft/2.
// istft_half_overlap: both fused for hop = n_fft/2 (the HPSS case): a warp walks a run of consecutive frames of one
//   clip, keeps the previous frame's second half in shared memory and writes finished output blocks directly, so the
//   [frames, n_fft] intermediate (2x the signal, written and read) never exists.
#include "fft_warp.cuh"

namespace {

// inverse untangle of one frame into the warp's natural-order buffer: z[k] = conj(Z[k]) with Z = E + iO,
// E = (X[k] + conj(X[M-k]))/2, O = conj(w^k) (X[k] - conj(X[M-k]))/2; imaginary parts of DC / Nyquist ignored (C2R)
template <int M>
__device__ __forceinline__ void inverse_untangle_one(float2 xk, float2 xm, float2 w, int k, float2* z) {
    if (k == 0) {
        xk.y = 0.f;
        xm.y = 0.f;
    }
    const float2 e = make_float2(0.5f * (xk.x + xm.x), 0.5f * (xk.y - xm.y));
    const float2 d = make_float2(0.5f * (xk.x - xm.x), 0.5f * (xk.y + xm.y));
    const float2 o = cmul(cconj(w), d);
    z[k] = make_float2(e.x - o.y, -(e.y + o.x));            // conj(Z[k])
    if (k > 0) z[M - k] = make_float2(e.x + o.y, e.y - o.x);  // conj(Z[M-k])
}

template <int M>
__device__ __forceinline__ void inverse_untangle(const float2* __restrict__ Xf, bool active, const float2* __restrict__ tw,
                                                 float2* z, int lane) {
    // the spectrum rows come from HBM: the loads of a batch of UB iterations are issued together (one exposed memory
    // latency per batch instead of one per iteration)
    constexpr int UJ = M / 64, UB = 8;  // k = lane + 32 j <= M / 2 - 1 for j < UJ and every lane; then k = M / 2 (lane 0)
    static_assert(UJ % UB == 0, "untangle batches");
    const float2 zero = make_float2(0.f, 0.f);
#pragma unroll
    for (int j0 = 0; j0 < UJ; j0 += UB) {
        float2 xk[UB], xm[UB], w[UB];
#pragma unroll
        for (int u = 0; u < UB; ++u) {
            const int k = lane + 32 * (j0 + u);
            xk[u] = active ? __ldg(&Xf[k]) : zero;
            xm[u] = active ? __ldg(&Xf[M - k]) : zero;
            w[u] = __ldg(&tw[k]);
        }
#pragma unroll
        for (int u = 0; u < UB; ++u) inverse_untangle_one<M>(xk[u], xm[u], w[u], lane + 32 * (j0 + u), z);
    }
    if (lane == 0) {
        const float2 x = active ? __ldg(&Xf[M / 2]) : zero;
        inverse_untangle_one<M>(x, x, __ldg(&tw[M / 2]), M / 2, z);
    }
}

template <int M>
__global__ void __launch_bounds__(256, 2) irfft_frames_kernel(const float2* __restrict__ X, long long ldx,
                                                              const float* __restrict__ win,
                                                              const float2* __restrict__ tw, long long total_frames,
                                                              float* __restrict__ out) {
    using L = WarpFft<M>;
    constexpr int N1 = L::N1, FPW = L::FPW, ZF = L::ZF, N = 2 * M;
    extern __shared__ float4 smem_raw[];
    float2* twT = reinterpret_cast<float2*>(smem_raw);
    float2* zall = twT + 32 * N1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h = lane / N1, n1 = lane % N1;
    warp_fft_fill_twiddles<M>(twT, tw);
    __syncthreads();
    float2* zw = zall + (size_t)warp * FPW * ZF;
    const float inv_m = 1.0f / (float)M;

    for (long long fbase = ((long long)blockIdx.x * L::WPB + warp) * FPW; fbase < total_frames;
         fbase += (long long)gridDim.x * L::WPB * FPW) {
#pragma unroll 1
        for (int fs = 0; fs < FPW; ++fs) {
            const long long f = fbase + fs;
            inverse_untangle<M>(X + (f < total_frames ? f : 0) * ldx, f < total_frames, tw, zw + fs * ZF, lane);
        }
        __syncwarp();
        float2 v[32];
        {
            const float2* z = zw + h * ZF;
#pragma unroll
            for (int r = 0; r < 32; ++r) v[r] = z[n1 + N1 * r];
        }
        __syncwarp();
        warp_fft_forward<M>(v, zw, twT, lane);
        // x[2n] = Re(conj(Y[n]))/M, x[2n+1] = Im(conj(Y[n]))/M, times the synthesis window
#pragma unroll 1
        for (int fs = 0; fs < FPW; ++fs) {
            const long long f = fbase + fs;
            if (f >= total_frames) break;
            const float2* z = zw + fs * ZF;
            float2* of = reinterpret_cast<float2*>(out + f * N);
#pragma unroll 4
            for (int n = lane; n < M; n += 32) {
                const float2 y = z[n];
                float a = y.x * inv_m, b = -y.y * inv_m;
                if (win) {
                    const float2 w = __ldg(reinterpret_cast<const float2*>(win) + n);
                    a *= w.x;
                    b *= w.y;
                }
                of[n] = make_float2(a, b);
            }
        }
        __syncwarp();
    }
}

// Ragged batches (all three overlap-add kernels and the fused inverse): clip c has clip_frames[c] <= n_frames frames and
// clip_len[c] <= length output samples; frames / samples beyond them do not exist (no contribution to the sums or the
// window envelope) and the tail of the output row is written as zeros.  Both NULL = uniform batch.
__global__ void overlap_add_kernel(const float* __restrict__ frames, const float* __restrict__ win, int n_fft, int hop,
                                   int cap_frames, long long length, float* __restrict__ out,
                                   const int* __restrict__ clip_frames, const long long* __restrict__ clip_len) {
    const int clip = blockIdx.y;
    const float* fc = frames + (long long)clip * cap_frames * n_fft;
    float* oc = out + (long long)clip * length;
    const int n_frames = clip_frames ? __ldg(clip_frames + clip) : cap_frames;
    const long long len_c = clip_len ? __ldg(clip_len + clip) : length;
    const long long total = (long long)n_fft + (long long)hop * (n_frames - 1);
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < length; n += (long long)gridDim.x * blockDim.x) {
        const long long p = n + n_fft / 2;
        float acc = 0.f, env = 0.f;
        if (p < total && n < len_c) {
            long long t_hi = p / hop;
            if (t_hi > n_frames - 1) t_hi = n_frames - 1;
            long long t_lo = (p - n_fft + hop) / hop;  // ceil((p - n_fft + 1)/hop) for p - n_fft + 1 > 0
            if (p - n_fft + 1 <= 0) t_lo = 0;
#pragma unroll 8  // hop 1 (PLP) sums n_fft frames per sample: keep 8 independent loads in flight, same summation order
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

// Dense overlap (n_fft / hop >= 64: PLP's hop-1 inverse, rosa/beat.py:72, sums 1024 frames per sample): the frame range
// of a sample is cut into OD_SEG consecutive segments summed by OD_SEG threads (8x the loads in flight of the serial
// gather, which is latency-bound at 46 k threads), and the partial sums are added in ascending frame order.
constexpr int OD_SEG = 8;
__global__ void __launch_bounds__(32 * OD_SEG) overlap_add_dense_kernel(const float* __restrict__ frames,
                                                                        const float* __restrict__ win, int n_fft, int hop,
                                                                        int cap_frames, long long length,
                                                                        float* __restrict__ out,
                                                                        const int* __restrict__ clip_frames,
                                                                        const long long* __restrict__ clip_len) {
    __shared__ float sacc[OD_SEG][32], senv[OD_SEG][32];
    const int clip = blockIdx.y, seg = threadIdx.y;
    const float* fc = frames + (long long)clip * cap_frames * n_fft;
    float* oc = out + (long long)clip * length;
    const int n_frames = clip_frames ? __ldg(clip_frames + clip) : cap_frames;
    const long long len_c = clip_len ? __ldg(clip_len + clip) : length;
    const long long total = (long long)n_fft + (long long)hop * (n_frames - 1);
    // Persistent over 32-sample blocks (one full wave of CTAs instead of 1.1).  A sample covered by all n_fft / hop
    // frames sees the window in the same order whatever its position, so this warp's share of the envelope is computed
    // once (same fmaf order as the general loop below) and reused for every such block.
    const int full = n_fft / hop;  // frames over an interior sample (hop divides n_fft in this path)
    float env_full = 0.f;
    bool have_env_full = false;
    for (long long blk = blockIdx.x; blk * 32 < length; blk += gridDim.x) {
        const long long n = blk * 32 + threadIdx.x;
        const long long p = n + n_fft / 2;
        float acc = 0.f, env = 0.f;
        if (n < len_c && p < total) {
            long long t_hi = p / hop;
            if (t_hi > n_frames - 1) t_hi = n_frames - 1;
            long long t_lo = (p - n_fft + hop) / hop;
            if (p - n_fft + 1 <= 0) t_lo = 0;
            const long long cnt = t_hi - t_lo + 1, per = (cnt + OD_SEG - 1) / OD_SEG;
            const long long a = t_lo + seg * per;
            long long b = a + per - 1;
            if (b > t_hi) b = t_hi;
            // hop == 1: i = p - t runs over the same values for every interior sample
            const bool interior = hop == 1 && cnt == full;
            const bool warp_interior = __all_sync(__activemask(), interior) && __activemask() == 0xffffffffu;
            if (warp_interior && have_env_full) {
#pragma unroll 8
                for (long long t = a; t <= b; ++t) acc += __ldg(fc + t * n_fft + (int)(p - t * hop));
                env = env_full;
            } else {
#pragma unroll 8
                for (long long t = a; t <= b; ++t) {
                    const int i = (int)(p - t * hop);
                    acc += __ldg(fc + t * n_fft + i);
                    const float w = win ? __ldg(win + i) : 1.f;
                    env = fmaf(w, w, env);
                }
                if (warp_interior) {
                    env_full = env;
                    have_env_full = true;
                }
            }
        }
        sacc[seg][threadIdx.x] = acc;
        senv[seg][threadIdx.x] = env;
        __syncthreads();
        if (seg == 0 && n < length) {
#pragma unroll
            for (int s2 = 1; s2 < OD_SEG; ++s2) {
                acc += sacc[s2][threadIdx.x];
                env += senv[s2][threadIdx.x];
            }
            oc[n] = env > 1e-11f ? acc / env : 0.f;
        }
        __syncthreads();
    }
}

// Same gather, four consecutive output samples per thread (128-bit loads / stores, 32-bit index arithmetic): used when
// hop, n_fft / 2 and the clip length are multiples of 4 and every index fits 31 bits — the HPSS case (hop = n_fft / 2).
// The four samples share their covering frames; per sample the frames are summed in the same order as above.
__global__ void __launch_bounds__(256) overlap_add_vec4_kernel(const float* __restrict__ frames, const float* __restrict__ win,
                                                               int n_fft, int hop, int n_frames, int length,
                                                               float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float* fc = frames + (long long)clip * n_frames * n_fft;
    float* oc = out + (long long)clip * length;
    const int total = n_fft + hop * (n_frames - 1);
    const int n4 = length >> 2;
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < n4; q += gridDim.x * blockDim.x) {
        const int p = 4 * q + n_fft / 2;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f), env = acc;
        if (p < total) {  // total is a multiple of 4 here, so the four samples are inside or outside together
            int t_hi = p / hop;
            if (t_hi > n_frames - 1) t_hi = n_frames - 1;
            const int t_lo = (p - n_fft + 1 <= 0) ? 0 : (p - n_fft + hop) / hop;
            for (int t = t_lo; t <= t_hi; ++t) {
                const int i = p - t * hop;  // multiple of 4, 0 <= i <= n_fft - 4
                const float4 v = __ldg(reinterpret_cast<const float4*>(fc + (long long)t * n_fft + i));
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                float4 w = make_float4(1.f, 1.f, 1.f, 1.f);
                if (win) w = __ldg(reinterpret_cast<const float4*>(win + i));
                env.x = fmaf(w.x, w.x, env.x); env.y = fmaf(w.y, w.y, env.y);
                env.z = fmaf(w.z, w.z, env.z); env.w = fmaf(w.w, w.w, env.w);
            }
        }
        float4 r;
        r.x = env.x > 1e-11f ? acc.x / env.x : 0.f;
        r.y = env.y > 1e-11f ? acc.y / env.y : 0.f;
        r.z = env.z > 1e-11f ? acc.z / env.z : 0.f;
        r.w = env.w > 1e-11f ? acc.w / env.w : 0.f;
        reinterpret_cast<float4*>(oc)[q] = r;
    }
}

template <int M>
int launch_irfft(const float* spec, long long ld, const float* win, const float* tw, long long total, float* frames,
                 cudaStream_t st) {
    using L = WarpFft<M>;
    const size_t smem = ((size_t)32 * L::N1 + (size_t)L::WPB * L::FPW * L::ZF) * sizeof(float2);
    cudaError_t e = cudaFuncSetAttribute(irfft_frames_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("irfft_frames: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    long long grid = (total + L::WPB * L::FPW - 1) / (L::WPB * L::FPW);
    if (grid > (long long)SSAR_NUM_SMS * 2) grid = (long long)SSAR_NUM_SMS * 2;
    irfft_frames_kernel<M><<<(unsigned)grid, 256, smem, st>>>(reinterpret_cast<const float2*>(spec), ld, win,
                                                             reinterpret_cast<const float2*>(tw), total, frames);
    return ssar_launch_status("irfft_frames");
}

// ---------------------------------------------------------------------------------------------------------------
// Fused inverse STFT for hop = n_fft / 2 = M (n_fft = 2048).  Output block b (samples n in [(b-1) M, b M), i.e. padded
// positions p = n + M in [b M, (b+1) M)) = (second half of frame b-1 + first half of frame b) / (w[i+M]^2 + w[i]^2),
// summed in ascending frame order like overlap_add_kernel.  A warp owns a run of RUN consecutive blocks of one clip: it
// transforms frames b0-1 .. b0+RUN-1 one after the other, parks each frame's second half in shared memory and emits a
// block per frame (one extra transform per run: 1/RUN overhead).
// ---------------------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256, 2) istft_half_overlap_kernel(const float2* __restrict__ X, long long ldx,
                                                                    const float* __restrict__ win,
                                                                    const float2* __restrict__ tw, int n_clips, int cap_frames,
                                                                    long long length, float* __restrict__ out, int OLA_RUN,
                                                                    const int* __restrict__ clip_frames,
                                                                    const long long* __restrict__ clip_len) {
    constexpr int M = 1024;
    using L = WarpFft<M>;
    constexpr int ZF = L::ZF;
    extern __shared__ float4 smem_raw[];
    float2* twT = reinterpret_cast<float2*>(smem_raw);          // [32][32]
    float2* zall = twT + 32 * 32;                                // per warp: frame buffer ZF + parked half M/2
    const float2* __restrict__ wins = reinterpret_cast<const float2*>(win);  // (w[2n], w[2n+1]); L1-resident
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    warp_fft_fill_twiddles<M>(twT, tw);
    __syncthreads();
    float2* zw = zall + (size_t)warp * (ZF + M / 2);
    float2* parked = zw + ZF;
    const float inv_m = 1.0f / (float)M;

    const long long n_blocks = (length + M - 1) / M;                    // output blocks per clip: b = 1 .. n_blocks
    const long long runs_per_clip = (n_blocks + OLA_RUN - 1) / OLA_RUN;
    const long long total_runs = runs_per_clip * n_clips;
    for (long long run = (long long)blockIdx.x * L::WPB + warp; run < total_runs; run += (long long)gridDim.x * L::WPB) {
        const int clip = (int)(run / runs_per_clip);
        const long long b0 = 1 + (run - (long long)clip * runs_per_clip) * OLA_RUN;
        long long b1 = b0 + OLA_RUN;  // exclusive
        if (b1 > n_blocks + 1) b1 = n_blocks + 1;
        const float2* Xc = X + (long long)clip * cap_frames * ldx;
        float* oc = out + (long long)clip * length;
        const int n_frames = clip_frames ? __ldg(clip_frames + clip) : cap_frames;
        const long long len_c = clip_len ? __ldg(clip_len + clip) : length;
        for (long long t = b0 - 1; t < b1; ++t) {
            const bool have = t < n_frames;          // frame t exists (warp-uniform)
            const bool have_prev = t - 1 < n_frames;  // frame t - 1 exists
            inverse_untangle<M>(Xc + (have ? t : 0) * ldx, have, tw, zw, lane);
            __syncwarp();
            float2 v[32];
#pragma unroll
            for (int r = 0; r < 32; ++r) v[r] = zw[lane + 32 * r];
            __syncwarp();
            warp_fft_forward<M>(v, zw, twT, lane);
            // windowed frame: x_t[2n] = Re(conj(Y[n]))/M * w[2n], x_t[2n+1] = Im(conj(Y[n]))/M * w[2n+1]
            if (t >= b0) {  // emit block t: first half of this frame + parked second half of frame t - 1
                const long long n0 = (t - 1) * M;
#pragma unroll 4
                for (int j = 0; j < 16; ++j) {
                    const int n = lane + 32 * j;  // packed index: samples i = 2n, 2n+1 < M
                    const float2 y = zw[n];
                    const float2 one = make_float2(1.f, 1.f);
                    const float2 wc = win ? __ldg(wins + n) : one, wp = win ? __ldg(wins + n + M / 2) : one;
                    const float2 cur = make_float2(y.x * inv_m * wc.x, -y.y * inv_m * wc.y);
                    const float2 prev = parked[n];
                    float2 acc = make_float2(prev.x, prev.y), env = make_float2(0.f, 0.f);
                    if (have_prev) env = make_float2(fmaf(wp.x, wp.x, 0.f), fmaf(wp.y, wp.y, 0.f));
                    if (have) {
                        acc.x += cur.x;
                        acc.y += cur.y;
                        env.x = fmaf(wc.x, wc.x, env.x);
                        env.y = fmaf(wc.y, wc.y, env.y);
                    }
                    const long long g = n0 + 2 * n;
                    const float r0 = (env.x > 1e-11f && g < len_c) ? acc.x / env.x : 0.f;
                    const float r1 = (env.y > 1e-11f && g + 1 < len_c) ? acc.y / env.y : 0.f;
                    if (g + 1 < length) {
                        *reinterpret_cast<float2*>(oc + g) = make_float2(r0, r1);
                    } else if (g < length) {
                        oc[g] = r0;
                    }
                }
            }
            __syncwarp();
            // park this frame's second half for the next block
#pragma unroll 4
            for (int j = 0; j < 16; ++j) {
                const int n = lane + 32 * j + M / 2;
                const float2 y = zw[n];
                const float2 w = win ? __ldg(wins + n) : make_float2(1.f, 1.f);
                parked[n - M / 2] = have ? make_float2(y.x * inv_m * w.x, -y.y * inv_m * w.y) : make_float2(0.f, 0.f);
            }
            __syncwarp();
        }
    }
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

extern "C" int ssar_istft_half_overlap(const float* spec, long long ld, int n_clips, int n_frames, int n_fft,
                                       const float* window, const float* twiddle, long long length, float* out,
                                       const int* clip_frames, const long long* clip_len, void* stream) {
    if (n_clips == 0 || length == 0) return SSAR_OK;
    SSAR_REQUIRE(spec && twiddle && out, "istft_half_overlap: null pointer");
    SSAR_REQUIRE(n_fft == 2048, "istft_half_overlap: n_fft must be 2048 (hop = 1024); use irfft_frames + overlap_add");
    SSAR_REQUIRE(n_frames > 0 && (length % 2 == 0 || n_clips == 1) && ((uintptr_t)out % 8 == 0) &&
                     (!window || (uintptr_t)window % 8 == 0),
                 "istft_half_overlap: needs an even length (or a single clip) and 8-byte aligned buffers");
    using L = WarpFft<1024>;
    const size_t smem = ((size_t)32 * 32 + (size_t)L::WPB * (L::ZF + 512)) * sizeof(float2);
    cudaError_t e = cudaFuncSetAttribute(istft_half_overlap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("istft_half_overlap: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    const long long n_blocks = (length + 1023) / 1024;
    // run length: a run of R blocks costs R + 1 transforms and the launch takes ceil(runs / resident warps) rounds, so
    // pick the R in [8, 32] with the cheapest rounds * (R + 1) (8 clips x 5760 blocks: R = 20 fills 288 of the 296 CTA
    // slots in one round; the fixed R = 24 of v1 left 19 % of them empty)
    const long long slots = (long long)SSAR_NUM_SMS * 2 * L::WPB;
    int OLA_RUN = 24;
    long long best = -1;
    for (int R = 8; R <= 32; ++R) {
        const long long r_ = ((n_blocks + R - 1) / R) * n_clips;
        const long long cost = ((r_ + slots - 1) / slots) * (R + 1);
        if (best < 0 || cost <= best) {
            best = cost;
            OLA_RUN = R;
        }
    }
    const long long runs = ((n_blocks + OLA_RUN - 1) / OLA_RUN) * n_clips;
    long long grid = (runs + L::WPB - 1) / L::WPB;
    if (grid > (long long)SSAR_NUM_SMS * 2) grid = (long long)SSAR_NUM_SMS * 2;
    istft_half_overlap_kernel<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(
        reinterpret_cast<const float2*>(spec), ld, window, reinterpret_cast<const float2*>(twiddle), n_clips, n_frames, length,
        out, OLA_RUN, clip_frames, clip_len);
    return ssar_launch_status("istft_half_overlap");
}

extern "C" int ssar_overlap_add(const float* frames, const float* window, int n_clips, int n_frames, int n_fft, int hop,
                                long long length, float* out, const int* clip_frames, const long long* clip_len,
                                void* stream) {
    if (n_clips == 0 || length == 0) return SSAR_OK;
    SSAR_REQUIRE(frames && out, "overlap_add: null pointer");
    SSAR_REQUIRE(hop > 0 && n_fft > 0 && n_frames > 0, "overlap_add: bad sizes");
    const long long span = (long long)n_fft + (long long)hop * (n_frames - 1);
    const bool vec4 = !clip_frames && !clip_len && hop % 4 == 0 && n_fft % 8 == 0 && length % 4 == 0 && span < (1LL << 30) &&
                      length < (1LL << 30) &&
                      ((uintptr_t)frames % 16 == 0) && ((uintptr_t)out % 16 == 0) && (!window || (uintptr_t)window % 16 == 0);
    if (vec4) {
        int g4 = ssar_div_up(length / 4, 256);
        if (g4 > SSAR_NUM_SMS * 16) g4 = SSAR_NUM_SMS * 16;
        overlap_add_vec4_kernel<<<dim3(g4, n_clips), 256, 0, (cudaStream_t)stream>>>(frames, window, n_fft, hop, n_frames,
                                                                                 (int)length, out);
        return ssar_launch_status("overlap_add");
    }
    if (n_fft / hop >= 64 && (length + 31) / 32 <= 0x7fffffffLL) {
        long long gd = (length + 31) / 32;
        const long long capd = ((long long)SSAR_NUM_SMS * 8 + n_clips - 1) / n_clips;  // 8 CTAs of 256 threads per SM
        if (gd > capd) gd = capd;
        overlap_add_dense_kernel<<<dim3((unsigned)gd, n_clips), dim3(32, OD_SEG), 0, (cudaStream_t)stream>>>(
            frames, window, n_fft, hop, n_frames, length, out, clip_frames, clip_len);
        return ssar_launch_status("overlap_add");
    }
    int gx = ssar_div_up(length, 256);
    if (gx > SSAR_NUM_SMS * 16) gx = SSAR_NUM_SMS * 16;
    overlap_add_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(frames, window, n_fft, hop, n_frames, length, out,
                                                                            clip_frames, clip_len);
    return ssar_launch_status("overlap_add");
}
