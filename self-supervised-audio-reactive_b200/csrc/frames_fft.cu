// Framed, windowed real FFT with fused epilogues (power/magnitude spectrum, complex spectrum, banded real
// filterbanks such as mel, banded complex projections such as the CQT basis, RMS, spectral flatness).
//
// v1: ONE WARP PER FRAME (two frames per warp for n_fft 1024).  Every lane holds 32 complex points in registers and the
// 1024-point (512-point) complex transform of the packed real frame is two register-resident radix-32 (radix-32 + 2 x
// radix-16) passes around one warp-private shared-memory transpose (fft_warp.cuh) — the frame loop has no
// __syncthreads, only __syncwarp.  One launch covers every clip of a batch; warps walk the (clip, frame) list
// persistently with neighbouring warps on neighbouring frames, so the 2x-4x overlapped waveform reads are served by
// L1/L2 and HBM sees each sample once.  The spectrum of a frame lives only in shared memory: |X|^2 -> filterbanks happen
// before anything is written, so the [frames, 1025] spectrogram never round-trips HBM unless it is a requested output.
#include <stdlib.h>

#include "fft_warp.cuh"

namespace {

constexpr int LANE_PAD = 64;  // zeroed floats behind the M + 1 bins of a power tile (lane-interleaved banks read into them)

template <int M, int WPB>
size_t frames_smem_bytes(bool window, bool separate_power) {
    using L = WarpFft<M>;
    size_t b = (size_t)32 * L::N1 * sizeof(float2);                  // twiddle table [k2][n1]
    if (window) b += (size_t)2 * M * sizeof(float);
    b += (size_t)WPB * L::FPW * L::ZF * sizeof(float2);
    if (separate_power) b += (size_t)WPB * L::FPW * (M + 4 + LANE_PAD) * sizeof(float);
    return b;
}

// WPB warps per CTA, CPS CTAs per SM: (8, 2) = 16 resident warps at <= 128 registers; (24, 1) = 24 resident warps at
// <= 85 registers (one CTA owns the SM's shared memory: 24 frame buffers of 8.4 KB + the tables)
template <int M, int WPB, int CPS>
__global__ void __launch_bounds__(32 * WPB, CPS) frames_fft_kernel(const ssar_frames_args a, const long long total_frames,
                                                                   const int separate_power) {
    using L = WarpFft<M>;
    constexpr int N1 = L::N1, FPW = L::FPW, TLD = L::TLD, ZF = L::ZF, N = 2 * M, NBINS = M + 1, NJ = L::NJ;
    extern __shared__ float4 smem_raw[];
    float2* twT = reinterpret_cast<float2*>(smem_raw);                 // [32][N1]: W_M^{n1 k2}
    float* wins = reinterpret_cast<float*>(twT + 32 * N1);             // [2M] or absent
    float2* zall = reinterpret_cast<float2*>(wins + (a.window ? N : 0));
    float* pall = reinterpret_cast<float*>(zall + WPB * FPW * ZF);  // only with separate_power

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h = lane / N1, n1 = lane % N1;  // frame slot inside the warp, point index inside the frame
    const int aux_frames = a.aux_frames > 0 ? a.aux_frames : a.n_frames;
    const int sub = a.sub_every > 1 ? a.sub_every : 1;  // complex / rms rows: every sub-th frame
    const int c_rows = (a.n_frames + sub - 1) / sub, r_rows = aux_frames / sub;
    const float2* __restrict__ tw = reinterpret_cast<const float2*>(a.twiddle);  // exp(-2 pi i k / 2M), k in [0, 2M)

    warp_fft_fill_twiddles<M>(twT, tw);
    if (a.window)
        for (int i = threadIdx.x; i < N; i += blockDim.x) wins[i] = __ldg(a.window + i);
    __syncthreads();

    float2* zw = zall + (size_t)warp * FPW * ZF;  // this warp's frame buffers
    const bool need_pw = a.out_spec || a.out_flatness || a.fb[0].n_bands > 0 || a.fb[1].n_bands > 0 || a.pip.count;

    for (long long fbase = ((long long)blockIdx.x * WPB + warp) * FPW; fbase < total_frames;
         fbase += (long long)gridDim.x * WPB * FPW) {
        // ------------------------------------------------------------------ load + window (lane: n = n1 + N1 r)
        const long long f = fbase + h;
        const bool in_grid = f < total_frames;
        const int clip = in_grid ? (int)(f / a.n_frames) : 0;
        const int t = in_grid ? (int)(f - (long long)clip * a.n_frames) : 0;
        // ragged batches: frames past the clip's own count transform zeros (and write zero rows below)
        const bool active = in_grid && (!a.clip_frames || t < __ldg(a.clip_frames + clip));
        const long long n_samples = a.clip_samples ? __ldg(a.clip_samples + clip) : a.n_samples;
        float2 v[32];
        float ss = 0.f;
        if (active) {
            const float* __restrict__ xc = a.x + (long long)clip * a.clip_stride;
            long long start = (long long)t * a.hop - M;
            bool interior = start >= 0 && start + N <= n_samples;
            if (a.clip_total > 0) {  // time segment of a longer clip: frame position and reflection in clip coordinates
                const long long gstart = (long long)(a.frame_offset + t) * a.hop - M;
                interior = gstart >= 0 && gstart + N <= a.clip_total;
                start = gstart - a.sample_offset;  // interior frames: plain segment-relative reads (the halo covers them)
            }
            const bool vec = interior && ((((long long)clip * a.clip_stride + start) & 1LL) == 0) &&
                             ((reinterpret_cast<uintptr_t>(a.x) & 7) == 0);
            if (vec) {
                const float2* __restrict__ xp = reinterpret_cast<const float2*>(xc + start);
#pragma unroll
                for (int r = 0; r < 32; ++r) v[r] = __ldg(xp + n1 + N1 * r);
            } else if (interior) {
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    const int n = n1 + N1 * r;
                    v[r] = make_float2(__ldg(xc + start + 2 * n), __ldg(xc + start + 2 * n + 1));
                }
            } else {
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    const int n = n1 + N1 * r;
                    if (a.clip_total > 0) {
                        const long long g = start + a.sample_offset + 2 * n;
                        v[r] = make_float2(__ldg(xc + (ssar_reflect(g, a.clip_total) - a.sample_offset)),
                                           __ldg(xc + (ssar_reflect(g + 1, a.clip_total) - a.sample_offset)));
                    } else {
                        v[r] = make_float2(__ldg(xc + ssar_reflect(start + 2 * n, n_samples)),
                                           __ldg(xc + ssar_reflect(start + 2 * n + 1, n_samples)));
                    }
                }
            }
            if (a.out_rms) {
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    ss = fmaf(v[r].x, v[r].x, ss);
                    ss = fmaf(v[r].y, v[r].y, ss);
                }
            }
            if (a.window) {
                const float2* wp = reinterpret_cast<const float2*>(wins);
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    v[r] = __fmul2_rn(v[r], wp[n1 + N1 * r]);  // packed: same bits as two FMULs
                }
            }
        } else {
#pragma unroll
            for (int r = 0; r < 32; ++r) v[r] = make_float2(0.f, 0.f);
        }

        // ------------------------------------------------------------------ the transform (fft_warp.cuh)
        warp_fft_forward<M>(v, zw, twT, lane);

        // rms: sum of squares over the N1 lanes of each frame
        if (a.out_rms) {
#pragma unroll
            for (int o = N1 / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            if (in_grid && n1 == 0 && t < aux_frames && t % sub == 0)
                a.out_rms[(long long)clip * r_rows + t / sub] = sqrtf(ss / (float)N);
        }

        // ------------------------------------------------------------------ per frame: untangle + epilogues, 32 lanes wide
#pragma unroll 1
        for (int fs = 0; fs < FPW; ++fs) {
            const long long ff = fbase + fs;
            if (ff >= total_frames) break;  // warp-uniform
            const int fclip = (int)(ff / a.n_frames);
            const int ft = (int)(ff - (long long)fclip * a.n_frames);
            const long long crow = (long long)fclip * c_rows + ft / sub;  // row of the complex output (if ft % sub == 0)
            const long long orow = (long long)fclip * aux_frames + ft;  // row of every other output
            // ragged: this clip's own frame count; its aux outputs stop (n_frames - aux_frames) frames earlier.  A frame past
            // the count carries an all-zero spectrum: it is written (zero rows) but feeds no maximum and no peak list
            const int cf = a.clip_frames ? __ldg(a.clip_frames + fclip) : a.n_frames;
            const bool valid = ft < cf - (a.n_frames - aux_frames);
            const bool aux = ft < aux_frames;                            // warp-uniform
            float2* z = zw + fs * ZF;

            // untangle the packed real transform in place: X[k] = E + w^k O, X[M-k] = conj(E - w^k O).  A lane's bins (k and
            // M - k, k = lane + 32 j) are touched by no other lane and no other j, so the loads of a batch of UB iterations
            // are issued together (the compiler cannot prove that and would serialise load -> store -> load)
            constexpr int UJ = M / 64;  // full iterations: k = lane + 32 j <= M / 2 - 1 for every lane; then k = M / 2 (lane 0)
            constexpr int UB = 4;
#pragma unroll
            for (int j0 = 0; j0 < UJ; j0 += UB) {
                float2 zk[UB], zm[UB], w[UB];
#pragma unroll
                for (int u = 0; u < UB; ++u) {
                    const int k = lane + 32 * (j0 + u);
                    zk[u] = z[k];
                    zm[u] = z[k == 0 ? 0 : M - k];
                    w[u] = __ldg(&tw[k]);
                }
#pragma unroll
                for (int u = 0; u < UB; ++u) {
                    const int k = lane + 32 * (j0 + u);
                    // packed (re, im) arithmetic where both components take the same operation; same bits as the scalar form
                    const float2 zc = make_float2(zm[u].x, -zm[u].y);  // conj(Z[M - k])
                    const float2 e = __fmul2_rn(make_float2(0.5f, 0.5f), __fadd2_rn(zk[u], zc));
                    const float2 d = __fadd2_rn(zk[u], make_float2(-zc.x, -zc.y));
                    const float2 o = make_float2(0.5f * d.y, -0.5f * d.x);
                    const float2 wo = cmul(w[u], o);
                    z[k] = __fadd2_rn(e, wo);
                    z[M - k] = cconj(__fadd2_rn(e, make_float2(-wo.x, -wo.y)));
                }
            }
            static_assert(UJ % UB == 0, "untangle batches");
            if (lane == 0) {  // k = M / 2: its own mirror
                const float2 zk = z[M / 2], zm = zk;
                const float2 w = __ldg(&tw[M / 2]);
                const float2 e = make_float2(0.5f * (zk.x + zm.x), 0.5f * (zk.y - zm.y));
                const float2 d = make_float2(zk.x - zm.x, zk.y + zm.y);
                const float2 o = make_float2(0.5f * d.y, -0.5f * d.x);
                const float2 wo = cmul(w, o);
                z[M / 2] = cconj(csub(e, wo));
            }
            __syncwarp();

            // epilogue 1: complex / power / magnitude outputs, flatness sums; the power tile goes to shared memory
            float* pw = separate_power ? pall + (size_t)(warp * FPW + fs) * (M + 4 + LANE_PAD) : reinterpret_cast<float*>(z);
            const bool cplx = a.out_complex && ft % sub == 0;
            if (cplx || (need_pw && aux)) {
                float2* oc = cplx ? reinterpret_cast<float2*>(a.out_complex) + crow * a.ldc : nullptr;
                float* os = (a.out_spec && aux) ? a.out_spec + orow * a.lds : nullptr;
                float pj[NJ];
                float slog = 0.f, slin = 0.f;
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    const int k = lane + 32 * j;
                    pj[j] = 0.f;
                    if (k < NBINS) {
                        const float2 X = z[k];
                        if (oc) oc[k] = X;
                        const float p = fmaf(X.x, X.x, X.y * X.y);
                        pj[j] = p;
                        if (os) os[k] = (a.spec_power == 2) ? p : sqrtf(p);
                        if (a.out_flatness) {
                            const float pt = fmaxf(1e-10f, p);
                            slog += logf(pt);
                            slin += pt;
                        }
                    }
                }
                if (a.out_flatness && aux) {
                    slog = warp_sum(slog);
                    slin = warp_sum(slin);
                    if (lane == 0) a.out_flatness[orow] = expf(slog / (float)NBINS) / (slin / (float)NBINS);
                }
                if (aux && (a.fb[0].n_bands > 0 || a.fb[1].n_bands > 0 || a.pip.count)) {
                    if (!separate_power) __syncwarp();  // every lane has read its bins: the tile may be overwritten
#pragma unroll
                    for (int j = 0; j < NJ; ++j) {
                        const int k = lane + 32 * j;
                        if (k < NBINS) pw[k] = pj[j];
                    }
                    static_assert(2 * ZF >= NBINS + 3 + LANE_PAD, "the zeroed tail must fit the aliased frame buffer");
                    for (int k = NBINS + lane; k < NBINS + 3 + LANE_PAD; k += 32) pw[k] = 0.f;
                    __syncwarp();
                }
            }

            // epilogue 2: banded real filterbanks on the shared power tile
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const ssar_filterbank& fb = a.fb[q];
                if (fb.n_bands <= 0 || !aux) continue;
                float bmax = -INFINITY;
                if (fb.lane_weights && fb.on_power) {
                    // lane-interleaved weights: one coalesced 512-byte row of weights and one 128-bit tile read per
                    // 4 bins, warp-uniform trip count per pass of 32 bands
                    const int n_pass = (fb.n_bands + 31) >> 5;
                    for (int p = 0; p < n_pass; ++p) {
                        const int b = 32 * p + lane;
                        const int n4 = __ldg(fb.lane_count4 + p);
                        const float4* __restrict__ w =
                            reinterpret_cast<const float4*>(fb.lane_weights) + (size_t)__ldg(fb.lane_offset4 + p) * 32 + lane;
                        const float4* pp = reinterpret_cast<const float4*>(pw + __ldg(fb.lane_start4 + b));
                        float acc = 0.f;
                        // weights two groups of 4 rows ahead of the sums (with ~200 KB of shared memory per CTA the L1 is
                        // small and a weight row is an L2 hit more often than not); same order of additions
                        constexpr int WG = 4;
                        float4 wv[WG];
#pragma unroll
                        for (int u = 0; u < WG; ++u)
                            wv[u] = u < n4 ? __ldg(w + (size_t)u * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
                        for (int k4 = 0; k4 < n4; k4 += WG) {
                            float4 nx[WG];
#pragma unroll
                            for (int u = 0; u < WG; ++u)
                                nx[u] = k4 + WG + u < n4 ? __ldg(w + (size_t)(k4 + WG + u) * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                            for (int u = 0; u < WG; ++u) {
                                if (k4 + u < n4) {  // warp-uniform
                                    const float4 pv = pp[k4 + u];
                                    acc = fmaf(wv[u].x, pv.x, acc);
                                    acc = fmaf(wv[u].y, pv.y, acc);
                                    acc = fmaf(wv[u].z, pv.z, acc);
                                    acc = fmaf(wv[u].w, pv.w, acc);
                                }
                            }
#pragma unroll
                            for (int u = 0; u < WG; ++u) wv[u] = nx[u];
                        }
                        if (b < fb.n_bands) {
                            fb.out[orow * fb.ld + b] = acc;
                            bmax = fmaxf(bmax, acc);
                        }
                    }
                } else {  // packed weights: one lane per band, a trip per bin
                    for (int b = lane; b < fb.n_bands; b += 32) {
                        const int st = __ldg(fb.start + b), cn = __ldg(fb.count + b);
                        const float* __restrict__ w = fb.weights + __ldg(fb.offset + b);
                        float acc = 0.f;
                        if (fb.on_power) {
                            for (int k = 0; k < cn; ++k) acc = fmaf(__ldg(w + k), pw[st + k], acc);
                        } else {
                            for (int k = 0; k < cn; ++k) acc = fmaf(__ldg(w + k), sqrtf(pw[st + k]), acc);
                        }
                        fb.out[orow * fb.ld + b] = acc;
                        bmax = fmaxf(bmax, acc);
                    }
                }
                if (fb.max_key && valid) {
                    bmax = warp_max(bmax);
                    if (lane == 0) atomic_max_key(fb.max_key + fclip, bmax);
                }
            }

            // epilogue 2b: piptrack on the magnitudes sqrt(power tile) (rosa/pitch.py:27-83; tuning.cu's piptrack_kernel
            // reads the same values back from a [frames, 1025] spectrogram in HBM)
            if (a.pip.count && aux && valid) {
                const ssar_piptrack_epilogue& pp = a.pip;
                float mxp = 0.f;
                for (int k = lane; k < NBINS; k += 32) mxp = fmaxf(mxp, pw[k]);
                const float ref = pp.threshold * sqrtf(warp_max(mxp));  // sqrt is monotone: max of the magnitudes
                for (int kb = pp.k_lo; kb < pp.k_hi; kb += 32) {
                    const int k = kb + lane;
                    float p = 0.f, m = 0.f;
                    bool is_peak = false;
                    if (k < pp.k_hi) {
                        const float c = sqrtf(pw[k]);
                        const float prev = k > 0 ? sqrtf(pw[k - 1]) : 0.f;
                        const float next = k + 1 < NBINS ? sqrtf(pw[k + 1]) : 0.f;
                        const float cm = c > ref ? c : 0.f;
                        const float pm = (k > 0 && prev > ref) ? prev : 0.f;
                        const float nm = (k + 1 < NBINS && next > ref) ? next : 0.f;
                        if (cm > pm && cm >= nm) {
                            float avg = 0.f, shift = 0.f;
                            if (k > 0 && k + 1 < NBINS) {
                                avg = 0.5f * (next - prev);
                                const float den = 2.f * c - next - prev;
                                shift = avg / (den + (fabsf(den) < 1.17549435e-38f ? 1.f : 0.f));
                            }
                            p = ((float)k + shift) * pp.sr / (float)N;
                            m = c + 0.5f * avg * shift;
                            is_peak = p > 0.f;
                        }
                    }
                    const unsigned ballot = __ballot_sync(0xffffffffu, is_peak);
                    if (ballot) {
                        int base = 0;
                        if (lane == 0) base = atomicAdd(pp.count + fclip, __popc(ballot));
                        base = __shfl_sync(0xffffffffu, base, 0);
                        if (is_peak) {
                            const long long slot = base + __popc(ballot & ((1u << lane) - 1u));
                            if (slot < pp.capacity) {
                                pp.pitch[fclip * pp.capacity + slot] = p;
                                pp.mag[fclip * pp.capacity + slot] = m;
                            }
                        }
                    }
                }
            }

            // epilogue 3: banded complex projection of the complex spectrum (CQT basis); needs z intact
            if (a.proj.n_rows > 0 && aux) {
                const ssar_projection& pjn = a.proj;
                const int slot = pjn.clip_slot ? __ldg(pjn.clip_slot + fclip) : 0;
                const int rb = slot * pjn.n_rows;
                for (int r = lane; r < pjn.n_rows; r += 32) {
                    const int st = __ldg(pjn.start + rb + r), cn = __ldg(pjn.count + rb + r);
                    const float2* __restrict__ w = reinterpret_cast<const float2*>(pjn.weights) + __ldg(pjn.offset + rb + r);
                    float2 acc = make_float2(0.f, 0.f);
                    for (int k = 0; k < cn; ++k) {
                        const float2 wk = __ldg(w + k);
                        const float2 X = z[st + k];
                        acc.x = fmaf(wk.x, X.x, acc.x);
                        acc.x = fmaf(-wk.y, X.y, acc.x);
                        acc.y = fmaf(wk.x, X.y, acc.y);
                        acc.y = fmaf(wk.y, X.x, acc.y);
                    }
                    const float sc = pjn.row_scale ? pjn.scale * __ldg(pjn.row_scale + slot * pjn.row_scale_ld + r) : pjn.scale;
                    reinterpret_cast<float2*>(pjn.out)[orow * pjn.ld + r] = make_float2(acc.x * sc, acc.y * sc);
                }
            }
        }
        __syncwarp();  // the next iteration overwrites the frame buffers
    }
}

template <int M, int WPB, int CPS>
int launch_frames_cfg(const ssar_frames_args& a, cudaStream_t st, long long total, int separate_power) {
    using L = WarpFft<M>;
    const size_t smem = frames_smem_bytes<M, WPB>(a.window != nullptr, separate_power != 0);
    cudaError_t e = cudaFuncSetAttribute(frames_fft_kernel<M, WPB, CPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("frames_fft: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    const long long groups = (total + WPB * L::FPW - 1) / (WPB * L::FPW);
    long long grid = (long long)SSAR_NUM_SMS * CPS;  // persistent grid: a multiple of the SM count
    if (groups < grid) grid = groups;
    frames_fft_kernel<M, WPB, CPS><<<(unsigned)grid, 32 * WPB, smem, st>>>(a, total, separate_power);
    return ssar_launch_status("frames_fft");
}

// 24 warps per SM at 80 registers (236 B of spills per thread) beat 16 warps at 128: cfg2 10.34 -> 8.99 ms per step, full
// workload 12.0 -> 11.0 ms of frames_fft per step.  SSAR_FRAMES_WIDE=0 selects the 2 x 8-warp configuration.
int g_frames_wide = -1;

template <int M>
int launch_frames(const ssar_frames_args& a, cudaStream_t st) {
    const long long total = (long long)a.n_clips * a.n_frames;
    if (total == 0) return SSAR_OK;
    const bool banks = a.fb[0].n_bands > 0 || a.fb[1].n_bands > 0 || a.pip.count;
    const int separate_power = (banks && a.proj.n_rows > 0) ? 1 : 0;  // the projection needs the complex tile intact
    if (a.pip.count) {
        cudaError_t em = cudaMemsetAsync(a.pip.count, 0, sizeof(int) * a.n_clips, st);
        if (em != cudaSuccess) {
            ssar_set_error("frames_fft: memset: %s", cudaGetErrorString(em));
            return SSAR_ERR_CUDA;
        }
    }
    if (g_frames_wide < 0) {
        const char* env = getenv("SSAR_FRAMES_WIDE");
        g_frames_wide = (env && env[0] == '0') ? 0 : 1;
    }
    if (g_frames_wide && !separate_power) return launch_frames_cfg<M, 24, 1>(a, st, total, separate_power);
    return launch_frames_cfg<M, 8, 2>(a, st, total, separate_power);
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
    SSAR_REQUIRE(a.n_samples > a.n_fft / 2 || a.clip_total > 0, "frames_fft: clip shorter than n_fft/2+1 samples cannot be reflect-padded");
    SSAR_REQUIRE(a.spec_power == 1 || a.spec_power == 2 || !a.out_spec, "frames_fft: spec_power must be 1 or 2");
    SSAR_REQUIRE(a.aux_frames >= 0 && a.aux_frames <= a.n_frames, "frames_fft: aux_frames=%d outside [0, n_frames]", a.aux_frames);
    SSAR_REQUIRE(a.sub_every >= 0 && (a.sub_every <= 1 || (!a.clip_frames && a.clip_total == 0 &&
                                                           (a.aux_frames > 0 ? a.aux_frames : a.n_frames) % a.sub_every == 0)),
                 "frames_fft: sub_every needs dense clips and (aux_)frames divisible by it");
    SSAR_REQUIRE(!a.pip.count || (a.pip.pitch && a.pip.mag && 0 <= a.pip.k_lo && a.pip.k_lo < a.pip.k_hi &&
                                  a.pip.k_hi <= a.n_fft / 2 + 1 &&
                                  a.pip.capacity >= ssar_piptrack_capacity(a.aux_frames > 0 ? a.aux_frames : a.n_frames, a.pip.k_lo, a.pip.k_hi)),
                 "frames_fft: bad piptrack epilogue (null lists, bin range or capacity)");
    if (a.clip_total > 0) {
        SSAR_REQUIRE(!a.clip_samples && a.sample_offset >= 0 && a.sample_offset + a.n_samples <= a.clip_total && a.frame_offset >= 0,
                     "frames_fft: bad time segment (offset / length outside the clip, or combined with ragged lengths)");
        SSAR_REQUIRE(a.clip_total > a.n_fft / 2 && a.n_frames >= 0 &&
                         (long long)(a.frame_offset + a.n_frames - 1) * a.hop <= a.clip_total,
                     "frames_fft: the segment's frames exceed 1 + clip_total / hop");
        // every sample a frame of the segment touches (after reflection at the clip's ends) must be resident
        if (a.n_frames > 0) {
            const long long first = (long long)a.frame_offset * a.hop - a.n_fft / 2;
            const long long last = (long long)(a.frame_offset + a.n_frames - 1) * a.hop + a.n_fft / 2 - 1;
            long long lo = first < 0 ? 0 : first, hi = last >= a.clip_total ? a.clip_total - 1 : last;
            if (first < 0 && -first > hi) hi = -first;                                       // mirror of the leading pad
            if (last >= a.clip_total && 2 * (a.clip_total - 1) - last < lo) lo = 2 * (a.clip_total - 1) - last;  // trailing
            SSAR_REQUIRE(a.sample_offset <= lo && hi < a.sample_offset + a.n_samples,
                         "frames_fft: the segment does not hold the n_fft / 2 halo its frames need");
        }
    } else
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
