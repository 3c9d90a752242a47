// Per-frame features computed from (mel / magnitude) spectra: power_to_db with the clip-global floor, onset
// strength (spectral flux), MFCC (dB + DCT-II), min/max reductions.
#include "block_utils.cuh"

namespace {

__device__ __forceinline__ float to_db(float x, float floor_db) {
    // rosa/convert.py:7-12 with ref_value = 1: 10*log10(max(amin, x)) - 10*log10(max(amin, 1)) floored at max-80
    return fmaxf(10.0f * log10f(fmaxf(1e-10f, x)), floor_db);
}
__device__ __forceinline__ float db_floor_from_key(uint32_t key, float top_db) {
    return 10.0f * log10f(fmaxf(1e-10f, key2f(key))) - top_db;
}

// one CTA per (clip, frame); aggregate 0 = mean, 1 = lower median (torch.median)
__global__ void onset_flux_kernel(const float* __restrict__ mel, const uint32_t* __restrict__ max_key, int T, int nb,
                                  int pad_width, int aggregate, float top_db, float* __restrict__ out) {
    extern __shared__ float sm[];
    float* scratch = sm;       // 32
    float* vals = sm + 32;     // next_pow2(nb) for the median path
    const int clip = blockIdx.y, t = blockIdx.x;
    const int j = t - pad_width;  // flux index: S[:, j+1] - S[:, j]
    float result = 0.f;
    if (j >= 0 && j + 1 < T) {  // uniform per block
        const float fl = db_floor_from_key(max_key[clip], top_db);
        const float* a = mel + ((long long)clip * T + j) * nb;
        const float* b = a + nb;
        if (aggregate == 0) {
            float acc = 0.f;
            for (int m = threadIdx.x; m < nb; m += blockDim.x) acc += fmaxf(0.f, to_db(b[m], fl) - to_db(a[m], fl));
            result = block_sum(acc, scratch) / (float)nb;
        } else {
            const int np = next_pow2(nb);
            for (int m = threadIdx.x; m < np; m += blockDim.x)
                vals[m] = m < nb ? fmaxf(0.f, to_db(b[m], fl) - to_db(a[m], fl)) : INFINITY;
            block_bitonic_sort(vals, np);
            result = vals[(nb - 1) / 2];
        }
    }
    if (threadIdx.x == 0) out[(long long)clip * T + t] = result;
}

// one warp per frame: dB then DCT-II rows; dct is [n_mfcc, nb] row-major
__global__ void mfcc_kernel(const float* __restrict__ mel, const uint32_t* __restrict__ max_key,
                            const float* __restrict__ dct, int T, int nb, int n_mfcc, float top_db,
                            float* __restrict__ out, int n_clips) {
    extern __shared__ float sm[];
    const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, l = threadIdx.x & 31;
    float* row = sm + w * nb;
    const long long total = (long long)n_clips * T;
    for (long long f = (long long)blockIdx.x * warps + w; f < total; f += (long long)gridDim.x * warps) {
        const int clip = (int)(f / T);
        const float fl = db_floor_from_key(max_key[clip], top_db);
        for (int m = l; m < nb; m += 32) row[m] = to_db(mel[f * nb + m], fl);
        __syncwarp();
        for (int k = l; k < n_mfcc; k += 32) {
            const float* d = dct + (long long)k * nb;
            float acc = 0.f;
            for (int m = 0; m < nb; ++m) acc = fmaf(__ldg(d + m), row[m], acc);
            out[f * n_mfcc + k] = acc;
        }
        __syncwarp();
    }
}

// v1 of the mean-aggregated onset envelope for 128 bands: one warp walks OF_RUN consecutive frames, each lane keeps the
// dB values of its 4 bands of the previous frame in registers, so every mel row is read (128-bit loads) and converted
// to dB once; the band mean is a shuffle reduction.  (v0 spends one 128-thread CTA and two block barriers per frame.)
constexpr int OF_RUN_MEAN = 16, OF_RUN_MEDIAN = 4;  // the median's 31 serial select rounds per frame want more, shorter runs
template <bool MEDIAN>
__global__ void __launch_bounds__(256) onset_flux_mean128_kernel(const float* __restrict__ mel,
                                                                 const uint32_t* __restrict__ max_key, int T, int pad_width,
                                                                 float top_db, float* __restrict__ out, int runs_per_clip) {
    const int lane = threadIdx.x & 31;
    const long long wid = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int clip = blockIdx.y;
    if (wid >= runs_per_clip) return;
    constexpr int OF_RUN = MEDIAN ? OF_RUN_MEDIAN : OF_RUN_MEAN;
    const int t0 = (int)wid * OF_RUN;
    const float fl = db_floor_from_key(max_key[clip], top_db);
    const float4* __restrict__ rows = reinterpret_cast<const float4*>(mel + (long long)clip * T * 128) + lane;
    float* __restrict__ oc = out + (long long)clip * T;
    float4 prev = make_float4(0.f, 0.f, 0.f, 0.f);
    int jprev = -2;  // flux index whose dB row `prev` holds
    // the run's OF_RUN + 1 mel rows are loaded up front (one exposed memory latency per run instead of one per frame)
    float4 raw[OF_RUN + 1];
#pragma unroll
    for (int u = 0; u <= OF_RUN; ++u) {
        const int jr = t0 - pad_width + u;
        raw[u] = (jr >= 0 && jr < T) ? __ldg(rows + (long long)jr * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < OF_RUN; ++u) {
        const int t = t0 + u;
        if (t >= T) break;
        const int j = t - pad_width;
        float result = 0.f;
        if (j >= 0 && j + 1 < T) {
            if (jprev != j) {
                const float4 a = raw[u];
                prev = make_float4(to_db(a.x, fl), to_db(a.y, fl), to_db(a.z, fl), to_db(a.w, fl));
            }
            const float4 b = raw[u + 1];
            const float4 cur = make_float4(to_db(b.x, fl), to_db(b.y, fl), to_db(b.z, fl), to_db(b.w, fl));
            const float d[4] = {fmaxf(0.f, cur.x - prev.x), fmaxf(0.f, cur.y - prev.y), fmaxf(0.f, cur.z - prev.z),
                                fmaxf(0.f, cur.w - prev.w)};
            if (!MEDIAN) {
                float acc = d[0];
                acc += d[1];
                acc += d[2];
                acc += d[3];
                result = warp_sum(acc) / 128.f;
            } else {
                // lower median (torch.median) = the value of rank 63 of the 128 non-negative flux values: bitwise radix
                // select on the float bit patterns (monotone for values >= 0) — the largest v with #(values < v) <= 63
                const unsigned u[4] = {__float_as_uint(d[0]) & 0x7fffffffu, __float_as_uint(d[1]) & 0x7fffffffu,
                                       __float_as_uint(d[2]) & 0x7fffffffu, __float_as_uint(d[3]) & 0x7fffffffu};
                unsigned prefix = 0;
#pragma unroll 1
                for (int bit = 30; bit >= 0; --bit) {
                    const unsigned cand = prefix | (1u << bit);
                    const unsigned c = (u[0] < cand) + (u[1] < cand) + (u[2] < cand) + (u[3] < cand);
                    if (__reduce_add_sync(0xffffffffu, c) <= 63u) prefix = cand;
                }
                result = __uint_as_float(prefix);
            }
            prev = cur;
            jprev = j + 1;
        }
        if (lane == 0) oc[t] = result;
    }
}

// v1 of the MFCC tail: one warp per 4 consecutive frames; the DCT matrix sits transposed in shared memory (lanes =
// coefficients: conflict-free), the 4 dB rows are read as 128-bit broadcasts.  Same ascending-band fmaf chain per
// coefficient as v0 (bit-identical), ~190 instead of ~2 700 shared/L1 wavefronts per frame.
constexpr int MF_FR = 4;
__global__ void __launch_bounds__(256) mfcc4_kernel(const float* __restrict__ mel, const uint32_t* __restrict__ max_key,
                                                    const float* __restrict__ dct, int T, int nb, int n_mfcc, float top_db,
                                                    float* __restrict__ out, int n_clips) {
    extern __shared__ float4 sm4[];
    float* dctT = reinterpret_cast<float*>(sm4);                      // [nb][n_mfcc]
    const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, l = threadIdx.x & 31;
    float* rows = dctT + ((nb * n_mfcc + 3) & ~3) + (size_t)w * MF_FR * nb;  // [MF_FR][nb], 16-byte aligned
    for (int i = threadIdx.x; i < nb * n_mfcc; i += blockDim.x) {
        const int k = i / nb, m = i - k * nb;
        dctT[m * n_mfcc + k] = __ldg(dct + i);
    }
    __syncthreads();
    const long long total = (long long)n_clips * T;
    const long long groups = (total + MF_FR - 1) / MF_FR;
    for (long long g = (long long)blockIdx.x * warps + w; g < groups; g += (long long)gridDim.x * warps) {
#pragma unroll
        for (int fr = 0; fr < MF_FR; ++fr) {
            const long long f = g * MF_FR + fr;
            if (f < total) {
                const float fl = db_floor_from_key(max_key[(int)(f / T)], top_db);
                for (int m = l; m < nb; m += 32) rows[fr * nb + m] = to_db(__ldg(mel + f * nb + m), fl);
            } else {
                for (int m = l; m < nb; m += 32) rows[fr * nb + m] = 0.f;
            }
        }
        __syncwarp();
        for (int k = l; k < n_mfcc; k += 32) {
            float acc[MF_FR];
#pragma unroll
            for (int fr = 0; fr < MF_FR; ++fr) acc[fr] = 0.f;
            for (int m = 0; m < nb; m += 4) {
                const float d0 = dctT[m * n_mfcc + k], d1 = dctT[(m + 1) * n_mfcc + k];
                const float d2 = dctT[(m + 2) * n_mfcc + k], d3 = dctT[(m + 3) * n_mfcc + k];
#pragma unroll
                for (int fr = 0; fr < MF_FR; ++fr) {
                    const float4 r = *reinterpret_cast<const float4*>(rows + fr * nb + m);
                    acc[fr] = fmaf(d0, r.x, acc[fr]);
                    acc[fr] = fmaf(d1, r.y, acc[fr]);
                    acc[fr] = fmaf(d2, r.z, acc[fr]);
                    acc[fr] = fmaf(d3, r.w, acc[fr]);
                }
            }
#pragma unroll
            for (int fr = 0; fr < MF_FR; ++fr) {
                const long long f = g * MF_FR + fr;
                if (f < total) out[f * n_mfcc + k] = acc[fr];
            }
        }
        __syncwarp();
    }
}

__global__ void power_to_db_kernel(const float* __restrict__ x, long long n_per_clip, const uint32_t* __restrict__ max_key,
                                   float top_db, int has_top, float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float fl = has_top ? db_floor_from_key(max_key[clip], top_db) : -INFINITY;
    const float* xc = x + (long long)clip * n_per_clip;
    float* oc = out + (long long)clip * n_per_clip;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x)
        oc[i] = to_db(xc[i], fl);
}

__global__ void minmax_kernel(const float* __restrict__ x, long long n_per_clip, uint32_t* min_key, uint32_t* max_key) {
    __shared__ float scratch[32];
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * n_per_clip;
    float mn = INFINITY, mx = -INFINITY;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x) {
        const float v = xc[i];
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = block_min(mn, scratch);
    mx = block_max(mx, scratch);
    if (threadIdx.x == 0) {
        if (min_key) atomic_min_key(min_key + clip, mn);
        if (max_key) atomic_max_key(max_key + clip, mx);
    }
}

}  // namespace

extern "C" int ssar_onset_flux(const float* mel, const uint32_t* max_key, int n_clips, int n_frames, int n_bands,
                               int pad_width, int aggregate, float top_db, float* out, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(mel && max_key && out, "onset_flux: null pointer");
    SSAR_REQUIRE(n_bands > 0 && n_bands <= 4096, "onset_flux: n_bands out of range");
    SSAR_REQUIRE(aggregate == 0 || aggregate == 1, "onset_flux: aggregate must be 0 (mean) or 1 (median)");
    if (n_bands == 128 && (reinterpret_cast<uintptr_t>(mel) & 15) == 0) {
        const int runs = ssar_div_up(n_frames, aggregate == 0 ? OF_RUN_MEAN : OF_RUN_MEDIAN);
        if (aggregate == 0)
            onset_flux_mean128_kernel<false><<<dim3(ssar_div_up(runs, 8), n_clips), 256, 0, (cudaStream_t)stream>>>(
                mel, max_key, n_frames, pad_width, top_db, out, runs);
        else
            onset_flux_mean128_kernel<true><<<dim3(ssar_div_up(runs, 8), n_clips), 256, 0, (cudaStream_t)stream>>>(
                mel, max_key, n_frames, pad_width, top_db, out, runs);
        return ssar_launch_status("onset_flux");
    }
    dim3 grid(n_frames, n_clips);
    const size_t smem = (32 + next_pow2(n_bands)) * sizeof(float);
    onset_flux_kernel<<<grid, 128, smem, (cudaStream_t)stream>>>(mel, max_key, n_frames, n_bands, pad_width, aggregate,
                                                                top_db, out);
    return ssar_launch_status("onset_flux");
}

extern "C" int ssar_mfcc(const float* mel, const uint32_t* max_key, const float* dct, int n_clips, int n_frames,
                         int n_bands, int n_mfcc, float top_db, float* out, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(mel && max_key && dct && out, "mfcc: null pointer");
    SSAR_REQUIRE(n_bands > 0 && n_bands <= 1024, "mfcc: n_bands out of range");
    const int warps = 8;
    const long long total = (long long)n_clips * n_frames;
    if (n_bands % 4 == 0) {
        const size_t smem4 = (((size_t)n_bands * n_mfcc + 3) & ~(size_t)3) * sizeof(float) +
                             (size_t)warps * MF_FR * n_bands * sizeof(float);
        if (smem4 <= 200 * 1024) {
            if (smem4 > 48 * 1024)
                cudaFuncSetAttribute(mfcc4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4);
            int g4 = ssar_div_up(ssar_div_up(total, MF_FR), warps);
            if (g4 > SSAR_NUM_SMS * 4) g4 = SSAR_NUM_SMS * 4;
            mfcc4_kernel<<<g4, warps * 32, smem4, (cudaStream_t)stream>>>(mel, max_key, dct, n_frames, n_bands, n_mfcc,
                                                                        top_db, out, n_clips);
            return ssar_launch_status("mfcc");
        }
    }
    int grid = ssar_div_up(total, warps);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    mfcc_kernel<<<grid, warps * 32, warps * n_bands * sizeof(float), (cudaStream_t)stream>>>(
        mel, max_key, dct, n_frames, n_bands, n_mfcc, top_db, out, n_clips);
    return ssar_launch_status("mfcc");
}

extern "C" int ssar_power_to_db(const float* x, int n_clips, long long n_per_clip, const uint32_t* max_key,
                                float top_db, int has_top, float* out, void* stream) {
    if (n_clips == 0 || n_per_clip == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out && (max_key || !has_top), "power_to_db: null pointer");
    int gx = ssar_div_up(n_per_clip, 256 * 4);
    if (gx > SSAR_NUM_SMS * 8) gx = SSAR_NUM_SMS * 8;
    power_to_db_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(x, n_per_clip, max_key, top_db, has_top, out);
    return ssar_launch_status("power_to_db");
}

extern "C" int ssar_minmax_keys(const float* x, int n_clips, long long n_per_clip, uint32_t* min_key, uint32_t* max_key,
                                void* stream) {
    if (n_clips == 0 || n_per_clip == 0) return SSAR_OK;
    SSAR_REQUIRE(x, "minmax_keys: null pointer");
    int gx = ssar_div_up(n_per_clip, 256 * 8);
    if (gx > SSAR_NUM_SMS * 4) gx = SSAR_NUM_SMS * 4;
    minmax_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(x, n_per_clip, min_key, max_key);
    return ssar_launch_status("minmax_keys");
}
