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
