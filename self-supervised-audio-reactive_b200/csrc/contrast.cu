// Spectral contrast (ssar/features/audio.py:72-123): per frame and per octave band, the mean of the q smallest
// (valley) and q largest (peak) magnitude bins, then power_to_db(peak) - power_to_db(valley) where each power_to_db
// floors at its own clip-global maximum - 80 dB (rosa/convert.py:7-12).
//
// Kernel 1: one warp per (frame, band).  The band's bins are staged in a per-warp shared-memory slab; the q extremes
//   are extracted by q rounds of warp arg-min / arg-max (q <= 32; 1..10 at the reference's settings), which is the
//   partial selection the reference's full torch.sort is used for.  Peak / valley maxima per clip are folded into
//   order-preserving keys with one atomic per warp.
// Kernel 2: elementwise dB difference with the two clip-global floors.
#include "common.cuh"

namespace {

constexpr int WARPS = 8;

__device__ __forceinline__ void warp_arg_reduce(float& v, int& i, bool want_min) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, i, o);
        const bool better = want_min ? (ov < v) : (ov > v);
        if (better || (ov == v && oi < i)) {
            v = ov;
            i = oi;
        }
    }
}

// ---- bands of at most 512 bins (every band at the reference's settings): SORTED LANE COLUMNS.
// Lane l holds bins l, l + 32, ... (PL of them, padded with +inf) as order-preserving keys, sorts them in registers
// (bitonic, compile-time indices) and parks the column in shared memory.  The q smallest of the band are then q rounds
// of {warp-wide min of the column heads (one redux), the first lane holding it advances its head}; the q largest walk
// the same columns from the back.  Only sums are wanted, so ties need no index.  Same values in the same order as the
// arg-min rounds of the general kernel below -> bit-identical sums, at ~1/4 of the instructions.
template <int PL>
__device__ __forceinline__ void sort_keys(uint32_t (&k)[PL]) {
#pragma unroll
    for (int len = 2; len <= PL; len <<= 1) {
#pragma unroll
        for (int j = len >> 1; j > 0; j >>= 1) {
#pragma unroll
            for (int i = 0; i < PL; ++i) {
                const int o = i ^ j;
                if (o > i) {
                    const uint32_t lo = min(k[i], k[o]), hi = max(k[i], k[o]);
                    const bool up = (i & len) == 0;
                    k[i] = up ? lo : hi;
                    k[o] = up ? hi : lo;
                }
            }
        }
    }
}

template <int PL>
__device__ __forceinline__ void band_extremes(const float* __restrict__ src, int n, int q, int l, uint32_t* col,
                                              float& valley_mean, float& peak_mean) {
    constexpr uint32_t PAD = 0xff800000u;  // f2key(+inf)
    uint32_t k[PL];
#pragma unroll
    for (int j = 0; j < PL; ++j) {
        const int i = l + 32 * j;
        k[j] = i < n ? f2key(__ldg(src + i)) : PAD;
    }
    sort_keys<PL>(k);
    if (PL > 1) {
#pragma unroll
        for (int j = 0; j < PL; ++j) col[j * 32] = k[j];  // private column: no warp sync needed
    }
    const int cnt = l < n ? (n - l + 31) / 32 : 0;
    float acc = 0.f;
    {
        uint32_t head = k[0];
        int pf = 0;
        for (int r = 0; r < q; ++r) {
            const uint32_t m = __reduce_min_sync(0xffffffffu, head);
            acc += key2f(m);
            const unsigned who = __ballot_sync(0xffffffffu, head == m);
            // branch-free advance: with the update inside `if (winner)` nvcc 12.9 feeds the redux a stale head
            const bool win = l == __ffs(who) - 1;
            pf += win ? 1 : 0;
            const uint32_t nxt = PL > 1 ? col[min(pf, PL - 1) * 32] : PAD;
            head = win ? (pf < PL ? nxt : PAD) : head;
        }
    }
    valley_mean = acc / (float)q;
    acc = 0.f;
    {
        int pb = cnt - 1;
        uint32_t head = 0u;  // below every real key
        if (PL > 1) {
            if (pb >= 0) head = col[pb * 32];
        } else if (pb >= 0) {
            head = k[0];
        }
        for (int r = 0; r < q; ++r) {
            const uint32_t m = __reduce_max_sync(0xffffffffu, head);
            acc += key2f(m);
            const unsigned who = __ballot_sync(0xffffffffu, head == m);
            const bool win = l == __ffs(who) - 1;
            pb -= win ? 1 : 0;
            const uint32_t nxt = PL > 1 ? col[max(pb, 0) * 32] : 0u;
            head = win ? (pb >= 0 ? nxt : 0u) : head;
        }
    }
    peak_mean = acc / (float)q;
}

__global__ void __launch_bounds__(WARPS * 32) contrast_bands_sorted_kernel(
    const float* __restrict__ mag, long long ld, long long n_rows, int n_frames, int nb, const int* __restrict__ start,
    const int* __restrict__ count, const int* __restrict__ qn, float* __restrict__ peak, float* __restrict__ valley,
    uint32_t* __restrict__ peak_key, uint32_t* __restrict__ valley_key) {
    __shared__ uint32_t cols[WARPS][16 * 32];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    uint32_t* col = &cols[w][l];
    const long long tasks = n_rows * nb;
    for (long long task = (long long)blockIdx.x * WARPS + w; task < tasks; task += (long long)gridDim.x * WARPS) {
        const long long row = task / nb;
        const int band = (int)(task - row * nb);
        const int st = __ldg(start + band), n = __ldg(count + band), q = __ldg(qn + band);
        const float* src = mag + row * ld + st;
        float vm, pm;
        if (n <= 32) band_extremes<1>(src, n, q, l, col, vm, pm);
        else if (n <= 64) band_extremes<2>(src, n, q, l, col, vm, pm);
        else if (n <= 128) band_extremes<4>(src, n, q, l, col, vm, pm);
        else if (n <= 256) band_extremes<8>(src, n, q, l, col, vm, pm);
        else band_extremes<16>(src, n, q, l, col, vm, pm);
        if (l == 0) {
            valley[row * nb + band] = vm;
            peak[row * nb + band] = pm;
            const int clip = (int)(row / n_frames);
            atomic_max_key(valley_key + clip, vm);
            atomic_max_key(peak_key + clip, pm);
        }
    }
}

__global__ void __launch_bounds__(WARPS * 32) contrast_bands_kernel(
    const float* __restrict__ mag, long long ld, long long n_rows, int n_frames, int nb, const int* __restrict__ start,
    const int* __restrict__ count, const int* __restrict__ qn, int slab, float* __restrict__ peak,
    float* __restrict__ valley, uint32_t* __restrict__ peak_key, uint32_t* __restrict__ valley_key) {
    extern __shared__ float sm[];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    float* s = sm + (size_t)w * slab;
    const long long tasks = n_rows * nb;
    for (long long task = (long long)blockIdx.x * WARPS + w; task < tasks; task += (long long)gridDim.x * WARPS) {
        const long long row = task / nb;
        const int band = (int)(task - row * nb);
        const int st = __ldg(start + band), n = __ldg(count + band), q = __ldg(qn + band);
        const float* src = mag + row * ld + st;
        for (int i = l; i < n; i += 32) s[i] = __ldg(src + i);
        __syncwarp();
        float sums[2];
#pragma unroll
        for (int side = 0; side < 2; ++side) {  // 0: q smallest (valley), 1: q largest (peak)
            const bool want_min = side == 0;
            const float sentinel = want_min ? INFINITY : -INFINITY;
            float acc = 0.f, keep_v = 0.f;
            int keep_i = -1;
            for (int r = 0; r < q; ++r) {
                float best = sentinel;
                int bi = 0x7fffffff;
                for (int i = l; i < n; i += 32) {
                    const float v = s[i];
                    if (want_min ? (v < best) : (v > best)) {
                        best = v;
                        bi = i;
                    }
                }
                warp_arg_reduce(best, bi, want_min);
                acc += best;
                if (l == r) {  // lane r remembers what round r removed so the slab can be restored
                    keep_v = best;
                    keep_i = bi;
                }
                if (l == 0 && bi != 0x7fffffff) s[bi] = sentinel;
                __syncwarp();
            }
            if (keep_i >= 0 && keep_i != 0x7fffffff) s[keep_i] = keep_v;
            __syncwarp();
            sums[side] = acc / (float)q;
        }
        if (l == 0) {
            valley[row * nb + band] = sums[0];
            peak[row * nb + band] = sums[1];
            const int clip = (int)(row / n_frames);
            atomic_max_key(valley_key + clip, sums[0]);
            atomic_max_key(peak_key + clip, sums[1]);
        }
        __syncwarp();
    }
}

__device__ __forceinline__ float db10(float x) { return 10.0f * log10f(fmaxf(1e-10f, x)); }

__global__ void contrast_finish_kernel(const float* __restrict__ peak, const float* __restrict__ valley,
                                       const uint32_t* __restrict__ peak_key, const uint32_t* __restrict__ valley_key,
                                       long long n_per_clip, int linear, float top_db, float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float pf = db10(key2f(peak_key[clip])) - top_db, vf = db10(key2f(valley_key[clip])) - top_db;
    const long long base = (long long)clip * n_per_clip;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_per_clip; i += (long long)gridDim.x * blockDim.x) {
        const float p = peak[base + i], v = valley[base + i];
        out[base + i] = linear ? (p - v) : (fmaxf(db10(p), pf) - fmaxf(db10(v), vf));
    }
}

}  // namespace

extern "C" int ssar_contrast_bands(const float* mag, long long ld, int n_clips, int n_frames, int n_bands,
                                   const int* band_start, const int* band_count, const int* band_q, int max_count,
                                   float* peak, float* valley, uint32_t* peak_key, uint32_t* valley_key, void* stream) {
    if (n_clips == 0 || n_frames == 0 || n_bands == 0) return SSAR_OK;
    SSAR_REQUIRE(mag && band_start && band_count && band_q && peak && valley && peak_key && valley_key,
                 "contrast_bands: null pointer");
    SSAR_REQUIRE(max_count > 0 && max_count <= 4096, "contrast_bands: band wider than 4096 bins");
    const long long rows_all = (long long)n_clips * n_frames;
    if (max_count <= 512) {
        long long g = (rows_all * n_bands + WARPS - 1) / WARPS;
        if (g > (long long)SSAR_NUM_SMS * 8) g = (long long)SSAR_NUM_SMS * 8;
        contrast_bands_sorted_kernel<<<(unsigned)g, WARPS * 32, 0, (cudaStream_t)stream>>>(
            mag, ld, rows_all, n_frames, n_bands, band_start, band_count, band_q, peak, valley, peak_key, valley_key);
        return ssar_launch_status("contrast_bands");
    }
    const size_t smem = (size_t)WARPS * max_count * sizeof(float);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(contrast_bands_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) {
            ssar_set_error("contrast_bands: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return SSAR_ERR_CUDA;
        }
    }
    const long long rows = (long long)n_clips * n_frames;
    long long grid = (rows * n_bands + WARPS - 1) / WARPS;
    if (grid > (long long)SSAR_NUM_SMS * 8) grid = (long long)SSAR_NUM_SMS * 8;
    contrast_bands_kernel<<<(unsigned)grid, WARPS * 32, smem, (cudaStream_t)stream>>>(
        mag, ld, rows, n_frames, n_bands, band_start, band_count, band_q, max_count, peak, valley, peak_key, valley_key);
    return ssar_launch_status("contrast_bands");
}

extern "C" int ssar_contrast_finish(const float* peak, const float* valley, const uint32_t* peak_key,
                                    const uint32_t* valley_key, int n_clips, long long n_per_clip, int linear,
                                    float top_db, float* out, void* stream) {
    if (n_clips == 0 || n_per_clip == 0) return SSAR_OK;
    SSAR_REQUIRE(peak && valley && peak_key && valley_key && out, "contrast_finish: null pointer");
    int gx = ssar_div_up(n_per_clip, 256);
    if (gx > SSAR_NUM_SMS * 4) gx = SSAR_NUM_SMS * 4;
    contrast_finish_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(peak, valley, peak_key, valley_key,
                                                                              n_per_clip, linear, top_db, out);
    return ssar_launch_status("contrast_finish");
}
