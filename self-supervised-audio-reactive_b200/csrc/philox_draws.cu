// Generator draws of the trajectory path, batched.  The reference draws every random number with torch calls on a
// seeded torch.Generator (ssar/random/patch.py:12-32, 41-66, 109-131; ssar/random/noise.py:15, 31, 46); on a CUDA
// generator each call is one tiny Philox kernel, ~430 per Patch.  These kernels issue the SAME Philox4x32-10 streams
// (same seed / subsequence / offset per element, same cuRAND transforms: this file includes cuRAND's own device
// headers) for MANY draws in one launch, so a seed gives the values the per-call sequence gives:
//   * ssar_philox_normal_segments — randn into a list of tensors (the noise bases of one Patch: 17 + up to 19 x 17
//     tensors), one launch instead of one per tensor;
//   * ssar_patch_sample — the whole config sampler of Patch.__init__ (randint / rand / exponential_ (multinomial) /
//     randn draws, whose count depends on two of the drawn values) for a batch of seeds, one warp per seed.
// The element -> (thread subsequence, call index, component) mapping and the per-call offset increment follow the
// launch geometry of ATen's nullary distribution kernels: block 256, grid min(ceil(numel / 256), SMs * (max threads per
// SM / 256)), unroll 4.  tests/test_gpu_philox.py checks both entry points bit for bit against the eager torch calls.
#include <curand_kernel.h>
#include <limits.h>

#include "common.cuh"

namespace {

constexpr int PB = 256;      // ATen's block size for nullary distribution kernels
constexpr int UNROLL = 4;    // curand4 engine calls per draw
constexpr int MAX_SEG = 96;  // segments per launch (descriptor array travels as a kernel parameter: 96 * 24 B)

struct SegBatch {
    float* dst[MAX_SEG];
    long long numel[MAX_SEG];
    unsigned long long offset[MAX_SEG];
    unsigned blk_start[MAX_SEG + 1];  // first CTA of every segment in the flattened grid (exclusive prefix of its blocks)
    int n;
};

__host__ __device__ inline unsigned seg_blocks(long long numel, unsigned max_blocks) {
    unsigned long long b = (unsigned long long)((numel + PB - 1) / PB);
    return (unsigned)(b < max_blocks ? b : max_blocks);
}

// randn: out[li] = (k-th component of the c-th curand_normal4 call of thread li mod S), S = 256 * blocks
__global__ void __launch_bounds__(PB) normal_segments_kernel(const SegBatch sb, unsigned long long seed, unsigned max_blocks) {
    // flattened grid: CTA -> (segment, block inside the segment's emulated ATen grid) by binary search over blk_start
    int lo = 0, hi = sb.n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (sb.blk_start[mid] <= blockIdx.x) lo = mid; else hi = mid;
    }
    const int s = lo;
    const long long numel = sb.numel[s];
    const unsigned blocks = seg_blocks(numel, max_blocks);
    const unsigned bx = blockIdx.x - sb.blk_start[s];
    const long long S = (long long)blocks * PB;
    const long long idx = (long long)bx * PB + threadIdx.x;
    curandStatePhilox4_32_10_t state;
    curand_init(seed, (unsigned long long)idx, sb.offset[s], &state);
    float* __restrict__ out = sb.dst[s];
    const long long rounded = ((numel - 1) / (S * UNROLL) + 1) * S * UNROLL;
    for (long long li = idx; li < rounded; li += S * UNROLL) {
        const float4 r = curand_normal4(&state);
        if (li < numel) out[li] = r.x;
        if (li + S < numel) out[li + S] = r.y;
        if (li + 2 * S < numel) out[li + 2 * S] = r.z;
        if (li + 3 * S < numel) out[li + 3 * S] = r.w;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Patch sampler.  A draw program is a list of ops (kind, n): every op is ONE torch call on a tensor of n <= 256
// elements, so element i is thread i's first engine call and the generator offset advances by 4 per op.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox_first(unsigned long long seed, unsigned idx, unsigned long long offset,
                                              curandStatePhilox4_32_10_t* st) {
    curand_init(seed, (unsigned long long)idx, offset, st);
}

// randint(lo, hi): ATen random_from_to, range < 2^32: (uint32 % range) + base
__device__ __forceinline__ int draw_randint(unsigned long long seed, unsigned long long offset, int lo, int hi) {
    curandStatePhilox4_32_10_t st;
    philox_first(seed, 0, offset, &st);
    const uint4 r = curand4(&st);
    return (int)(r.x % (unsigned)(hi - lo)) + lo;
}
// rand(): curand_uniform4 in (0, 1] -> value * (to - from) + from, the closed end folded to `from`
__device__ __forceinline__ float draw_rand(unsigned long long seed, unsigned long long offset) {
    curandStatePhilox4_32_10_t st;
    philox_first(seed, 0, offset, &st);
    const float4 u = curand_uniform4(&st);
    const float v = u.x * 1.0f + 0.0f;
    return v == 1.0f ? 0.0f : v;
}
__device__ __forceinline__ float draw_randn(unsigned long long seed, unsigned idx, unsigned long long offset) {
    curandStatePhilox4_32_10_t st;
    philox_first(seed, idx, offset, &st);
    const float4 r = curand_normal4(&st);
    return r.x * 1.0f + 0.0f;
}
// exponential_(1): ATen transformation::exponential on curand_uniform4's (0, 1] value
__device__ __forceinline__ float draw_exponential(unsigned long long seed, unsigned idx, unsigned long long offset) {
    curandStatePhilox4_32_10_t st;
    philox_first(seed, idx, offset, &st);
    const float4 u = curand_uniform4(&st);
    const float eps = 1.1920928955078125e-07f;  // std::numeric_limits<float>::epsilon()
    const float lg = (u.x >= 1.0f - eps / 2) ? -eps / 2 : logf(u.x);
    return (-1.0f / 1.0f) * lg;
}

__device__ void run_program(const int* __restrict__ kinds, const int* __restrict__ sizes, int n_ops, unsigned long long seed,
                            unsigned long long& offset, float* __restrict__ out, int& slot, int cap, int lane) {
    for (int o = 0; o < n_ops; ++o) {
        const int kind = kinds[o], n = sizes[o];
        if (kind == SSAR_DRAW_RANDN) {
            if (lane == 0 && slot < cap) out[slot] = draw_randn(seed, 0, offset);
            slot += 1;
        } else {  // SSAR_DRAW_EXPONENTIAL: n elements, element i on thread i
            for (int i = lane; i < n; i += 32)
                if (slot + i < cap) out[slot + i] = draw_exponential(seed, (unsigned)i, offset);
            slot += n;
        }
        offset += 4;
    }
}

// one warp per seed
__global__ void __launch_bounds__(128) patch_sample_kernel(const long long* __restrict__ seeds, int n_seeds,
                                                           const int* __restrict__ prog, int n_head, int n_lat_ops,
                                                           int n_noise_ops, int min_sub, int max_sub, int cap,
                                                           float* __restrict__ out, int* __restrict__ counts,
                                                           long long* __restrict__ end_offset) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n_seeds) return;
    const unsigned long long seed = (unsigned long long)seeds[w];
    float* o = out + (long long)w * cap;
    // program layout: [kinds | sizes] for head, latent config, noise config (concatenated)
    const int n_all = n_head + n_lat_ops + n_noise_ops;
    const int* kinds = prog;
    const int* sizes = prog + n_all;
    unsigned long long offset = 0;
    int slot = 0;
    // patch.py:41-43: num_base_latents = randint(3, 15); sigma_base_noise from rand(); loops_base_noise = choice(7)
    const int num_base = draw_randint(seed, offset, 3, 15);
    offset += 4;
    const float sigma_u = draw_rand(seed, offset);
    offset += 4;
    if (lane == 0) {
        o[0] = (float)num_base;
        o[1] = sigma_u;
    }
    slot = 2;
    run_program(kinds, sizes, n_head, seed, offset, o, slot, cap, lane);
    const int n_lat = draw_randint(seed, offset, min_sub, max_sub);
    offset += 4;
    for (int c = 0; c < n_lat; ++c) run_program(kinds + n_head, sizes + n_head, n_lat_ops, seed, offset, o, slot, cap, lane);
    const int n_noise = draw_randint(seed, offset, min_sub, max_sub);
    offset += 4;
    for (int c = 0; c < n_noise; ++c)
        run_program(kinds + n_head + n_lat_ops, sizes + n_head + n_lat_ops, n_noise_ops, seed, offset, o, slot, cap, lane);
    if (lane == 0) {
        counts[3 * w] = n_lat;
        counts[3 * w + 1] = n_noise;
        counts[3 * w + 2] = slot;  // floats produced (> cap means the caller's buffer was too small)
        end_offset[w] = (long long)offset;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// torch.randperm(n) on a CUDA generator, n <= 1024 (ATen randperm_out_cuda): int32 keys random_(INT_MIN, INT_MAX)
// (element i = the 64-bit value of thread i's first two engine words, modulo 2^32 - 1), stable ascending sort of the indices by the low `bits`
// key bits, then every run of equal keys is Fisher-Yates shuffled by its first thread from a second Philox state
// (subsequence = position of the run, offset = key offset + 4).  One CTA per permutation; the stable sort is a
// rank count (n^2 / CTA comparisons, n = 180 on the path).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) randperm_kernel(unsigned long long seed, unsigned long long first_offset,
                                                       unsigned long long stride, int n, int bits,
                                                       long long* __restrict__ out) {
    extern __shared__ int rp_sm[];
    int* key = rp_sm;           // masked keys in element order
    int* skey = rp_sm + n;      // masked keys in sorted order
    int* data = rp_sm + 2 * n;  // permutation
    const unsigned long long off = first_offset + stride * blockIdx.x;
    const unsigned mask = bits >= 32 ? 0xffffffffu : ((1u << bits) - 1u);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        curandStatePhilox4_32_10_t st;
        curand_init(seed, (unsigned long long)i, off, &st);
        const uint4 r = curand4(&st);
        // random_from_to with range = 2^32 - 1 takes ATen's 64-bit branch: two engine words form one 64-bit value
        const unsigned long long r64 = (unsigned long long)r.x | ((unsigned long long)r.y << 32);
        const long long v = (long long)(r64 % 4294967295ull) + (long long)INT_MIN;  // uniform_int_from_to(rand, range, base)
        key[i] = (int)((unsigned)(int)v & mask);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int k = key[i];
        int pos = 0;
        for (int j = 0; j < n; ++j) {
            const int kj = key[j];
            pos += (kj < k) || (kj == k && j < i);
        }
        skey[pos] = k;
        data[pos] = i;
    }
    __syncthreads();
    for (int tid = threadIdx.x; tid < n - 1; tid += blockDim.x) {
        if (skey[tid] != skey[tid + 1]) continue;               // not in a run
        if (tid != 0 && skey[tid] == skey[tid - 1]) continue;   // not the beginning of a run
        int island = 0;
        do { ++island; } while (tid + island < n && skey[tid + island] == skey[tid]);
        curandStatePhilox4_32_10_t st;
        curand_init(seed, (unsigned long long)tid, off + 4, &st);
        int* d = data + tid;
        for (int i = island - 1; i > 0; --i) {
            const unsigned r = curand(&st) % (unsigned)(i + 1);
            if ((unsigned)i != r) {
                const int tmp = d[i];
                d[i] = d[r];
                d[r] = tmp;
            }
        }
    }
    __syncthreads();
    long long* o = out + (long long)blockIdx.x * n;
    for (int i = threadIdx.x; i < n; i += blockDim.x) o[i] = (long long)data[i];
}

unsigned max_nullary_blocks() {
    static unsigned cached = 0;
    if (cached == 0) {
        int dev = 0, sms = 0, thr = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&thr, cudaDevAttrMaxThreadsPerMultiProcessor, dev);
        cached = (unsigned)(sms * (thr / PB));
    }
    return cached;
}

}  // namespace

extern "C" long long ssar_philox_normal_increment(long long numel) {
    if (numel <= 0) return 0;
    const unsigned blocks = seg_blocks(numel, max_nullary_blocks());
    return ((numel - 1) / ((long long)PB * blocks * UNROLL) + 1) * UNROLL;
}

extern "C" int ssar_philox_normal_segments(float* const* dst, const long long* numel, int n_segments, long long seed,
                                           long long first_offset, long long* end_offset, void* stream) {
    SSAR_REQUIRE(n_segments >= 0 && (n_segments == 0 || (dst && numel)), "philox_normal_segments: null pointer");
    SSAR_REQUIRE(first_offset >= 0 && first_offset % 4 == 0, "philox_normal_segments: offset must be a multiple of 4");
    const unsigned maxb = max_nullary_blocks();
    unsigned long long off = (unsigned long long)first_offset;
    int done = 0;
    while (done < n_segments) {
        SegBatch sb;
        sb.n = 0;
        unsigned gx = 0;
        while (done < n_segments && sb.n < MAX_SEG) {
            const long long ne = numel[done];
            SSAR_REQUIRE(ne >= 0, "philox_normal_segments: negative numel");
            if (ne > 0) {
                SSAR_REQUIRE(dst[done] != nullptr, "philox_normal_segments: null segment");
                sb.dst[sb.n] = dst[done];
                sb.numel[sb.n] = ne;
                sb.offset[sb.n] = off;
                sb.blk_start[sb.n] = gx;
                gx += seg_blocks(ne, maxb);
                off += (unsigned long long)ssar_philox_normal_increment(ne);
                ++sb.n;
            }
            ++done;
        }
        if (sb.n > 0) {
            sb.blk_start[sb.n] = gx;
            normal_segments_kernel<<<gx, PB, 0, (cudaStream_t)stream>>>(sb, (unsigned long long)seed, maxb);
            const int rc = ssar_launch_status("philox_normal_segments");
            if (rc != SSAR_OK) return rc;
        }
    }
    if (end_offset) *end_offset = (long long)off;
    return SSAR_OK;
}

extern "C" int ssar_patch_sample(const long long* seeds, int n_seeds, const int* program, int n_head_ops, int n_latent_ops,
                                 int n_noise_ops, int min_subpatches, int max_subpatches, int capacity, float* out,
                                 int* counts, long long* end_offset, void* stream) {
    if (n_seeds == 0) return SSAR_OK;
    SSAR_REQUIRE(seeds && program && out && counts && end_offset, "patch_sample: null pointer");
    SSAR_REQUIRE(max_subpatches > min_subpatches && min_subpatches >= 0, "patch_sample: bad sub-patch range");
    SSAR_REQUIRE(capacity >= 2, "patch_sample: capacity too small");
    const int warps_per_block = 4;
    patch_sample_kernel<<<ssar_div_up(n_seeds, warps_per_block), warps_per_block * 32, 0, (cudaStream_t)stream>>>(
        seeds, n_seeds, program, n_head_ops, n_latent_ops, n_noise_ops, min_subpatches, max_subpatches, capacity, out,
        counts, end_offset);
    return ssar_launch_status("patch_sample");
}

extern "C" int ssar_philox_randperm(long long seed, long long first_offset, long long offset_stride, int n, int bits,
                                    int count, long long* out, void* stream) {
    if (count == 0 || n == 0) return SSAR_OK;
    SSAR_REQUIRE(out != nullptr, "philox_randperm: null pointer");
    SSAR_REQUIRE(n > 0 && n <= 1024 && count > 0, "philox_randperm: n must be in [1, 1024]");
    SSAR_REQUIRE(bits > 0 && bits <= 32, "philox_randperm: bits must be in [1, 32]");
    SSAR_REQUIRE(first_offset >= 0 && first_offset % 4 == 0 && offset_stride % 4 == 0,
                 "philox_randperm: offsets must be multiples of 4");
    randperm_kernel<<<count, 256, 3 * n * sizeof(int), (cudaStream_t)stream>>>(
        (unsigned long long)seed, (unsigned long long)first_offset, (unsigned long long)offset_stride, n, bits, out);
    return ssar_launch_status("philox_randperm");
}
