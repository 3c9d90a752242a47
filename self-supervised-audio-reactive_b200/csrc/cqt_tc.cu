// Constant-Q response on the 5th-generation tensor cores (rosa/constantq.py:189-191, __cqt_response, all 7 octaves of
// rosa/constantq.py:77-100 in ONE launch).
//
// The reference forms, per octave i, C_i = fft_basis [36 x 513] . rfft(frames_i) [513 x T] with rectangular 1024-sample
// frames of the octave's decimated signal y_i (hop_i = hop / 2^i, centre reflect-padded).  The DFT and the (sparsified)
// basis are both linear, so in the TIME domain
//     C_i[t, r] = sum_n y_i[t hop_i - 512 + n] . b[r, n],      b[r, n] = sum_k fft_basis[r, k] e^{-2 pi i k n / 1024}
// is a dense K = 1024 contraction of the frame matrix with a [1024 x 72] real matrix (Re b, Im b interleaved) that is the
// same for every octave (sr and fmin halve together).  The frame matrix is never materialised: it is a 3-D TMA tensor map
// over the padded signal whose ROW STRIDE IS hop_i SAMPLES, i.e. overlapping rows (hop 16 .. 1024 against 1024-sample rows).
//
//   * operands: fp32 signal / basis split into fp16 (hi, lo) pairs, each scaled by a power of two into fp16's normal range
//     (signal: per clip from its own amax, basis: one constant), D = hi.hi + hi.lo + lo.hi with fp32 accumulation in TMEM
//     (the dropped lo.lo term is 2^-22 relative); the scales are undone in the epilogue.  cqt_split_kernel writes the
//     reflect-padded fp16 signals of all octaves in one launch.
//   * tcgen05.mma kind::f16, M = 128 frames x N = 80 (72 used) x K = 16 per instruction; 16 k-blocks of 64 samples per
//     tile, each a 4-stage TMA ring slot of (A hi | A lo | B hi | B lo) = 52 KB, 128-byte swizzle; accumulators double
//     buffered in TMEM so the epilogue of tile i overlaps the MMAs of tile i + 1.
//   * warp roles as in traj_mix_tc.cu: warp 0 TMA producer, warp 1 MMA issuer, warps 2..5 epilogue (tcgen05.ld, scale by
//     sqrt(2^i) / sqrt(filter length) (constantq.py:97, :113), 128-bit stores of the (re, im) pairs of a frame's 36 bins).
//   * persistent grid over (octave, clip, 128-frame tile) items; the basis panel (320 KB per tuning) stays in L2.
// Bound: L2 -> shared-memory operand traffic (832 KB per tile) and the tensor pipe are roughly balanced; HBM sees the
// fp16 signals once (2 x 2 B per sample and octave) and the [T, 252] complex output once.
#include "tc_common.cuh"

namespace {

constexpr int CQ_BM = 128, CQ_BN = 80, CQ_BK = 64, CQ_NFFT = 1024, CQ_ROWS = 36;
constexpr int CQ_STAGES = 4, CQ_MAX_OCT = 8;
constexpr int CQ_A_BYTES = CQ_BM * CQ_BK * 2;        // 16 KB
constexpr int CQ_B_BYTES = CQ_BN * CQ_BK * 2;        // 10 KB
constexpr int CQ_STAGE_BYTES = 2 * CQ_A_BYTES + 2 * CQ_B_BYTES;  // 52 KB, every part 1024-byte aligned
constexpr int CQ_THREADS = 192;
constexpr int CQ_ACC_COLS = 128;                      // TMEM columns reserved per accumulator (80 used)
constexpr uint32_t CQ_IDESC = (1u << 4) | ((uint32_t)(CQ_BN >> 3) << 17) | ((uint32_t)(CQ_BM >> 4) << 24);

struct CqtMaps {
    CUtensorMap a[CQ_MAX_OCT][2];  // per octave: (hi, lo) frame views [clip][frame][sample] with overlapping rows
    CUtensorMap b[2];              // basis panels (hi, lo) [slot * 80 + row][1024]
};

struct CqtOctaves {
    int n_oct;
    long long n_samples[CQ_MAX_OCT];  // per clip
    long long padded[CQ_MAX_OCT];     // row length of the split buffers (samples)
    long long offset[CQ_MAX_OCT];     // start of the octave's [n_clips, padded] block inside the hi (and lo) buffer
    const float* signal[CQ_MAX_OCT];  // dense [n_clips, n_samples]
};

// power-of-two scale that brings a clip's amax into [128, 256) (fp16 hi AND lo halves stay normal numbers down to
// ~1e-3 of the clip's peak, with 2^4 of headroom for the sqrt(2) per octave gain and the decimator's overshoot)
__device__ __forceinline__ float clip_scale(uint32_t amax_bits) {
    const float amax = __uint_as_float(amax_bits);
    if (!(amax > 0.f) || !isfinite(amax)) return 1.f;
    return exp2f((float)(7 - ilogbf(amax)));
}

__global__ void cqt_amax_kernel(const float* __restrict__ x, long long n, uint32_t* __restrict__ amax) {
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * n;
    float m = 0.f;
    const long long head = min(n, (long long)((4 - ((reinterpret_cast<uintptr_t>(xc) >> 2) & 3)) & 3));  // to 16-byte alignment
    const long long n4 = (n - head) >> 2;
    const float4* __restrict__ x4 = reinterpret_cast<const float4*>(xc + head);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 v = __ldg(x4 + i);
        m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
    }
    if (blockIdx.x == 0 && threadIdx.x < 8) {  // the unaligned head and the tail
        const long long i = threadIdx.x < 4 ? threadIdx.x : head + 4 * n4 + (threadIdx.x - 4);
        if ((threadIdx.x < 4 && i < head) || (threadIdx.x >= 4 && i < n)) m = fmaxf(m, fabsf(__ldg(xc + i)));
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0 && __float_as_uint(m) > __ldcg(amax + clip)) atomicMax(amax + clip, __float_as_uint(m));
}

// y_i [clip, n] fp32 -> hi / lo fp16 [clip, padded]: sample j of a row is y_i[reflect(j - 512)] * scale (zeros beyond
// n + 1024), i.e. torch.stft's centre padding baked in, so frame t starts at row offset t * hop_i.
// Four samples per thread (64-bit stores); the interior takes a path without the reflection arithmetic.
__global__ void cqt_split_kernel(const CqtOctaves oc, int n_clips, const uint32_t* __restrict__ amax,
                                 __half* __restrict__ hi, __half* __restrict__ lo, const long long* __restrict__ clip_samples) {
    const int o = blockIdx.z, clip = blockIdx.y;
    const long long lp = oc.padded[o];
    const float* __restrict__ src = oc.signal[o] + (long long)clip * oc.n_samples[o];
    // ragged batches: the clip's own length at this octave (each decimation halves, rounding up) — the reflection
    // happens at ITS end
    long long n = oc.n_samples[o];
    if (clip_samples) {
        n = __ldg(clip_samples + clip);
        for (int q = 0; q < o; ++q) n = (n + 1) >> 1;
    }
    const float s = clip_scale(__ldg(amax + clip));
    __half* h = hi + oc.offset[o] + (long long)clip * lp;
    __half* l = lo + oc.offset[o] + (long long)clip * lp;
    constexpr int HALF = CQ_NFFT / 2;
    const long long lp4 = lp >> 2;  // padded rows are multiples of 1024 samples
    for (long long j4 = (long long)blockIdx.x * blockDim.x + threadIdx.x; j4 < lp4; j4 += (long long)gridDim.x * blockDim.x) {
        const long long j = 4 * j4;
        float v[4];
        if (j >= HALF && j + 3 < n + HALF) {
            const float* p = src + (j - HALF);
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldg(p + u) * s;
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = (j + u < n + CQ_NFFT) ? __ldg(src + ssar_reflect(j + u - HALF, n)) * s : 0.f;
        }
        __half hh[4], ll[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) split_half(v[u], hh[u], ll[u]);
        __half2 h01 = __halves2half2(hh[0], hh[1]), h23 = __halves2half2(hh[2], hh[3]);
        __half2 l01 = __halves2half2(ll[0], ll[1]), l23 = __halves2half2(ll[2], ll[3]);
        *reinterpret_cast<uint2*>(h + j) = make_uint2(*reinterpret_cast<uint32_t*>(&h01), *reinterpret_cast<uint32_t*>(&h23));
        *reinterpret_cast<uint2*>(l + j) = make_uint2(*reinterpret_cast<uint32_t*>(&l01), *reinterpret_cast<uint32_t*>(&l23));
    }
}

struct CqShared {
    uint64_t full[CQ_STAGES], empty[CQ_STAGES];
    uint64_t acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(CQ_THREADS, 1)
cqt_tc_kernel(const __grid_constant__ CqtMaps maps, int n_oct, int n_clips, int n_frames, const int* __restrict__ clip_slot,
              const uint32_t* __restrict__ amax, const float* __restrict__ row_scale, long long row_scale_ld,
              float basis_scale_inv, float* __restrict__ out, const int* __restrict__ clip_frames) {
    extern __shared__ __align__(1024) uint8_t cq_smem_raw[];
    uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(cq_smem_raw) + 1023) & ~(uintptr_t)1023);
    CqShared* sh = reinterpret_cast<CqShared*>(ring + (size_t)CQ_STAGES * CQ_STAGE_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (n_frames + CQ_BM - 1) / CQ_BM;
    const int n_items = n_oct * n_clips * m_tiles;
    const int n_bins = n_oct * CQ_ROWS;
    constexpr int KB = CQ_NFFT / CQ_BK;  // 16 k-blocks per tile

    if (threadIdx.x == 0) {
        for (int s = 0; s < CQ_STAGES; ++s) {
            mbar_init(&sh->full[s], 1);
            mbar_init(&sh->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sh->acc_full[a], 1);
            mbar_init(&sh->acc_empty[a], 4);  // one arrive per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sh->tmem_base)), "n"(256)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = sh->tmem_base;

    if (warp == 0) {
        // ===================================================================== TMA producer
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int mt = item % m_tiles, oc = item / m_tiles;
                const int clip = oc % n_clips, o = oc / n_clips;
                const int slot = clip_slot ? __ldg(clip_slot + clip) : 0;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(&sh->empty[stage], phase ^ 1);
                    mbar_expect_tx(&sh->full[stage], CQ_STAGE_BYTES);
                    uint8_t* dst = ring + (size_t)stage * CQ_STAGE_BYTES;
                    tma_load_3d(dst, &maps.a[o][0], &sh->full[stage], kb * CQ_BK, mt * CQ_BM, clip);
                    tma_load_3d(dst + CQ_A_BYTES, &maps.a[o][1], &sh->full[stage], kb * CQ_BK, mt * CQ_BM, clip);
                    tma_load_2d(dst + 2 * CQ_A_BYTES, &maps.b[0], &sh->full[stage], kb * CQ_BK, slot * CQ_BN);
                    tma_load_2d(dst + 2 * CQ_A_BYTES + CQ_B_BYTES, &maps.b[1], &sh->full[stage], kb * CQ_BK, slot * CQ_BN);
                    if (++stage == CQ_STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (one thread)
        if (lane == 0) {
            uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                mbar_wait(&sh->acc_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * CQ_ACC_COLS;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(&sh->full[stage], phase);
                    tc_fence_after();
                    const uint32_t ah = smem_u32(ring + (size_t)stage * CQ_STAGE_BYTES);
                    const uint32_t al = ah + CQ_A_BYTES, bh = ah + 2 * CQ_A_BYTES, bl = bh + CQ_B_BYTES;
#pragma unroll
                    for (int ks = 0; ks < CQ_BK / 16; ++ks) {
                        const uint32_t ko = ks * 32;  // 16 fp16 = 32 bytes along the swizzled row
                        const uint64_t dah = umma_desc_sw128(ah + ko), dal = umma_desc_sw128(al + ko);
                        const uint64_t dbh = umma_desc_sw128(bh + ko), dbl = umma_desc_sw128(bl + ko);
                        umma_f16(d_tmem, dah, dbh, CQ_IDESC, (kb | ks) != 0);
                        umma_f16(d_tmem, dah, dbl, CQ_IDESC, 1);
                        umma_f16(d_tmem, dal, dbh, CQ_IDESC, 1);
                    }
                    umma_commit(&sh->empty[stage]);  // frees the ring slot when these MMAs retire
                    if (++stage == CQ_STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&sh->acc_full[acc]);
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else {
        // ===================================================================== epilogue warps 2..5
        const int quad = warp & 3;  // TMEM lane quadrant this warp may read
        uint32_t acc = 0, acc_phase = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int mt = item % m_tiles, oc = item / m_tiles;
            const int clip = oc % n_clips, o = oc / n_clips;
            const int slot = clip_slot ? __ldg(clip_slot + clip) : 0;
            const int col0 = n_bins - CQ_ROWS * (o + 1);  // octave o fills bins [col0, col0 + 36) (constantq.py:100-104)
            const int t = mt * CQ_BM + quad * 32 + lane;
            // sqrt(2^o) (basis gain, :97) / sqrt(filter length) (:113), and the operand scales undone
            // (ragged batches: frames past the clip's own count are written as zeros)
            const float gain = (clip_frames && t >= __ldg(clip_frames + clip))
                                   ? 0.f
                                   : sqrtf((float)(1 << o)) * basis_scale_inv / clip_scale(__ldg(amax + clip));
            const float* __restrict__ rs = row_scale + (long long)slot * row_scale_ld + col0;
            mbar_wait(&sh->acc_full[acc], acc_phase);
            tc_fence_after();
            float4* dst = reinterpret_cast<float4*>(out + (((long long)clip * n_frames + t) * n_bins + col0) * 2);
#pragma unroll 1
            for (int c = 0; c < 72 / 8; ++c) {  // 8 accumulator columns = 4 complex bins per step
                uint32_t r[8];
                const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * CQ_ACC_COLS + c * 8;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                             : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                             : "r"(taddr)
                             : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (t < n_frames) {
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const float s0 = gain * __ldg(rs + 4 * c + 2 * q), s1 = gain * __ldg(rs + 4 * c + 2 * q + 1);
                        dst[2 * c + q] = make_float4(__uint_as_float(r[4 * q]) * s0, __uint_as_float(r[4 * q + 1]) * s0,
                                                     __uint_as_float(r[4 * q + 2]) * s1, __uint_as_float(r[4 * q + 3]) * s1);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sh->acc_empty[acc]);
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(256) : "memory");
    }
}

inline long long cq_padded(long long n_samples) { return (n_samples + CQ_NFFT + 1023) / 1024 * 1024; }

}  // namespace

extern "C" long long ssar_cqt_tc_scratch_bytes(int n_clips, const long long* n_samples, int n_octaves) {
    long long total = 0;
    for (int o = 0; o < n_octaves; ++o) total += (long long)n_clips * cq_padded(n_samples[o]);
    return 1024LL * ((n_clips * 4 + 1023) / 1024) + 2 * total * 2 + 2048;  // per-clip amax words, hi + lo fp16 buffers, slack
}

extern "C" int ssar_cqt_tc(const float* const* signals, const long long* n_samples, int n_octaves, int n_clips, int hop,
                           int n_frames, const void* basis_hi, const void* basis_lo, int n_slots, const int* clip_slot,
                           const float* row_scale, long long row_scale_ld, float basis_scale, void* scratch, float* out,
                           const long long* clip_samples, const int* clip_frames, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(signals && n_samples && basis_hi && basis_lo && row_scale && scratch && out, "cqt_tc: null pointer");
    SSAR_REQUIRE(n_octaves >= 1 && n_octaves <= CQ_MAX_OCT, "cqt_tc: 1..%d octaves", CQ_MAX_OCT);
    SSAR_REQUIRE(n_clips <= 1024 || clip_slot, "cqt_tc: bad clip count");
    SSAR_REQUIRE(hop % (8 << (n_octaves - 1)) == 0, "cqt_tc: hop must be a multiple of 8 * 2^(octaves - 1) (16-byte TMA row strides)");
    SSAR_REQUIRE(((uintptr_t)out & 15) == 0, "cqt_tc: out must be 16-byte aligned");
    SSAR_REQUIRE(basis_scale > 0.f && n_slots >= 1, "cqt_tc: bad basis scale / slot count");
    cudaStream_t st = (cudaStream_t)stream;
    CqtOctaves oc;
    oc.n_oct = n_octaves;
    long long total = 0;
    for (int o = 0; o < n_octaves; ++o) {
        const int hop_o = hop >> o;
        SSAR_REQUIRE(signals[o] != nullptr && n_samples[o] > CQ_NFFT / 2, "cqt_tc: octave %d signal missing or shorter than 513 samples", o);
        SSAR_REQUIRE((long long)(n_frames - 1) * hop_o <= n_samples[o], "cqt_tc: n_frames=%d exceeds octave %d's length", n_frames, o);
        oc.n_samples[o] = n_samples[o];
        oc.padded[o] = cq_padded(n_samples[o]);
        oc.offset[o] = total;
        oc.signal[o] = signals[o];
        total += (long long)n_clips * oc.padded[o];
    }
    uint8_t* sb = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(scratch) + 1023) & ~(uintptr_t)1023);
    uint32_t* amax = reinterpret_cast<uint32_t*>(sb);
    __half* hi = reinterpret_cast<__half*>(sb + 1024 * ((n_clips * 4 + 1023) / 1024));
    __half* lo = hi + total;

    cudaError_t e = cudaMemsetAsync(amax, 0, sizeof(uint32_t) * n_clips, st);
    if (e != cudaSuccess) {
        ssar_set_error("cqt_tc: memset: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    {
        int gx = ssar_div_up(n_samples[0], 256 * 16);
        if (gx > SSAR_NUM_SMS * 2) gx = SSAR_NUM_SMS * 2;
        cqt_amax_kernel<<<dim3(gx, n_clips), 256, 0, st>>>(signals[0], n_samples[0], amax);
        int sx = ssar_div_up(oc.padded[0] / 4, 256);
        if (sx > SSAR_NUM_SMS * 8) sx = SSAR_NUM_SMS * 8;
        cqt_split_kernel<<<dim3(sx, n_clips, n_octaves), 256, 0, st>>>(oc, n_clips, amax, hi, lo, clip_samples);
    }
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        ssar_set_error("cqt_tc: cuTensorMapEncodeTiled unavailable");
        return SSAR_ERR_CUDA;
    }
    CqtMaps maps;
    memset(&maps, 0, sizeof(maps));
    const cuuint32_t estr[3] = {1, 1, 1};
    for (int o = 0; o < n_octaves; ++o) {
        const int hop_o = hop >> o;
        // frames[clip][t][n] = padded[clip][t * hop_o + n]: rows overlap (stride hop_o samples, extent 1024 samples)
        const cuuint64_t dims[3] = {(cuuint64_t)CQ_NFFT, (cuuint64_t)n_frames, (cuuint64_t)n_clips};
        const cuuint64_t strides[2] = {(cuuint64_t)hop_o * 2, (cuuint64_t)oc.padded[o] * 2};
        const cuuint32_t box[3] = {(cuuint32_t)CQ_BK, (cuuint32_t)CQ_BM, 1};
        for (int p = 0; p < 2; ++p) {
            void* base = (p == 0 ? hi : lo) + oc.offset[o];
            if (enc(&maps.a[o][p], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
                ssar_set_error("cqt_tc: cuTensorMapEncodeTiled failed for octave %d", o);
                return SSAR_ERR_CUDA;
            }
        }
    }
    {
        const cuuint64_t dims[2] = {(cuuint64_t)CQ_NFFT, (cuuint64_t)n_slots * CQ_BN};
        const cuuint64_t strides[1] = {(cuuint64_t)CQ_NFFT * 2};
        const cuuint32_t box[2] = {(cuuint32_t)CQ_BK, (cuuint32_t)CQ_BN};
        for (int p = 0; p < 2; ++p) {
            if (enc(&maps.b[p], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(p == 0 ? basis_hi : basis_lo), dims, strides, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
                ssar_set_error("cqt_tc: cuTensorMapEncodeTiled failed for the basis");
                return SSAR_ERR_CUDA;
            }
        }
    }
    const size_t smem = (size_t)CQ_STAGES * CQ_STAGE_BYTES + sizeof(CqShared) + 1024;
    e = cudaFuncSetAttribute(cqt_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("cqt_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    const long long items = (long long)n_octaves * n_clips * ((n_frames + CQ_BM - 1) / CQ_BM);
    const int grid = (int)(items < SSAR_NUM_SMS ? items : SSAR_NUM_SMS);
    cqt_tc_kernel<<<grid, CQ_THREADS, smem, st>>>(maps, n_octaves, n_clips, n_frames, clip_slot, amax, row_scale, row_scale_ld,
                                                 1.f / basis_scale, out, clip_frames);
    return ssar_launch_status("cqt_tc");
}
