// Trajectory mix on the 5th-generation tensor cores: out[t, n] = sum_k W_{g(n)}[t, k] * palette[k, n]  (the one dense
// contraction of the path: einsum("TN,NWL->TWL") and everything LatentWeights folds into it, ssar/random/latent.py:7-78).
//
//   * tcgen05.mma kind::f16, M = 128 x N = 128 x K = 16 per instruction, fp32 accumulators in TMEM (2 x 128 columns,
//     double buffered so the epilogue of tile i overlaps the MMAs of tile i + 1);
//   * fp32 accuracy from fp16 operands by a 3-term split: a = a_hi + a_lo, b = b_hi + b_lo (a_hi = fp16(a),
//     a_lo = fp16(a - a_hi)), D = a_hi b_hi + a_hi b_lo + a_lo b_hi with fp32 accumulation — the dropped a_lo b_lo term
//     is 2^-22 relative; measured rel-L2 vs the fp32 CUDA-core kernel ~1e-7 (tests/test_gpu_traj.py);
//   * operands are pre-split once per call into K-major fp16 panels (K padded to a multiple of 64) and fetched by TMA
//     (cp.async.bulk.tensor.2d, 128-byte swizzle) straight into the UMMA shared-memory layout;
//   * A-STATIONARY: a CTA owns a (128-row block, layer group) pair, keeps that A panel (hi + lo, <= 96 KB) in shared
//     memory and streams the palette columns of the group through a 3-stage ring (32 KB per stage), so A is read once
//     and the palette panel (7 MB, L2 resident) once per row block;
//   * warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM allocator), warps 2..5 = epilogue
//     (tcgen05.ld 32x32b -> registers -> swizzled shared tile -> TMA store), all synchronised with mbarriers only.
// The kernel is write bound: 36.9 kB of output per frame against 2 * 3 * K * 9216 tensor FLOP.
#include "block_utils.cuh"
#include "tc_common.cuh"

namespace {

constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 64;      // CTA tile; BK fp16 = one 128-byte swizzle row
constexpr int TC_STAGES = 3;                              // palette ring
constexpr int TC_MAX_KB = 3;                              // K <= 192 (A panel 96 KB)
constexpr int TC_TILE_BYTES = TC_BM * TC_BK * 2;          // 16 KB: one 128 x 64 fp16 tile (A or B, hi or lo)
constexpr int TC_THREADS = 192;
constexpr int TC_CSTAGE_BYTES = 32 * 32 * 4;              // one [32 x 32] fp32 store tile
constexpr int TC_CBUFS = 2;                               // store tiles per epilogue warp (the next chunk is staged while the
                                                          // TMA unit still reads the previous one)

// ---------------------------------------------------------------------------------------------------------------
// operand split: fp32 -> (hi, lo) fp16, K-major, K zero-padded to KP
// ---------------------------------------------------------------------------------------------------------------
// Power-of-two operand scale from the operand's amax: brings it into [2^9, 2^10), so hi AND lo stay normal fp16 numbers
// for every element down to 2^-12 of the peak (no overflow for operands >= 65504, no subnormal lo terms for tiny ones).
__device__ __forceinline__ float operand_scale(uint32_t amax_bits) {
    const float amax = __uint_as_float(amax_bits);
    if (!(amax > 0.f) || !isfinite(amax)) return 1.f;
    return exp2f((float)(9 - ilogbf(amax)));
}

__global__ void basis_amax_kernel(const float* __restrict__ B, long long ldb, int P, int NC, uint32_t* __restrict__ amax) {
    const long long n = (long long)P * NC;
    float m = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        m = fmaxf(m, fabsf(__ldg(B + (i / NC) * ldb + (i % NC))));
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0 && __float_as_uint(m) > __ldcg(amax)) atomicMax(amax, __float_as_uint(m));
}

// W [T, G, P] (row stride ldw, group stride gs) -> Ah / Al [G][Tpad][KP].  One WARP per (group, row): the row's own amax
// gives its power-of-two scale (amax bits written to scales[g * Tpad + t]; the epilogue, where a lane owns an output
// row, undoes it), so there is no separate reduction pass, no atomic, and every row keeps full fp16 x 2 precision
// whatever the other rows hold.  Rows beyond T are written as zeros.
__global__ void __launch_bounds__(256) split_weights_kernel(const float* __restrict__ W, long long ldw, long long gs, int T,
                                                            int Tpad, int G, int P, int KP, uint32_t* __restrict__ scales,
                                                            __half* __restrict__ Ah, __half* __restrict__ Al) {
    const int lane = threadIdx.x & 31;
    const long long n_rows = (long long)G * Tpad;
    for (long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n_rows;
         r += (long long)gridDim.x * (blockDim.x >> 5)) {
        const int g = (int)(r / Tpad), t = (int)(r - (long long)g * Tpad);
        const float* __restrict__ row = W + (long long)t * ldw + (long long)g * gs;
        float v[6];  // KP <= 192: six values per lane
        float m = 0.f;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const int k = lane + 32 * j;
            v[j] = (t < T && k < P) ? __ldg(row + k) : 0.f;
            m = fmaxf(m, fabsf(v[j]));
        }
        m = warp_max(m);
        if (lane == 0) scales[r] = __float_as_uint(m);
        const float sc = operand_scale(__float_as_uint(m));
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            const int k = lane + 32 * j;
            if (k < KP) {
                __half h, l;
                split_half(v[j] * sc, h, l);
                Ah[r * KP + k] = h;
                Al[r * KP + k] = l;
            }
        }
    }
}

// palette [P, NC] (row stride ldb) -> Bh / Bl [NC][KP] (transposed through a 32 x 32 shared tile), scaled likewise
__global__ void split_palette_kernel(const float* __restrict__ B, long long ldb, int P, int NC, int KP,
                                     const uint32_t* __restrict__ amax, __half* __restrict__ Bh, __half* __restrict__ Bl) {
    __shared__ float tile[32][33];
    const float sc = operand_scale(__ldg(amax));
    const int n0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int k = k0 + r, n = n0 + threadIdx.x;
        tile[r][threadIdx.x] = (k < P && n < NC) ? __ldg(B + (long long)k * ldb + n) * sc : 0.f;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int n = n0 + r, k = k0 + threadIdx.x;
        if (n < NC && k < KP) {
            __half h, l;
            split_half(tile[threadIdx.x][r], h, l);
            Bh[(long long)n * KP + k] = h;
            Bl[(long long)n * KP + k] = l;
        }
    }
}

// kind::f16: D = F32 (bit 4), A = B = F16 (0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t TC_IDESC = (1u << 4) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

struct TcShared {
    uint64_t a_full, a_empty;
    uint64_t b_full[TC_STAGES], b_empty[TC_STAGES];
    uint64_t acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(TC_THREADS, 1)
mix_gemm_tc_kernel(const __grid_constant__ CUtensorMap map_ah, const __grid_constant__ CUtensorMap map_al,
                   const __grid_constant__ CUtensorMap map_bh, const __grid_constant__ CUtensorMap map_bl,
                   const __grid_constant__ CUtensorMap map_out, int Tpad, int n_groups, int group_cols, int KB, int n_splits,
                   const uint32_t* __restrict__ amax_w, const uint32_t* __restrict__ amax_b) {
    extern __shared__ __align__(1024) uint8_t tc_smem_raw[];
    // carve: [A panel: KB x (hi 16 KB | lo 16 KB)] [B ring: STAGES x (hi 16 KB | lo 16 KB)] [barriers]
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* a_panel = base;
    uint8_t* b_ring = a_panel + (size_t)TC_MAX_KB * 2 * TC_TILE_BYTES;
    uint8_t* c_stage = b_ring + (size_t)TC_STAGES * 2 * TC_TILE_BYTES;  // 4 epilogue warps x [32 rows x 128 B], swizzled
    TcShared* sh = reinterpret_cast<TcShared*>(c_stage + 4 * TC_CBUFS * TC_CSTAGE_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = Tpad / TC_BM;
    // work item = (layer group, 128-row block, column split): the A panel stays resident while the item walks its share
    // of the group's 128-column tiles (column splits only exist to fill the SMs when there are few row blocks)
    const int tiles_per_group = group_cols / TC_BN;
    const int n_tiles = (tiles_per_group + n_splits - 1) / n_splits;  // per item (the last split may have fewer)
    const int n_items = m_tiles * n_groups * n_splits;

    if (threadIdx.x == 0) {
        mbar_init(&sh->a_full, 1);
        mbar_init(&sh->a_empty, 1);
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&sh->b_full[s], 1);
            mbar_init(&sh->b_empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&sh->acc_full[a], 1);
            mbar_init(&sh->acc_empty[a], 4);  // one arrive per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // TMEM: 2 accumulators x 128 columns
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
            uint32_t stage = 0, b_phase = 0, a_phase = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int sp = item % n_splits, gm = item / n_splits;
                const int g = gm / m_tiles, mt = gm - g * m_tiles;
                const int row0 = g * Tpad + mt * TC_BM;
                const int nt0 = sp * n_tiles, nt1 = min(nt0 + n_tiles, tiles_per_group);
                mbar_wait(&sh->a_empty, a_phase ^ 1);
                mbar_expect_tx(&sh->a_full, (uint32_t)KB * 2 * TC_TILE_BYTES);
                for (int kb = 0; kb < KB; ++kb) {
                    tma_load_2d(a_panel + (size_t)(2 * kb) * TC_TILE_BYTES, &map_ah, &sh->a_full, kb * TC_BK, row0);
                    tma_load_2d(a_panel + (size_t)(2 * kb + 1) * TC_TILE_BYTES, &map_al, &sh->a_full, kb * TC_BK, row0);
                }
                a_phase ^= 1;
                for (int nt = nt0; nt < nt1; ++nt) {
                    const int col0 = g * group_cols + nt * TC_BN;
                    for (int kb = 0; kb < KB; ++kb) {
                        mbar_wait(&sh->b_empty[stage], b_phase ^ 1);
                        mbar_expect_tx(&sh->b_full[stage], 2 * TC_TILE_BYTES);
                        uint8_t* dst = b_ring + (size_t)stage * 2 * TC_TILE_BYTES;
                        tma_load_2d(dst, &map_bh, &sh->b_full[stage], kb * TC_BK, col0);
                        tma_load_2d(dst + TC_TILE_BYTES, &map_bl, &sh->b_full[stage], kb * TC_BK, col0);
                        if (++stage == TC_STAGES) {
                            stage = 0;
                            b_phase ^= 1;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (one thread)
        if (lane == 0) {
            uint32_t stage = 0, b_phase = 0, a_phase = 0, acc = 0, acc_phase = 0;
            for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
                const int sp = item % n_splits;
                const int nt0 = sp * n_tiles, nt1 = min(nt0 + n_tiles, tiles_per_group);
                mbar_wait(&sh->a_full, a_phase);
                a_phase ^= 1;
                tc_fence_after();
                for (int nt = nt0; nt < nt1; ++nt) {
                    mbar_wait(&sh->acc_empty[acc], acc_phase ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * TC_BN;
                    for (int kb = 0; kb < KB; ++kb) {
                        mbar_wait(&sh->b_full[stage], b_phase);
                        tc_fence_after();
                        const uint32_t ah = smem_u32(a_panel + (size_t)(2 * kb) * TC_TILE_BYTES);
                        const uint32_t al = ah + TC_TILE_BYTES;
                        const uint32_t bh = smem_u32(b_ring + (size_t)stage * 2 * TC_TILE_BYTES);
                        const uint32_t bl = bh + TC_TILE_BYTES;
#pragma unroll
                        for (int ks = 0; ks < TC_BK / 16; ++ks) {
                            const uint32_t ko = ks * 32;  // 16 fp16 = 32 bytes along the swizzled row
                            const uint64_t dah = umma_desc_sw128(ah + ko), dal = umma_desc_sw128(al + ko);
                            const uint64_t dbh = umma_desc_sw128(bh + ko), dbl = umma_desc_sw128(bl + ko);
                            umma_f16(d_tmem, dah, dbh, TC_IDESC, (kb | ks) != 0);
                            umma_f16(d_tmem, dah, dbl, TC_IDESC, 1);
                            umma_f16(d_tmem, dal, dbh, TC_IDESC, 1);
                        }
                        umma_commit(&sh->b_empty[stage]);  // frees the ring slot when these MMAs retire
                        if (++stage == TC_STAGES) {
                            stage = 0;
                            b_phase ^= 1;
                        }
                    }
                    umma_commit(&sh->acc_full[acc]);  // accumulator complete -> epilogue
                    if (++acc == 2) {
                        acc = 0;
                        acc_phase ^= 1;
                    }
                }
                umma_commit(&sh->a_empty);  // the A panel may be overwritten once every MMA of this item retired
            }
        }
    } else {
        // ===================================================================== epilogue warps 2..5
        const int quad = warp & 3;  // TMEM lane quadrant this warp may read
        uint8_t* stage_base = c_stage + (size_t)quad * TC_CBUFS * TC_CSTAGE_BYTES;
        uint32_t cbuf = 0;
        const float unscale_b = 1.f / operand_scale(__ldg(amax_b));  // powers of two: exact
        uint32_t acc = 0, acc_phase = 0;
        for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int sp = item % n_splits, gm = item / n_splits;
            const int g = gm / m_tiles, mt = gm - g * m_tiles;
            const int nt0 = sp * n_tiles, nt1 = min(nt0 + n_tiles, tiles_per_group);
            const int row0 = mt * TC_BM + quad * 32;  // first output row of this warp's 32-lane slice
            // this lane's output row carries its own operand scale (rows past the tensor's end read a zero-row scale of 1)
            const float unscale = unscale_b / operand_scale(__ldg(amax_w + (long long)g * Tpad + row0 + lane));
            for (int nt = nt0; nt < nt1; ++nt) {
                mbar_wait(&sh->acc_full[acc], acc_phase);
                tc_fence_after();
                const int col0 = g * group_cols + nt * TC_BN;
#pragma unroll 1
                for (int c = 0; c < TC_BN / 32; ++c) {
                    uint32_t r[32];
                    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * TC_BN + c * 32;
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                        : "r"(taddr)
                        : "memory");
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    // the TMA store issued TC_CBUFS chunks ago must have finished READING this staging tile before it is
                    // rewritten (the last one may still be in flight)
                    uint8_t* stage_buf = stage_base + (size_t)cbuf * TC_CSTAGE_BYTES;
                    cbuf = (cbuf + 1) % TC_CBUFS;
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(TC_CBUFS - 1) : "memory");
                    __syncwarp();
                    // lane = row: 32 floats = one 128-byte row of the [32 x 32] tile, 16-byte chunks XOR-swizzled by (row % 8)
                    // (the layout a SWIZZLE_128B tensor map expects; 4-way instead of 32-way bank conflicts)
                    uint8_t* rowp = stage_buf + lane * 128;
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        *reinterpret_cast<float4*>(rowp + ((q ^ (lane & 7)) << 4)) =
                            make_float4(__uint_as_float(r[4 * q]) * unscale, __uint_as_float(r[4 * q + 1]) * unscale,
                                        __uint_as_float(r[4 * q + 2]) * unscale, __uint_as_float(r[4 * q + 3]) * unscale);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) {  // rows beyond the tensor's extent are clipped by the TMA unit
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                                         reinterpret_cast<uint64_t>(&map_out)),
                                     "r"(smem_u32(stage_buf)), "r"(col0 + c * 32), "r"(row0)
                                     : "memory");
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
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
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // all output writes complete
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(256) : "memory");
    }
}

// fp16 [rows, KP] K-major panel, box 64 (K) x 128 (rows), 128-byte swizzle
bool make_panel_map(CUtensorMap* map, const void* ptr, long long rows, int KP) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)KP, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)KP * 2};
    const cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)TC_BM};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// fp32 [rows, ld] output, box 32 (columns) x 32 (rows), 128-byte swizzle on the shared-memory side
bool make_out_map(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {32, 32};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

static inline long long tc_kp(int n_basis) { return ((long long)n_basis + TC_BK - 1) / TC_BK * TC_BK; }
static inline long long tc_tpad(int n_rows) { return ((long long)n_rows + TC_BM - 1) / TC_BM * TC_BM; }
static inline long long tc_scale_bytes(int n_groups, long long Tpad) {
    return ((long long)n_groups * Tpad * 4 + 1023) / 1024 * 1024;  // one amax word per (group, row)
}

// split basis block: [amax word, padded to 1024 B][Bh [n_cols][KP]][Bl [n_cols][KP]]
extern "C" long long ssar_mix_gemm_tc_basis_bytes(int n_basis, int n_cols) {
    return 1024 + 2 * (long long)n_cols * tc_kp(n_basis) * 2 + 1024;
}

static int tc_split_basis(const float* basis, long long ldb, int n_basis, int n_cols, uint8_t* blk, cudaStream_t st) {
    const int KP = (int)tc_kp(n_basis);
    uint32_t* amax = reinterpret_cast<uint32_t*>(blk);
    __half* Bh = reinterpret_cast<__half*>(blk + 1024);
    __half* Bl = Bh + (long long)n_cols * KP;
    cudaError_t e = cudaMemsetAsync(amax, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) {
        ssar_set_error("mix_gemm_tc: memset: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    int grid = ssar_div_up((long long)n_basis * n_cols, 256 * 8);
    if (grid > SSAR_NUM_SMS * 4) grid = SSAR_NUM_SMS * 4;
    basis_amax_kernel<<<grid, 256, 0, st>>>(basis, ldb, n_basis, n_cols, amax);
    split_palette_kernel<<<dim3(ssar_div_up(n_cols, 32), ssar_div_up(KP, 32)), dim3(32, 8), 0, st>>>(basis, ldb, n_basis, n_cols,
                                                                                                  KP, amax, Bh, Bl);
    return ssar_launch_status("mix_gemm_tc_split_basis");
}

extern "C" int ssar_mix_gemm_tc_split_basis(const float* basis, long long ldb, int n_basis, int n_cols, void* basis_split,
                                            void* stream) {
    SSAR_REQUIRE(basis && basis_split, "mix_gemm_tc_split_basis: null pointer");
    SSAR_REQUIRE(n_basis > 0 && n_basis <= TC_MAX_KB * TC_BK && n_cols > 0, "mix_gemm_tc_split_basis: n_basis must be in [1, 192]");
    uint8_t* blk = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(basis_split) + 1023) & ~(uintptr_t)1023);
    return tc_split_basis(basis, ldb, n_basis, n_cols, blk, (cudaStream_t)stream);
}

extern "C" long long ssar_mix_gemm_tc_scratch_bytes(int n_rows, int n_groups, int n_basis, int n_cols) {
    const long long KP = tc_kp(n_basis), Tpad = tc_tpad(n_rows);
    return 1024 + tc_scale_bytes(n_groups, Tpad) + 2 * (long long)n_groups * Tpad * KP * 2 +
           ssar_mix_gemm_tc_basis_bytes(n_basis, n_cols) + 1024;
}

extern "C" int ssar_mix_gemm_tc(const float* weights, long long ldw, long long group_stride, int group_cols,
                                const float* basis, long long ldb, float* out, long long ldo, int n_rows, int n_basis,
                                int n_cols, const void* basis_split, void* scratch, void* stream) {
    if (n_rows == 0 || n_cols == 0) return SSAR_OK;
    SSAR_REQUIRE(weights && (basis || basis_split) && out && scratch, "mix_gemm_tc: null pointer");
    SSAR_REQUIRE(group_cols > 0 && group_cols % TC_BN == 0 && n_cols % group_cols == 0,
                 "mix_gemm_tc: group_cols must be a multiple of 128 that divides n_cols");
    SSAR_REQUIRE(n_basis > 0 && n_basis <= TC_MAX_KB * TC_BK, "mix_gemm_tc: n_basis must be in [1, 192] (use ssar_mix_gemm)");
    SSAR_REQUIRE(ldo % 4 == 0 && ((uintptr_t)out % 16 == 0), "mix_gemm_tc: out rows must be 16-byte aligned");
    SSAR_REQUIRE(n_cols % 32 == 0, "mix_gemm_tc: n_cols must be a multiple of 32");
    cudaStream_t st = (cudaStream_t)stream;
    const int G = n_cols / group_cols;
    const int KP = (int)tc_kp(n_basis), Tpad = (int)tc_tpad(n_rows), KB = KP / TC_BK;
    uint8_t* sb = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(scratch) + 1023) & ~(uintptr_t)1023);
    uint32_t* amax_w = reinterpret_cast<uint32_t*>(sb);  // one amax word per (group, row)
    __half* Ah = reinterpret_cast<__half*>(sb + tc_scale_bytes(G, Tpad));
    __half* Al = Ah + (long long)G * Tpad * KP;
    const uint8_t* bblk;
    if (basis_split) {  // the basis was split once for many calls (ssar_mix_gemm_tc_split_basis)
        bblk = reinterpret_cast<const uint8_t*>((reinterpret_cast<uintptr_t>(basis_split) + 1023) & ~(uintptr_t)1023);
    } else {
        uint8_t* mine = reinterpret_cast<uint8_t*>(Al + (long long)G * Tpad * KP);
        mine = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(mine) + 1023) & ~(uintptr_t)1023);
        const int rc = tc_split_basis(basis, ldb, n_basis, n_cols, mine, st);
        if (rc != SSAR_OK) return rc;
        bblk = mine;
    }
    const uint32_t* amax_b = reinterpret_cast<const uint32_t*>(bblk);
    const __half* Bh = reinterpret_cast<const __half*>(bblk + 1024);
    const __half* Bl = Bh + (long long)n_cols * KP;

    {
        int grid = ssar_div_up((long long)G * Tpad, 8);
        if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
        split_weights_kernel<<<grid, 256, 0, st>>>(weights, ldw, group_stride, n_rows, Tpad, G, n_basis, KP, amax_w, Ah, Al);
    }
    CUtensorMap mah, mal, mbh, mbl, mout;
    if (!make_panel_map(&mah, Ah, (long long)G * Tpad, KP) || !make_panel_map(&mal, Al, (long long)G * Tpad, KP) ||
        !make_panel_map(&mbh, Bh, n_cols, KP) || !make_panel_map(&mbl, Bl, n_cols, KP) ||
        !make_out_map(&mout, out, n_rows, n_cols, ldo)) {
        ssar_set_error("mix_gemm_tc: cuTensorMapEncodeTiled unavailable or failed");
        return SSAR_ERR_CUDA;
    }
    const size_t smem = (size_t)(TC_MAX_KB + TC_STAGES) * 2 * TC_TILE_BYTES + 4 * TC_CBUFS * TC_CSTAGE_BYTES + sizeof(TcShared) + 1024;
    cudaError_t e = cudaFuncSetAttribute(mix_gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        ssar_set_error("mix_gemm_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    const int base_items = (Tpad / TC_BM) * G, tiles_per_group = group_cols / TC_BN;
    int n_splits = SSAR_NUM_SMS / base_items;  // fill the SMs when there are few (row block, group) pairs
    if (n_splits < 1) n_splits = 1;
    if (n_splits > tiles_per_group) n_splits = tiles_per_group;
    int grid = base_items * n_splits;
    if (grid > SSAR_NUM_SMS) grid = SSAR_NUM_SMS;
    mix_gemm_tc_kernel<<<grid, TC_THREADS, smem, st>>>(mah, mal, mbh, mbl, mout, Tpad, G, group_cols, KB, n_splits, amax_w,
                                                       amax_b);
    return ssar_launch_status("mix_gemm_tc");
}
