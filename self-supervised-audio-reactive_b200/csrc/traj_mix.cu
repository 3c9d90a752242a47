// Latent-trajectory kernels: feature-weighted sums of latent bases written straight into [T, num_ws, 512].
//
// Every latent sequence the reference builds is linear in the latent palette with time-varying scalar weights:
//   einsum("TN,NWL->TWL") (ssar/random/latent.py:48, features/generate.py:167,216, models/latent_n_noise2.py:159),
//   the 2-latent lerp (latent.py:45), segmentation gathers (latent.py:39-41), natural-cubic-spline loops
//   (latent.py:7-13: sum_k B_k(t) y_k) and the Gaussian smoothing / average / modulate merges that follow them.
// So the host keeps only the small weight matrices  A_g [T, K]  (one per layer group g) and this kernel does
//   C[t, n] = alpha[t] * C[t, n] + beta[t] * sum_k A_{g(n)}[t, k] * B[k, n],     n = layer * 512 + l
// once, with B = the palette [K, num_ws * 512] and g(n) = n / group_cols.
//
// Tiling: CTA = 64 rows x 128 columns, K in slabs of 16 staged in shared memory, 256 threads each owning an 8 x 4
// register tile (float4 along n: coalesced 512-byte rows for B loads and C stores).  The op is bound by the C write
// (36.9 kB per frame); K is small (<= a few hundred).
#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 128, BK = 16;
constexpr int AS_LD = BM + 4;

__global__ void __launch_bounds__(256) mix_gemm_kernel(const float* __restrict__ A, long long lda, long long group_stride,
                                                       int group_cols, const float* __restrict__ B, long long ldb,
                                                       float* __restrict__ C, long long ldc, int M, int K, int col0,
                                                       int col1, const float* __restrict__ row_alpha,
                                                       const float* __restrict__ row_beta) {
    __shared__ __align__(16) float As[BK * AS_LD];
    __shared__ __align__(16) float Bs[BK * BN];
    const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
    const int m0 = blockIdx.y * BM;
    const int n0 = col0 + blockIdx.x * BN;
    const float* Ag = A + (long long)(n0 / group_cols) * group_stride;

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += BK) {
        // A slab: 64 rows x 16 k (row-major in global) -> As[k][row]
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * 256;
            const int r = idx >> 4, k = idx & 15;
            const int m = m0 + r;
            As[k * AS_LD + r] = (m < M && k0 + k < K) ? __ldg(Ag + (long long)m * lda + k0 + k) : 0.f;
        }
        // B slab: 16 k x 128 columns
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int idx = tid + i * 256;
            const int k = idx >> 5, c4 = idx & 31;
            const int n = n0 + c4 * 4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k0 + k < K && n + 3 < col1) v = __ldg(reinterpret_cast<const float4*>(B + (long long)(k0 + k) * ldb + n));
            *reinterpret_cast<float4*>(&Bs[k * BN + c4 * 4]) = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 b = *reinterpret_cast<const float4*>(&Bs[k * BN + tx * 4]);
            const float4 a0 = *reinterpret_cast<const float4*>(&As[k * AS_LD + ty * 8]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[k * AS_LD + ty * 8 + 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                acc[i][0] = fmaf(a[i], b.x, acc[i][0]);
                acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
                acc[i][2] = fmaf(a[i], b.z, acc[i][2]);
                acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
            }
        }
        __syncthreads();
    }

    const int n = n0 + tx * 4;
    if (n + 3 >= col1) return;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + ty * 8 + i;
        if (m >= M) continue;
        float4* dst = reinterpret_cast<float4*>(C + (long long)m * ldc + n);
        float4 r = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
        if (row_beta) {
            const float bt = __ldg(row_beta + m);
            r.x *= bt; r.y *= bt; r.z *= bt; r.w *= bt;
        }
        if (row_alpha) {
            const float al = __ldg(row_alpha + m);
            const float4 old = *dst;
            r.x = fmaf(al, old.x, r.x);
            r.y = fmaf(al, old.y, r.y);
            r.z = fmaf(al, old.z, r.z);
            r.w = fmaf(al, old.w, r.w);
        }
        *dst = r;
    }
}

// weights[t, g, k] = alpha[t] * weights[t, g, k] + beta[t] * S[t, k]   for g in [g0, g1)
// (the average / modulate / overwrite merges of latent.py:68-78 applied to the weight matrices)
__global__ void weights_merge_kernel(float* __restrict__ Wt, int T, int G, int K, int g0, int g1,
                                     const float* __restrict__ S, const float* __restrict__ alpha,
                                     const float* __restrict__ beta, float alpha_s, float beta_s) {
    const long long n = (long long)T * (g1 - g0) * K;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % K);
        const long long r = i / K;
        const int g = g0 + (int)(r % (g1 - g0));
        const int t = (int)(r / (g1 - g0));
        const float al = alpha ? alpha[t] : alpha_s, be = beta ? beta[t] : beta_s;
        float* w = Wt + ((long long)t * G + g) * K + k;
        *w = fmaf(al, *w, be * S[(long long)t * K + k]);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// One latent sub-patch (ssar/random/latent.py:35-78) applied to the weight matrices in ONE launch:
//   raw[t, k]   = src_scale * src[t, k]                       (src_mode 0: einsum weights / spline weights / smoothed one-hots)
//               = (f, 1 - f), f = src_scale * src[t, 0]       (src_mode 1: the 2-latent lerp of latent.py:44-45)
//   S[t, k]     = sum_i taps[i] * raw[(t + i - radius) mod T, k]    (the circular sigma = 1 smoothing of latent.py:54)
//   beta[t]     = mod_scale * mod[t] (alpha = 1 - beta) when mod != NULL, else the scalars
//   W[t, g, sel[k]] = alpha * W[t, g, sel[k]] + beta * S[t, k],   W[t, g, p] = alpha * W[t, g, p] elsewhere,  g in [g0, g1)
// Accumulation order of the taps equals gauss1d_kernel's (ascending tap index, fmaf), products that the reference
// rounds separately (feature * weight, 1 - f, 1 - m) are rounded separately here (__fmul_rn / __fsub_rn).
// CTA = SP_TT frames; the inverse selection map (palette column -> k) and the smoothed rows live in shared memory.
// ---------------------------------------------------------------------------------------------------------------
constexpr int SP_TT = 8;

__device__ __forceinline__ float subpatch_raw(const float* __restrict__ src, long long ld, float scale, int mode, int t, int k) {
    if (mode == 0) return __fmul_rn(scale, __ldg(src + (long long)t * ld + k));
    const float f = __fmul_rn(scale, __ldg(src + (long long)t * ld));
    return k == 0 ? f : __fsub_rn(1.f, f);
}

__global__ void __launch_bounds__(256) subpatch_merge_kernel(float* __restrict__ Wt, int T, int G, int P, int g0, int g1,
                                                             const long long* __restrict__ sel, int K,
                                                             const float* __restrict__ src, long long src_ld, float src_scale,
                                                             int src_mode, const float* __restrict__ taps, int radius,
                                                             const float* __restrict__ mod, long long mod_ld, float mod_scale,
                                                             float alpha_s, float beta_s) {
    extern __shared__ float sp_smem[];
    float* S = sp_smem;                                   // [SP_TT][K]
    float* ab = S + SP_TT * K;                            // [SP_TT][2]
    int* inv = reinterpret_cast<int*>(ab + SP_TT * 2);    // [P]
    const int t0 = blockIdx.x * SP_TT;
    for (int p = threadIdx.x; p < P; p += blockDim.x) inv[p] = -1;
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        const long long p = sel[k];
        if (p >= 0 && p < P) inv[(int)p] = k;
    }
    for (int i = threadIdx.x; i < SP_TT * K; i += blockDim.x) {
        const int tt = i / K, k = i - tt * K;
        const int t = t0 + tt;
        float acc = 0.f;
        if (t < T) {
            if (radius == 0) {
                acc = subpatch_raw(src, src_ld, src_scale, src_mode, t, k);
            } else {
                int ti = (t - radius) % T;
                if (ti < 0) ti += T;
                for (int j = 0; j <= 2 * radius; ++j) {
                    acc = fmaf(__ldg(taps + j), subpatch_raw(src, src_ld, src_scale, src_mode, ti, k), acc);
                    if (++ti == T) ti = 0;
                }
            }
        }
        S[i] = acc;
    }
    if (threadIdx.x < SP_TT) {
        const int t = t0 + threadIdx.x;
        float al = alpha_s, be = beta_s;
        if (mod && t < T) {
            be = __fmul_rn(mod_scale, __ldg(mod + (long long)t * mod_ld));
            al = __fsub_rn(1.f, be);
        }
        ab[2 * threadIdx.x] = al;
        ab[2 * threadIdx.x + 1] = be;
    }
    __syncthreads();
    const int Gr = g1 - g0;
    const bool overwrite = (mod == nullptr) && alpha_s == 0.f;
    for (int i = threadIdx.x; i < SP_TT * Gr * P; i += blockDim.x) {
        const int p = i % P;
        const int r = i / P;
        const int g = g0 + r % Gr;
        const int tt = r / Gr;
        const int t = t0 + tt;
        if (t >= T) break;
        const int k = inv[p];
        const float s = k >= 0 ? S[tt * K + k] : 0.f;
        float* w = Wt + ((long long)t * G + g) * P + p;
        const float al = ab[2 * tt], be = ab[2 * tt + 1];
        *w = overwrite ? be * s : fmaf(al, *w, be * s);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// The whole sub-patch stack of one clip in one launch.  W[t] after sub-patch j depends only on W[t] after j - 1 and on
// the (time-smoothed) source of sub-patch j around t, so a CTA that owns SP_TT frames can walk the stack with its
// weight rows resident in shared memory: [T, G, P] is written once instead of read + written per sub-patch, and the
// ~13 dependent launches per clip become one.  Arithmetic per element is subpatch_merge_kernel's.
// ---------------------------------------------------------------------------------------------------------------
constexpr int STACK_MAX = 24;
constexpr int STACK_KMAX = 32;
struct StackArgs {
    ssar_subpatch_desc d[STACK_MAX];
    int n;
};

// v2: the sources of ALL sub-patches are independent of W, so they are evaluated side by side first (phase A: every
// thread has its global loads of ~n sub-patches in flight at once, one barrier), and a weight element then walks the
// stack in a register (phase B: element (t, g, p) after sub-patch j depends only on itself and S_j[t][inv_j[p]]), so
// the per-sub-patch barriers and their serialised load latencies (v1: 5 barriers x n, barrier stall 5-6 per issue) are
// gone and W no longer needs shared memory.  Per-element arithmetic is unchanged.
__global__ void __launch_bounds__(256) latent_stack_kernel(float* __restrict__ Wt, int T, int G, int P, const StackArgs sa,
                                                           const float* __restrict__ taps, int load_initial, int n_label) {
    extern __shared__ float st_smem[];
    const int n = sa.n;
    const int GP = G * P;
    float* S = st_smem;                                          // [n][SP_TT][STACK_KMAX]
    float* ab = S + n * SP_TT * STACK_KMAX;                      // [n][SP_TT][2]
    float* OH = ab + n * SP_TT * 2;                              // [n_label][SP_TT + 16][STACK_KMAX] (radius <= 8)
    signed char* inv = reinterpret_cast<signed char*>(OH + n_label * (SP_TT + 16) * STACK_KMAX);  // [n][P]
    __shared__ int label_slot[STACK_MAX], item_off[STACK_MAX + 1];
    const int t0 = blockIdx.x * SP_TT;
    const int rows = min(SP_TT, T - t0);

    for (int i = threadIdx.x; i < n * P; i += blockDim.x) inv[i] = -1;
    if (threadIdx.x == 0) {
        int c = 0, o = 0;
        for (int j = 0; j < n; ++j) {
            label_slot[j] = sa.d[j].src_mode == 2 ? c++ : -1;
            item_off[j] = o;
            o += SP_TT * sa.d[j].n_select;
        }
        item_off[n] = o;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n * STACK_KMAX; i += blockDim.x) {
        const int j = i / STACK_KMAX, k = i - j * STACK_KMAX;
        if (k < sa.d[j].n_select) {
            const long long p = sa.d[j].selection[k];
            if (p >= 0 && p < P) inv[j * P + (int)p] = (signed char)k;
        }
    }
    // smoothed one-hot rows of the label sub-patches for t0 - radius .. t0 + SP_TT + radius (onehot_gauss_kernel's order)
    for (int j = 0; j < n; ++j) {
        const ssar_subpatch_desc& d = sa.d[j];
        if (d.src_mode != 2) continue;
        const long long* __restrict__ labels = reinterpret_cast<const long long*>(d.src);
        const int K = d.n_select, nr = SP_TT + 2 * d.radius, pr = d.pre_radius;
        float* oh = OH + label_slot[j] * (SP_TT + 16) * STACK_KMAX;
        for (int i = threadIdx.x; i < nr * K; i += blockDim.x) {
            const int rr = i / K, k = i - rr * K;
            int t = (t0 - d.radius + rr) % T;
            if (t < 0) t += T;
            int ti = (t - pr) % T;
            if (ti < 0) ti += T;
            float acc = 0.f;
            for (int q = 0; q <= 2 * pr; ++q) {
                acc = fmaf(__ldg(d.pre_taps + q), (__ldg(labels + ti) == k) ? 1.f : 0.f, acc);
                if (++ti == T) ti = 0;
            }
            oh[rr * STACK_KMAX + k] = acc;
        }
    }
    if (n_label > 0) __syncthreads();
    // phase A: time-smoothed sources S_j[tt][k] of every sub-patch; the (j, tt, k < K_j) items are packed so that all
    // 256 threads hold loads of different sub-patches at once, and an item's taps are fetched before its fmaf chain
    const int n_items = item_off[n];
    for (int i = threadIdx.x; i < n_items; i += blockDim.x) {
        int j = 0;
        while (i >= item_off[j + 1]) ++j;
        const ssar_subpatch_desc& d = sa.d[j];
        const int K = d.n_select, radius = d.radius;
        const int li = i - item_off[j];
        const int tt = li / K, k = li - tt * K;
        const int t = t0 + tt;
        float acc = 0.f;
        if (t < T) {
            const float* __restrict__ srcf = reinterpret_cast<const float*>(d.src);
            if (d.src_mode == 2) {
                const float* oh = OH + label_slot[j] * (SP_TT + 16) * STACK_KMAX;
                if (radius == 0) {
                    acc = __fmul_rn(d.src_scale, oh[tt * STACK_KMAX + k]);
                } else {
                    for (int q = 0; q <= 2 * radius; ++q)
                        acc = fmaf(__ldg(taps + q), __fmul_rn(d.src_scale, oh[(tt + q) * STACK_KMAX + k]), acc);
                }
            } else if (radius == 0) {
                acc = subpatch_raw(srcf, d.src_ld, d.src_scale, d.src_mode, t, k);
            } else {
                int ti = (t - radius) % T;
                if (ti < 0) ti += T;
                float v[17];
#pragma unroll
                for (int q = 0; q < 17; ++q) {
                    v[q] = q <= 2 * radius ? subpatch_raw(srcf, d.src_ld, d.src_scale, d.src_mode, ti, k) : 0.f;
                    if (++ti == T) ti = 0;
                }
#pragma unroll
                for (int q = 0; q < 17; ++q)
                    if (q <= 2 * radius) acc = fmaf(__ldg(taps + q), v[q], acc);
            }
        }
        S[(j * SP_TT + tt) * STACK_KMAX + k] = acc;
    }
    for (int i = threadIdx.x; i < n * SP_TT; i += blockDim.x) {
        const int j = i / SP_TT, t = t0 + (i - j * SP_TT);
        const ssar_subpatch_desc& d = sa.d[j];
        float al = d.alpha, be = d.beta;
        if (d.mod && t < T) {
            be = __fmul_rn(d.mod_scale, __ldg(d.mod + (long long)t * d.mod_ld));
            al = __fsub_rn(1.f, be);
        }
        ab[2 * i] = al;
        ab[2 * i + 1] = be;
    }
    __syncthreads();
    // phase B: a thread owns weight columns (g, p) for the CTA's SP_TT frames and walks them through the stack in
    // registers; per (sub-patch, column) one inv lookup, per frame one S and one (alpha, beta) read
    const float2* ab2 = reinterpret_cast<const float2*>(ab);
    for (int r = threadIdx.x; r < GP; r += blockDim.x) {
        const int g = r / P, p = r - g * P;
        float* wp = Wt + (long long)t0 * GP + r;
        float w[SP_TT];
#pragma unroll
        for (int tt = 0; tt < SP_TT; ++tt) w[tt] = (load_initial && tt < rows) ? wp[(long long)tt * GP] : 0.f;
        for (int j = 0; j < n; ++j) {
            const ssar_subpatch_desc& d = sa.d[j];
            if (g < d.g_begin || g >= d.g_end) continue;
            const int k = inv[j * P + p];
            const bool overwrite = (d.mod == nullptr) && d.alpha == 0.f;
            const float* Sj = S + j * SP_TT * STACK_KMAX + max(k, 0);
#pragma unroll
            for (int tt = 0; tt < SP_TT; ++tt) {
                const float sv = k >= 0 ? Sj[tt * STACK_KMAX] : 0.f;
                const float2 q = ab2[j * SP_TT + tt];
                w[tt] = overwrite ? q.y * sv : fmaf(q.x, w[tt], q.y * sv);
            }
        }
#pragma unroll
        for (int tt = 0; tt < SP_TT; ++tt)
            if (tt < rows) wp[(long long)tt * GP] = w[tt];
    }
}

// out[t, k] = sum_i taps[i] * (labels[(t + i - radius) mod T] == k): the one-hot gather of latent.py:39-40 followed by
// the circular Gaussian of :41, without materialising the one-hot matrix (same tap order as gauss1d_kernel).
__global__ void onehot_gauss_kernel(const long long* __restrict__ labels, int T, int K, const float* __restrict__ taps,
                                    int radius, float* __restrict__ out) {
    const long long n = (long long)T * K;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int t = (int)(i / K), k = (int)(i - (long long)t * K);
        int ti = (t - radius) % T;
        if (ti < 0) ti += T;
        float acc = 0.f;
        for (int j = 0; j <= 2 * radius; ++j) {
            acc = fmaf(__ldg(taps + j), (__ldg(labels + ti) == k) ? 1.f : 0.f, acc);
            if (++ti == T) ti = 0;
        }
        out[i] = acc;
    }
}

// out[t, j] = mu[t] + sig[t] * x[t, j]   (models/latent_n_noise2.py:169-174) — mu / sig may be NULL (0 / 1)
__global__ void rows_affine_kernel(const float* __restrict__ x, long long rows, long long inner, const float* __restrict__ mu,
                                   const float* __restrict__ sig, float* __restrict__ out) {
    const long long n = rows * inner;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long t = i / inner;
        const float m = mu ? mu[t] : 0.f, s = sig ? sig[t] : 1.f;
        out[i] = m + s * x[i];
    }
}

}  // namespace

extern "C" int ssar_mix_gemm(const float* weights, long long ldw, long long group_stride, int group_cols,
                             const float* basis, long long ldb, float* out, long long ldo, int n_rows, int n_basis,
                             int col_begin, int col_end, const float* row_alpha, const float* row_beta, void* stream) {
    if (n_rows == 0 || col_end <= col_begin) return SSAR_OK;
    SSAR_REQUIRE(weights && basis && out, "mix_gemm: null pointer");
    SSAR_REQUIRE(n_basis >= 0 && group_cols > 0, "mix_gemm: bad sizes");
    SSAR_REQUIRE(col_begin % BN == 0 && (col_end - col_begin) % 4 == 0 && group_cols % BN == 0,
                 "mix_gemm: col_begin and group_cols must be multiples of 128 and the column count a multiple of 4");
    SSAR_REQUIRE((ldb % 4 == 0) && (ldo % 4 == 0) && ((uintptr_t)basis % 16 == 0) && ((uintptr_t)out % 16 == 0),
                 "mix_gemm: basis / out rows must be 16-byte aligned");
    dim3 grid(ssar_div_up(col_end - col_begin, BN), ssar_div_up(n_rows, BM));
    SSAR_REQUIRE(grid.y <= 65535, "mix_gemm: too many rows per launch");
    mix_gemm_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(weights, ldw, group_stride, group_cols, basis, ldb, out, ldo,
                                                           n_rows, n_basis, col_begin, col_end, row_alpha, row_beta);
    return ssar_launch_status("mix_gemm");
}

extern "C" int ssar_weights_merge(float* weights, int n_rows, int n_groups, int n_basis, int g_begin, int g_end,
                                  const float* seq, const float* alpha, const float* beta, float alpha_scalar,
                                  float beta_scalar, void* stream) {
    if (n_rows == 0 || n_basis == 0 || g_end <= g_begin) return SSAR_OK;
    SSAR_REQUIRE(weights && seq, "weights_merge: null pointer");
    SSAR_REQUIRE(0 <= g_begin && g_end <= n_groups, "weights_merge: bad group range");
    const long long n = (long long)n_rows * (g_end - g_begin) * n_basis;
    int grid = ssar_div_up(n, 256);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    weights_merge_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(weights, n_rows, n_groups, n_basis, g_begin, g_end, seq,
                                                                alpha, beta, alpha_scalar, beta_scalar);
    return ssar_launch_status("weights_merge");
}

extern "C" int ssar_subpatch_merge(float* weights, int n_rows, int n_groups, int n_basis, int g_begin, int g_end,
                                   const long long* selection, int n_select, const float* src, long long src_ld,
                                   float src_scale, int src_mode, const float* taps, int radius, const float* mod,
                                   long long mod_ld, float mod_scale, float alpha_scalar, float beta_scalar, void* stream) {
    if (n_rows == 0 || n_basis == 0 || g_end <= g_begin) return SSAR_OK;
    SSAR_REQUIRE(weights && selection && src, "subpatch_merge: null pointer");
    SSAR_REQUIRE(0 <= g_begin && g_end <= n_groups, "subpatch_merge: bad group range");
    SSAR_REQUIRE(n_select > 0 && (src_mode == 0 || (src_mode == 1 && n_select == 2)),
                 "subpatch_merge: src_mode 1 (lerp) needs exactly 2 selected latents");
    SSAR_REQUIRE(radius >= 0 && (radius == 0 || taps), "subpatch_merge: taps required when radius > 0");
    const size_t smem = ((size_t)SP_TT * n_select + SP_TT * 2 + n_basis) * sizeof(float);
    SSAR_REQUIRE(smem <= 200 * 1024, "subpatch_merge: palette / selection too large for one CTA's shared memory");
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(subpatch_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    subpatch_merge_kernel<<<ssar_div_up(n_rows, SP_TT), 256, smem, (cudaStream_t)stream>>>(
        weights, n_rows, n_groups, n_basis, g_begin, g_end, selection, n_select, src, src_ld, src_scale, src_mode, taps,
        radius, mod, mod_ld, mod_scale, alpha_scalar, beta_scalar);
    return ssar_launch_status("subpatch_merge");
}

extern "C" int ssar_latent_stack(float* weights, int n_rows, int n_groups, int n_basis, const ssar_subpatch_desc* descs,
                                 int n_descs, const float* taps, int load_initial, void* stream) {
    if (n_rows == 0 || n_basis == 0 || n_descs == 0) return SSAR_OK;
    SSAR_REQUIRE(weights && descs, "latent_stack: null pointer");
    SSAR_REQUIRE(n_descs > 0 && n_descs <= STACK_MAX, "latent_stack: at most 24 sub-patches per launch");
    StackArgs sa;
    sa.n = n_descs;
    for (int j = 0; j < n_descs; ++j) {
        const ssar_subpatch_desc& d = descs[j];
        SSAR_REQUIRE(d.selection && d.src, "latent_stack: null selection / src in sub-patch %d", j);
        SSAR_REQUIRE(0 <= d.g_begin && d.g_begin < d.g_end && d.g_end <= n_groups, "latent_stack: bad group range");
        SSAR_REQUIRE(d.n_select > 0 && d.n_select <= STACK_KMAX, "latent_stack: 1..32 selected latents per sub-patch");
        SSAR_REQUIRE(d.src_mode == 0 || d.src_mode == 2 || (d.src_mode == 1 && d.n_select == 2),
                     "latent_stack: src_mode 1 (lerp) needs exactly 2 selected latents");
        SSAR_REQUIRE(d.radius >= 0 && d.radius <= 8 && (d.radius == 0 || taps), "latent_stack: radius must be in [0, 8]");
        SSAR_REQUIRE(d.src_mode != 2 || (d.pre_taps && d.pre_radius >= 0), "latent_stack: labels need pre_taps");
        sa.d[j] = d;
    }
    int n_label = 0;
    for (int j = 0; j < n_descs; ++j) n_label += descs[j].src_mode == 2 ? 1 : 0;
    const size_t smem = ((size_t)n_descs * SP_TT * STACK_KMAX + (size_t)n_descs * SP_TT * 2 +
                         (size_t)n_label * (SP_TT + 16) * STACK_KMAX) * sizeof(float) + (size_t)n_descs * n_basis;
    SSAR_REQUIRE(smem <= 200 * 1024 && n_basis <= 32767, "latent_stack: too many basis latents for one CTA's shared memory");
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(latent_stack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    latent_stack_kernel<<<ssar_div_up(n_rows, SP_TT), 256, smem, (cudaStream_t)stream>>>(weights, n_rows, n_groups, n_basis,
                                                                                       sa, taps, load_initial, n_label);
    return ssar_launch_status("latent_stack");
}

extern "C" int ssar_onehot_gauss(const long long* labels, int n_rows, int n_classes, const float* taps, int radius,
                                 float* out, void* stream) {
    if (n_rows == 0 || n_classes == 0) return SSAR_OK;
    SSAR_REQUIRE(labels && out && (radius == 0 || taps), "onehot_gauss: null pointer");
    SSAR_REQUIRE(radius >= 0, "onehot_gauss: bad radius");
    int grid = ssar_div_up((long long)n_rows * n_classes, 256);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    onehot_gauss_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(labels, n_rows, n_classes, taps, radius, out);
    return ssar_launch_status("onehot_gauss");
}

extern "C" int ssar_rows_affine(const float* x, long long n_rows, long long inner, const float* mu, const float* sig,
                                float* out, void* stream) {
    if (n_rows == 0 || inner == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out, "rows_affine: null pointer");
    int grid = ssar_div_up(n_rows * inner, 256 * 4);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    if (grid < 1) grid = 1;
    rows_affine_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, n_rows, inner, mu, sig, out);
    return ssar_launch_status("rows_affine");
}
