// Second-order IIR (biquad) over whole clips as a three-phase blocked scan.
// Replaces torchaudio.functional.lowpass_biquad / highpass_biquad -> biquad -> lfilter(clamp=True) behind
// features/processing.py:141-150 (low_pass / mid_pass / high_pass):
//     v[n] = b0 x[n] + b1 x[n-1] + b2 x[n-2]          (coefficients already divided by a0)
//     y[n] = v[n] - (a2 y[n-2] + a1 y[n-1]),  y[-1] = y[-2] = 0,  output clamped to [-1, 1] at the very end.
//
// Phase A: every chunk of L samples runs the recurrence from a ZERO state and keeps its final state z_c; the CTA (128
//          chunks) also scans its z_c in fp64 and emits the zero-state end state of its whole span.
// Phase B: the true end-of-span states follow the affine recurrence S_c = A^(128 L) S_{c-1} + Z_c (A = companion
//          matrix); one CTA per clip scans the ~720 spans in fp64 (one element per thread, coalesced).
// Phase C: every CTA rebuilds the true state entering each of its chunks (fp64 block scan of z_c + the state entering
//          the span), then every chunk re-runs the SAME fp32 recurrence from that state and writes clamp(y), so within
//          a chunk the operation order is the sequential reference's.
// Chunks are staged through shared memory so global reads and writes stay coalesced.
#include "common.cuh"

namespace {

constexpr int L = 64;        // samples per chunk
constexpr int CPB = 128;     // chunks (= threads) per CTA
constexpr int LDT = L + 1;   // padded chunk stride in shared memory

struct Coef {
    float b0, b1, b2, a1, a2;
};

__device__ __forceinline__ float step(const Coef& c, float x0, float x1, float x2, float y1, float y2) {
    // explicit roundings: the contraction of these sums must not depend on what else the compiler sees in the loop (the
    // single- and the dual-filter instantiation have to give the same bits)
    const float v = __fmaf_rn(c.b0, x0, __fmaf_rn(c.b1, x1, __fmul_rn(c.b2, x2)));
    return __fsub_rn(v, __fmaf_rn(c.a1, y1, __fmul_rn(c.a2, y2)));
}

struct M2 {
    double a, b, c, d;  // [[a, b], [c, d]]
};
__device__ __forceinline__ M2 mm(const M2& x, const M2& y) {
    return {x.a * y.a + x.b * y.c, x.a * y.b + x.b * y.d, x.c * y.a + x.d * y.c, x.c * y.b + x.d * y.d};
}
struct V2 {
    double x, y;
};
__device__ __forceinline__ V2 mv(const M2& m, const V2& v) { return {m.a * v.x + m.b * v.y, m.c * v.x + m.d * v.y}; }

// powers AL^(2^s), s = 0..7, of AL = A^L (A = companion matrix on (y[n-1], y[n-2])), in fp64
struct ALPowers {
    M2 p[8];
};
__device__ __forceinline__ ALPowers al_powers(double a1, double a2) {
    ALPowers r;
    M2 AL = {-a1, -a2, 1.0, 0.0};
#pragma unroll
    for (int i = 0; i < 6; ++i) AL = mm(AL, AL);  // L = 64 = 2^6
    r.p[0] = AL;
#pragma unroll
    for (int i = 1; i < 8; ++i) r.p[i] = mm(r.p[i - 1], r.p[i - 1]);
    return r;
}

// inclusive scan over the CTA's CPB chunks of q_j = sum_{i <= j} AL^(j - i) z_i (Hillis-Steele, fp64, shared memory)
__device__ __forceinline__ V2 block_scan_states(V2 q, const ALPowers& pw, V2* sq) {
    const int tid = threadIdx.x;
    sq[tid] = q;
    __syncthreads();
#pragma unroll
    for (int s = 0; s < 7; ++s) {  // CPB = 128 = 2^7
        const int d = 1 << s;
        V2 add = {0.0, 0.0};
        if (tid >= d) add = mv(pw.p[s], sq[tid - d]);
        __syncthreads();
        q.x += add.x;
        q.y += add.y;
        sq[tid] = q;
        __syncthreads();
    }
    return q;
}

// Per-filter arguments of the chunk kernel: NF = 1, or 2 filters applied to the SAME input in one pass (the low-pass and the
// high-pass of features/processing.py:141-150 read their common input once; two independent recurrences per thread also
// double the instruction-level parallelism of the 64-step chain).
struct FilterArgs {
    Coef c;
    const float2* state_in;   // [clip, chunk] zero-state end states (phase C)
    float2* state_out;        // [clip, chunk] (phase A)
    float* y;                 // output signal or NULL
    V2* cta_z;                // per-CTA aggregates (phase A)
    const V2* cta_s;          // true state at the end of every CTA span (phase C)
    double a1d, a2d;
    float* block_ss;          // per-1024-sample sums of squares of the output, or NULL
};

template <bool WRITE, int NF>
__global__ void __launch_bounds__(CPB) biquad_chunks_kernel(const float* __restrict__ x, long long n, long long stride_x,
                                                            const FilterArgs fa0, const FilterArgs fa1, long long stride_y,
                                                            long long n_chunks, int clamp, long long n_ss_blocks) {
    const FilterArgs* fas[2] = {&fa0, &fa1};
    __shared__ float tile[CPB * LDT];
    extern __shared__ float otile[];  // (WRITE && NF > 1): the second filter's output tile, CPB * LDT floats
    __shared__ V2 sq[CPB];
    // powers of the chunk transition matrices: computed once per CTA (while the other threads load the tile) instead of by
    // every thread — 64 registers of fp64 state and ~100 fp64 instructions per thread less, same values
    __shared__ ALPowers pw[NF];
    if (threadIdx.x == 0) pw[0] = al_powers(fa0.a1d, fa0.a2d);
    if (NF > 1 && threadIdx.x == 32) pw[NF - 1] = al_powers(fa1.a1d, fa1.a2d);
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * stride_x;
    const long long tile0 = (long long)blockIdx.x * CPB * L;
    const bool vec_ok = tile0 + (long long)CPB * L <= n && ((reinterpret_cast<uintptr_t>(xc + tile0) & 15) == 0);
    if (vec_ok) {  // whole tile inside the clip and 16-byte aligned: 128-bit loads, 8 in flight per thread
        const float4* __restrict__ x4 = reinterpret_cast<const float4*>(xc + tile0);
#pragma unroll 8
        for (int q = threadIdx.x; q < CPB * L / 4; q += CPB) {
            const float4 v = __ldg(x4 + q);
            float* d = tile + ((4 * q) / L) * LDT + ((4 * q) % L);
            d[0] = v.x;
            d[1] = v.y;
            d[2] = v.z;
            d[3] = v.w;
        }
    } else {
#pragma unroll 4
        for (int i = threadIdx.x; i < CPB * L; i += CPB) {
            const long long g = tile0 + i;
            tile[(i / L) * LDT + (i % L)] = g < n ? __ldg(xc + g) : 0.f;
        }
    }
    __syncthreads();
    const long long chunk = (long long)blockIdx.x * CPB + threadIdx.x;
    float y1[NF], y2[NF], ssq[NF];
#pragma unroll
    for (int f = 0; f < NF; ++f) y1[f] = y2[f] = ssq[f] = 0.f;
    if (WRITE) {
        // true state entering every chunk of this CTA: S_j = AL^(j+1) S_in + sum_{i<=j} AL^(j-i) z_i (fp64), chunk j starts
        // from S_{j-1}; S_in = the state entering the CTA from the carry kernel
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            V2 z = {0.0, 0.0};
            if (chunk < n_chunks) {
                const float2 zz = fas[f]->state_in[(long long)clip * n_chunks + chunk];
                z = {(double)zz.x, (double)zz.y};
            }
            V2 S = block_scan_states(z, pw[f], sq);
            V2 sin = {0.0, 0.0};
            if (blockIdx.x > 0) sin = fas[f]->cta_s[(long long)clip * gridDim.x + blockIdx.x - 1];
            V2 prop = sin;
            const int e = threadIdx.x + 1;
#pragma unroll
            for (int b = 0; b < 8; ++b)
                if (e & (1 << b)) prop = mv(pw[f].p[b], prop);
            S.x += prop.x;
            S.y += prop.y;
            __syncthreads();
            sq[threadIdx.x] = S;
            __syncthreads();
            const V2 init = threadIdx.x > 0 ? sq[threadIdx.x - 1] : sin;
            y1[f] = (float)init.x;
            y2[f] = (float)init.y;
            __syncthreads();  // sq is reused by the next filter's scan
        }
    }
    if (chunk < n_chunks) {
        const long long n0 = chunk * L;
        float x1 = n0 >= 1 ? __ldg(xc + n0 - 1) : 0.f;
        float x2 = n0 >= 2 ? __ldg(xc + n0 - 2) : 0.f;
        float* row = tile + threadIdx.x * LDT;
        float* orow = otile + ((WRITE && NF > 1) ? threadIdx.x * LDT : 0);
#pragma unroll 8
        for (int i = 0; i < L; ++i) {
            const float x0 = row[i];
#pragma unroll
            for (int f = 0; f < NF; ++f) {
                const float yy = step(fas[f]->c, x0, x1, x2, y1[f], y2[f]);
                if (WRITE) {
                    const float yo = clamp ? fminf(fmaxf(yy, -1.f), 1.f) : yy;
                    if (f == 0) row[i] = yo;
                    else orow[i] = yo;
                    if (n0 + i < n) ssq[f] = fmaf(yo, yo, ssq[f]);
                    if (i == 0 && fas[f]->block_ss && (n0 == 0 || n0 == 1024))  // y[0], y[1024]: reflected edge of rms frame 0
                        fas[f]->block_ss[(long long)clip * (n_ss_blocks + 2) + n_ss_blocks + (n0 == 0 ? 0 : 1)] = yo;
                }
                y2[f] = y1[f];
                y1[f] = yy;
            }
            x2 = x1;
            x1 = x0;
        }
        if (!WRITE) {
#pragma unroll
            for (int f = 0; f < NF; ++f) fas[f]->state_out[(long long)clip * n_chunks + chunk] = make_float2(y1[f], y2[f]);
        }
    }
    if (!WRITE) {
        // CTA aggregate: zero-state end state of the CTA's whole span (chunks past the end contribute zero)
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            V2 z = {0.0, 0.0};
            if (chunk < n_chunks) z = {(double)y1[f], (double)y2[f]};
            const V2 q = block_scan_states(z, pw[f], sq);
            if (threadIdx.x == CPB - 1) fas[f]->cta_z[(long long)clip * gridDim.x + blockIdx.x] = q;
            __syncthreads();
        }
    }
    if (WRITE) {
        // sum of squares of the (clamped) output per 1024-sample block = 16 consecutive chunks = 16 lanes
        static_assert(L * 16 == 1024, "block_ss assumes 16 chunks per 1024 samples");
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            if (!fas[f]->block_ss) continue;
            float v = ssq[f];
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            const long long blk = (long long)blockIdx.x * (CPB / 16) + (threadIdx.x >> 4);
            if ((threadIdx.x & 15) == 0 && blk < n_ss_blocks) fas[f]->block_ss[(long long)clip * (n_ss_blocks + 2) + blk] = v;
        }
    }
    if (WRITE) {
        __syncthreads();
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            if (!fas[f]->y) continue;
            const float* src = f == 0 ? tile : otile;
            float* yc = fas[f]->y + (long long)clip * stride_y;
            if (tile0 + (long long)CPB * L <= n && ((reinterpret_cast<uintptr_t>(yc + tile0) & 15) == 0)) {
                float4* __restrict__ y4 = reinterpret_cast<float4*>(yc + tile0);
#pragma unroll 4
                for (int q = threadIdx.x; q < CPB * L / 4; q += CPB) {
                    const float* d = src + ((4 * q) / L) * LDT + ((4 * q) % L);
                    y4[q] = make_float4(d[0], d[1], d[2], d[3]);
                }
            } else {
                for (int i = threadIdx.x; i < CPB * L; i += CPB) {
                    const long long g = tile0 + i;
                    if (g < n) yc[g] = src[(i / L) * LDT + (i % L)];
                }
            }
        }
    }
}

// One CTA (1024 threads) per clip over the per-CTA aggregates of the chunk kernel (n_chunks here = number of chunk-kernel
// CTAs per clip, 720 for a 4-minute clip: one element per thread, coalesced).  z (in): zero-state end state of every
// span of CPB chunks; s (out): true state at the end of every span.
__global__ void __launch_bounds__(1024) biquad_carry_kernel(const V2* __restrict__ z, V2* __restrict__ s,
                                                            long long n_chunks, double a1, double a2) {
    __shared__ M2 sP[1024];
    __shared__ V2 sq[1024];
    const int clip = blockIdx.x, tid = threadIdx.x;
    const V2* zc = z + (long long)clip * n_chunks;
    V2* sc = s + (long long)clip * n_chunks;
    // A = companion matrix on (y[n-1], y[n-2]);  AL = A^(L * CPB) by repeated squaring (64 * 128 = 2^13)
    M2 AL = {-a1, -a2, 1.0, 0.0};
#pragma unroll
    for (int i = 0; i < 13; ++i) AL = mm(AL, AL);
    const long long run = (n_chunks + 1023) / 1024;
    const long long c0 = (long long)tid * run;
    long long c1 = c0 + run;
    if (c1 > n_chunks) c1 = n_chunks;
    // local aggregate of this thread's run: S_out = P S_in + q
    M2 P = {1.0, 0.0, 0.0, 1.0};
    V2 q = {0.0, 0.0};
    for (long long c = c0; c < c1; ++c) {
        const V2 zz = zc[c];
        q = mv(AL, q);
        q.x += zz.x;
        q.y += zz.y;
        P = mm(AL, P);
    }
    sP[tid] = P;
    sq[tid] = q;
    __syncthreads();
    // inclusive Hillis-Steele scan of affine maps: (P2,q2) o (P1,q1) = (P2 P1, P2 q1 + q2)
    for (int d = 1; d < 1024; d <<= 1) {
        M2 Pn = sP[tid];
        V2 qn = sq[tid];
        if (tid >= d) {
            const M2 Pl = sP[tid - d];
            const V2 ql = sq[tid - d];
            const V2 t = mv(Pn, ql);
            qn = {t.x + qn.x, t.y + qn.y};
            Pn = mm(Pn, Pl);
        }
        __syncthreads();
        sP[tid] = Pn;
        sq[tid] = qn;
        __syncthreads();
    }
    // state entering this thread's run = inclusive result of the previous thread applied to the zero state = its q
    V2 S = {0.0, 0.0};
    if (tid > 0) S = sq[tid - 1];
    for (long long c = c0; c < c1; ++c) {
        const V2 zz = zc[c];
        S = mv(AL, S);
        S.x += zz.x;
        S.y += zz.y;
        sc[c] = S;
    }
}

// rms(y, frame_length 2048, hop 1024, centre, reflect) (features/audio.py:31-37) from the per-1024-sample sums of squares
// the biquad's output pass leaves behind: frame t >= 1 covers blocks t - 1 and t; frame 0 covers the reflection
// y[1024..1] followed by y[0..1023].  y is only read for that one edge frame (two samples per clip).
__global__ void rms_from_block_ss_kernel(const float* __restrict__ ss, long long n_blocks, int n_frames,
                                         float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float* s = ss + (long long)clip * (n_blocks + 2);  // [n_blocks sums | y[0] | y[1024]]
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n_frames; t += gridDim.x * blockDim.x) {
        float acc;
        if (t == 0) {
            const float a = s[n_blocks], b = s[n_blocks + 1];
            acc = s[0] + ((s[0] - a * a) + b * b);
        } else {
            acc = s[t - 1] + s[t];
        }
        out[(long long)clip * n_frames + t] = sqrtf(acc / 2048.f);
    }
}

}  // namespace

static long long biquad_scratch_one(int n_clips, long long n_samples) {
    const long long n_chunks = (n_samples + L - 1) / L;
    const long long n_ctas = (n_chunks + CPB - 1) / CPB;
    return (((long long)n_clips * n_chunks * (long long)sizeof(float2) + 15) / 16) * 16 + 2LL * n_clips * n_ctas * 16;
}

extern "C" long long ssar_biquad_scratch_bytes(int n_clips, long long n_samples) {
    return 2 * biquad_scratch_one(n_clips, n_samples) + 16;  // room for a filter pair
}

struct FilterSpec {
    float b0, b1, b2, a1, a2;
    float* y;
    float* block_ss;
};

// n_filters = 1 or 2 filters over the same input: phase A (zero-state chunks), the per-clip carry scan(s), phase C
static int biquad_impl(const float* x, long long n_samples, long long stride_x, int n_clips, const FilterSpec* specs,
                       int n_filters, int clamp, long long stride_y, void* scratch, void* stream) {
    if (n_clips == 0 || n_samples == 0) return SSAR_OK;
    SSAR_REQUIRE(x && scratch, "biquad: null pointer");
    SSAR_REQUIRE(n_clips <= 65535, "biquad: too many clips per launch");
    cudaStream_t st = (cudaStream_t)stream;
    const long long n_chunks = (n_samples + L - 1) / L;
    const long long n_ctas = (n_chunks + CPB - 1) / CPB;
    const long long n_ss = (n_samples + 1023) / 1024;
    // scratch per filter: z [n_clips, n_chunks] float2 | per-CTA aggregates [n_clips, n_ctas] double2 | per-CTA states
    FilterArgs fa[2];
    for (int f = 0; f < 2; ++f) {
        const FilterSpec& sp = specs[f < n_filters ? f : 0];
        SSAR_REQUIRE(f >= n_filters || sp.y || sp.block_ss, "biquad: a filter needs an output (signal or block sums)");
        char* base = reinterpret_cast<char*>(scratch) + (long long)f * biquad_scratch_one(n_clips, n_samples);
        float2* z = reinterpret_cast<float2*>(base);
        V2* cz = reinterpret_cast<V2*>(base + (((long long)n_clips * n_chunks * (long long)sizeof(float2) + 15) / 16) * 16);
        V2* cs = cz + (long long)n_clips * n_ctas;
        fa[f].c = {sp.b0, sp.b1, sp.b2, sp.a1, sp.a2};
        fa[f].state_in = z;
        fa[f].state_out = z;
        fa[f].y = sp.y;
        fa[f].cta_z = cz;
        fa[f].cta_s = cs;
        fa[f].a1d = (double)sp.a1;
        fa[f].a2d = (double)sp.a2;
        fa[f].block_ss = sp.block_ss;
    }
    dim3 grid((unsigned)n_ctas, n_clips);
    if (n_filters == 2) biquad_chunks_kernel<false, 2><<<grid, CPB, 0, st>>>(x, n_samples, stride_x, fa[0], fa[1], stride_y, n_chunks, 0, n_ss);
    else biquad_chunks_kernel<false, 1><<<grid, CPB, 0, st>>>(x, n_samples, stride_x, fa[0], fa[0], stride_y, n_chunks, 0, n_ss);
    for (int f = 0; f < n_filters; ++f)
        biquad_carry_kernel<<<n_clips, 1024, 0, st>>>(fa[f].cta_z, const_cast<V2*>(fa[f].cta_s), n_ctas, fa[f].a1d, fa[f].a2d);
    if (n_filters == 2) {
        const size_t osm = (size_t)CPB * LDT * sizeof(float);  // + 35.6 KB static: above the 48 KB default
        cudaError_t e = cudaFuncSetAttribute(biquad_chunks_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)osm);
        if (e != cudaSuccess) {
            ssar_set_error("biquad: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return SSAR_ERR_CUDA;
        }
        biquad_chunks_kernel<true, 2><<<grid, CPB, osm, st>>>(x, n_samples, stride_x, fa[0], fa[1], stride_y, n_chunks, clamp, n_ss);
    }
    else biquad_chunks_kernel<true, 1><<<grid, CPB, 0, st>>>(x, n_samples, stride_x, fa[0], fa[0], stride_y, n_chunks, clamp, n_ss);
    return ssar_launch_status("biquad");
}

extern "C" int ssar_biquad(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                           float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* scratch,
                           void* stream) {
    SSAR_REQUIRE(y != nullptr, "biquad: null pointer");
    const FilterSpec sp = {b0, b1, b2, a1, a2, y, nullptr};
    return biquad_impl(x, n_samples, stride_x, n_clips, &sp, 1, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_biquad_ss(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                              float b2, float a1, float a2, int clamp, float* y, long long stride_y, float* block_ss,
                              void* scratch, void* stream) {
    SSAR_REQUIRE(block_ss != nullptr, "biquad_ss: null block_ss");
    const FilterSpec sp = {b0, b1, b2, a1, a2, y, block_ss};
    return biquad_impl(x, n_samples, stride_x, n_clips, &sp, 1, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_biquad_pair(const float* x, long long n_samples, long long stride_x, int n_clips, const float* coeffs,
                                int clamp, float* y0, float* y1, long long stride_y, float* block_ss0, float* block_ss1,
                                void* scratch, void* stream) {
    SSAR_REQUIRE(coeffs != nullptr, "biquad_pair: null coefficients");
    const FilterSpec sp[2] = {{coeffs[0], coeffs[1], coeffs[2], coeffs[3], coeffs[4], y0, block_ss0},
                              {coeffs[5], coeffs[6], coeffs[7], coeffs[8], coeffs[9], y1, block_ss1}};
    return biquad_impl(x, n_samples, stride_x, n_clips, sp, 2, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_rms_from_block_ss(const float* block_ss, long long n_blocks, int n_clips, int n_frames, float* out,
                                      void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(block_ss && out, "rms_from_block_ss: null pointer");
    SSAR_REQUIRE(n_frames <= n_blocks && n_blocks >= 2, "rms_from_block_ss: frame t needs blocks t - 1 and t");
    rms_from_block_ss_kernel<<<dim3(ssar_div_up(n_frames, 256), n_clips), 256, 0, (cudaStream_t)stream>>>(
        block_ss, n_blocks, n_frames, out);
    return ssar_launch_status("rms_from_block_ss");
}

// ---------------------------------------------------------------------------------------------------------------
// ssar_biquad_static: ssar_biquad for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.  The scan scratch
// lives in a file-scope device array of SSAR_BIQUAD_STATIC_SCRATCH_BYTES (16 MiB; every GPU that loads the library holds
// it until exit).  Every call shares the array and it keeps the last call's contents: the scan writes every word it
// later reads, and anything that has to start from zero is cleared on the stream before the kernels; calls that overlap
// on one GPU (two streams, two threads) race on it.
// ---------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace {
constexpr long long STATIC_SCRATCH_BYTES = SSAR_BIQUAD_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_biquad_static_scratch[STATIC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_static_mu;
void* g_static_ptr[MAX_DEVICES] = {};  // device address of the array on every device, resolved once per device

// In host code the array's name is a host-side stand-in, not a device address: resolve it with the device current.
int static_scratch_ptr(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        ssar_set_error("biquad_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return SSAR_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_biquad_static_scratch);
        if (e != cudaSuccess || !p) {
            ssar_set_error("biquad_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return SSAR_ERR_CUDA;
        }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return SSAR_OK;
}
}  // namespace

extern "C" int ssar_biquad_static(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                                  float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* stream) {
    if (n_clips == 0 || n_samples == 0) return SSAR_OK;
    SSAR_REQUIRE(x != nullptr && y != nullptr, "biquad_static: null pointer");
    SSAR_REQUIRE(n_clips > 0 && n_samples > 0, "biquad_static: negative size");
    const long long need = ssar_biquad_scratch_bytes(n_clips, n_samples);
    SSAR_REQUIRE(need <= STATIC_SCRATCH_BYTES,
                 "biquad_static: %d clips x %lld samples need %lld bytes of scratch, the static array holds %lld "
                 "(use ssar_biquad with caller-provided scratch)", n_clips, n_samples, need, STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    const int rc = static_scratch_ptr(&scratch);
    if (rc != SSAR_OK) return rc;
    const FilterSpec sp = {b0, b1, b2, a1, a2, y, nullptr};
    return biquad_impl(x, n_samples, stride_x, n_clips, &sp, 1, clamp, stride_y, scratch, stream);
}
