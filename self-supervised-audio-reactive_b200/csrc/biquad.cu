// Second-order IIR (biquad) over whole clips as a three-phase blocked scan.
// Replaces torchaudio.functional.lowpass_biquad / highpass_biquad -> biquad -> lfilter(clamp=True) behind
// features/processing.py:141-150 (low_pass / mid_pass / high_pass):
//     v[n] = b0 x[n] + b1 x[n-1] + b2 x[n-2]          (coefficients already divided by a0)
//     y[n] = v[n] - (a2 y[n-2] + a1 y[n-1]),  y[-1] = y[-2] = 0,  output clamped to [-1, 1] at the very end.
//
// Phase A: every chunk of L samples runs the recurrence from a ZERO state and keeps only its final state z_c.
// Phase B: the true end-of-chunk states follow the affine recurrence S_c = A^L S_{c-1} + z_c (A = companion
//          matrix); one CTA per clip scans it in fp64 (per-thread runs + Hillis-Steele over (P, q) pairs).
// Phase C: every chunk re-runs the SAME fp32 recurrence from its true initial state S_{c-1} and writes clamp(y),
//          so within a chunk the operation order is the sequential reference's.
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
    const float v = c.b2 * x2 + c.b1 * x1 + c.b0 * x0;
    return v - (c.a2 * y2 + c.a1 * y1);
}

template <bool WRITE>
__global__ void __launch_bounds__(CPB) biquad_chunks_kernel(const float* __restrict__ x, long long n, long long stride_x,
                                                            Coef c, const float2* __restrict__ state_in,
                                                            float2* __restrict__ state_out, float* __restrict__ y,
                                                            long long stride_y, long long n_chunks, int clamp) {
    __shared__ float tile[CPB * LDT];
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * stride_x;
    const long long tile0 = (long long)blockIdx.x * CPB * L;
    for (int i = threadIdx.x; i < CPB * L; i += CPB) {
        const long long g = tile0 + i;
        tile[(i / L) * LDT + (i % L)] = g < n ? __ldg(xc + g) : 0.f;
    }
    __syncthreads();
    const long long chunk = (long long)blockIdx.x * CPB + threadIdx.x;
    if (chunk < n_chunks) {
        const long long n0 = chunk * L;
        float x1 = n0 >= 1 ? __ldg(xc + n0 - 1) : 0.f;
        float x2 = n0 >= 2 ? __ldg(xc + n0 - 2) : 0.f;
        float y1 = 0.f, y2 = 0.f;
        if (WRITE && chunk > 0) {
            const float2 s = state_in[(long long)clip * n_chunks + chunk - 1];
            y1 = s.x;
            y2 = s.y;
        }
        float* row = tile + threadIdx.x * LDT;
#pragma unroll 8
        for (int i = 0; i < L; ++i) {
            const float x0 = row[i];
            const float yy = step(c, x0, x1, x2, y1, y2);
            if (WRITE) row[i] = clamp ? fminf(fmaxf(yy, -1.f), 1.f) : yy;
            x2 = x1;
            x1 = x0;
            y2 = y1;
            y1 = yy;
        }
        if (!WRITE) state_out[(long long)clip * n_chunks + chunk] = make_float2(y1, y2);
    }
    if (WRITE) {
        __syncthreads();
        float* yc = y + (long long)clip * stride_y;
        for (int i = threadIdx.x; i < CPB * L; i += CPB) {
            const long long g = tile0 + i;
            if (g < n) yc[g] = tile[(i / L) * LDT + (i % L)];
        }
    }
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

// One CTA (1024 threads) per clip.  z (in): zero-state end states per chunk; s (out): true end states per chunk.
__global__ void __launch_bounds__(1024) biquad_carry_kernel(const float2* __restrict__ z, float2* __restrict__ s,
                                                            long long n_chunks, double a1, double a2) {
    __shared__ M2 sP[1024];
    __shared__ V2 sq[1024];
    const int clip = blockIdx.x, tid = threadIdx.x;
    const float2* zc = z + (long long)clip * n_chunks;
    float2* sc = s + (long long)clip * n_chunks;
    // A = companion matrix on (y[n-1], y[n-2]);  AL = A^L by repeated squaring (L = 64 = 2^6)
    M2 AL = {-a1, -a2, 1.0, 0.0};
#pragma unroll
    for (int i = 0; i < 6; ++i) AL = mm(AL, AL);
    const long long run = (n_chunks + 1023) / 1024;
    const long long c0 = (long long)tid * run;
    long long c1 = c0 + run;
    if (c1 > n_chunks) c1 = n_chunks;
    // local aggregate of this thread's run: S_out = P S_in + q
    M2 P = {1.0, 0.0, 0.0, 1.0};
    V2 q = {0.0, 0.0};
    for (long long c = c0; c < c1; ++c) {
        const float2 zz = zc[c];
        q = mv(AL, q);
        q.x += (double)zz.x;
        q.y += (double)zz.y;
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
        const float2 zz = zc[c];
        S = mv(AL, S);
        S.x += (double)zz.x;
        S.y += (double)zz.y;
        sc[c] = make_float2((float)S.x, (float)S.y);
    }
}

}  // namespace

extern "C" long long ssar_biquad_scratch_bytes(int n_clips, long long n_samples) {
    const long long n_chunks = (n_samples + L - 1) / L;
    return 2LL * n_clips * n_chunks * (long long)sizeof(float2);
}

extern "C" int ssar_biquad(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                           float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* scratch,
                           void* stream) {
    if (n_clips == 0 || n_samples == 0) return SSAR_OK;
    SSAR_REQUIRE(x && y && scratch, "biquad: null pointer");
    SSAR_REQUIRE(n_clips <= 65535, "biquad: too many clips per launch");
    cudaStream_t st = (cudaStream_t)stream;
    const long long n_chunks = (n_samples + L - 1) / L;
    float2* z = reinterpret_cast<float2*>(scratch);
    float2* s = z + (long long)n_clips * n_chunks;
    const Coef c = {b0, b1, b2, a1, a2};
    dim3 grid((unsigned)((n_chunks + CPB - 1) / CPB), n_clips);
    biquad_chunks_kernel<false><<<grid, CPB, 0, st>>>(x, n_samples, stride_x, c, nullptr, z, nullptr, 0, n_chunks, 0);
    biquad_carry_kernel<<<n_clips, 1024, 0, st>>>(z, s, n_chunks, (double)a1, (double)a2);
    biquad_chunks_kernel<true><<<grid, CPB, 0, st>>>(x, n_samples, stride_x, c, s, nullptr, y, stride_y, n_chunks, clamp);
    return ssar_launch_status("biquad");
}
