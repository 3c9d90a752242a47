// In-shared-memory Stockham FFT building blocks (radix 16/4/2 butterflies held in registers,
// exchanged through padded shared memory).  A "frame group" of TPF threads owns one frame:
//   M = n_fft/2 complex points (real input packed as z[n] = x[2n] + i x[2n+1]),
//   M = 1024 -> passes 16,16,4 with TPF = 64;  M = 512 -> passes 16,16,2 with TPF = 32.
// All frame groups of a CTA run the passes in lock-step, so plain __syncthreads() is the exchange barrier.
#pragma once
#include "common.cuh"

// padded shared-memory index: one extra float2 every 16 keeps the stride-16 Stockham writes conflict-free
__device__ __forceinline__ int PADI(int i) { return i + (i >> 4); }
__host__ __device__ constexpr int fft_padded_len(int n) { return n + (n >> 4) + 2; }

__device__ __forceinline__ void fft2(float2& a, float2& b) {
    float2 t = a;
    a = cadd(t, b);
    b = csub(t, b);
}

// forward 4-point DFT (e^{-i...}); outputs in natural order
__device__ __forceinline__ void fft4(float2& a, float2& b, float2& c, float2& d) {
    float2 t0 = cadd(a, c), t1 = csub(a, c), t2 = cadd(b, d);
    float2 bd = csub(b, d);
    float2 t3 = make_float2(bd.y, -bd.x);  // (b - d) * (-i)
    a = cadd(t0, t2);
    b = cadd(t1, t3);
    c = csub(t0, t2);
    d = csub(t1, t3);
}

// forward 16-point DFT as 4x4; v[] in natural order on entry and exit
__device__ __forceinline__ void fft16(float2* v) {
    // W16^m = exp(-2*pi*i*m/16)
    const float C1 = 0.92387953251128673848f, S1 = 0.38268343236508978178f, H = 0.70710678118654752440f;
#pragma unroll
    for (int c = 0; c < 4; ++c) fft4(v[c], v[c + 4], v[c + 8], v[c + 12]);
    // v[n2 + 4*k1] *= W16^(n2*k1)
    v[5] = cmul(v[5], make_float2(C1, -S1));    // 1*1
    v[6] = cmul(v[6], make_float2(H, -H));      // 2*1
    v[7] = cmul(v[7], make_float2(S1, -C1));    // 3*1
    v[9] = cmul(v[9], make_float2(H, -H));      // 1*2
    v[10] = make_float2(v[10].y, -v[10].x);     // 2*2 -> W16^4 = -i
    v[11] = cmul(v[11], make_float2(-H, -H));   // 3*2 -> W16^6
    v[13] = cmul(v[13], make_float2(S1, -C1));  // 1*3
    v[14] = cmul(v[14], make_float2(-H, -H));   // 2*3 -> W16^6
    v[15] = cmul(v[15], make_float2(-C1, S1));  // 3*3 -> W16^9
#pragma unroll
    for (int k1 = 0; k1 < 4; ++k1) fft4(v[4 * k1], v[4 * k1 + 1], v[4 * k1 + 2], v[4 * k1 + 3]);
    // X[k1 + 4*k2] sits at v[4*k1 + k2]: transpose the 4x4 in registers
    float2 t;
#define SSAR_SWAP(i, j) t = v[i]; v[i] = v[j]; v[j] = t;
    SSAR_SWAP(1, 4) SSAR_SWAP(2, 8) SSAR_SWAP(3, 12) SSAR_SWAP(6, 9) SSAR_SWAP(7, 13) SSAR_SWAP(11, 14)
#undef SSAR_SWAP
}

template <int R>
__device__ __forceinline__ void butterfly(float2* v) {
    if (R == 16) fft16(v);
    else if (R == 4) fft4(v[0], v[1], v[2], v[3]);
    else fft2(v[0], v[1]);
}

// Twiddle and butterfly the R values of Stockham butterfly j (stride Ns), then scatter to shared memory.
// tw[k] = exp(-2*pi*i*k/(2M)), k in [0, 2M)
template <int M, int R>
__device__ __forceinline__ void stockham_finish(float2* v, float2* s, const float2* __restrict__ tw, int j, int Ns) {
    const int k = j & (Ns - 1);  // Ns is a power of two
    if (Ns > 1) {
        const int step = 2 * (M / R) / Ns * k;  // exp(-2 pi i k r/(Ns R)) = tw[r*step]
#pragma unroll
        for (int r = 1; r < R; ++r) v[r] = cmul(v[r], __ldg(&tw[r * step]));
    }
    butterfly<R>(v);
    const int j0 = (j - k) * R + k;
#pragma unroll
    for (int r = 0; r < R; ++r) s[PADI(j0 + r * Ns)] = v[r];
}

// One Stockham pass whose inputs already sit in shared memory.  Ends with the data written and a barrier.
template <int M, int R, int TPF>
__device__ __forceinline__ void stockham_pass(float2* s, const float2* __restrict__ tw, int lt, int Ns) {
    constexpr int NB = M / R;
    constexpr int BPT = NB / TPF;
    static_assert(BPT >= 1 && BPT * TPF == NB, "pass shape");
    float2 v[BPT][R];
#pragma unroll
    for (int b = 0; b < BPT; ++b) {
        const int j = lt + b * TPF;
#pragma unroll
        for (int r = 0; r < R; ++r) v[b][r] = s[PADI(j + r * NB)];
    }
    __syncthreads();
#pragma unroll
    for (int b = 0; b < BPT; ++b) stockham_finish<M, R>(v[b], s, tw, lt + b * TPF, Ns);
    __syncthreads();
}

template <int M> struct FftShape;
template <> struct FftShape<1024> { static constexpr int TPF = 64, R1 = 16, R2 = 16, R3 = 4; };
template <> struct FftShape<512> { static constexpr int TPF = 32, R1 = 16, R2 = 16, R3 = 2; };

// Passes 2 and 3 (pass 1 is done by the caller because its inputs come straight from registers).
template <int M>
__device__ __forceinline__ void fft_tail_passes(float2* s, const float2* __restrict__ tw, int lt) {
    using S = FftShape<M>;
    stockham_pass<M, S::R2, S::TPF>(s, tw, lt, S::R1);
    stockham_pass<M, S::R3, S::TPF>(s, tw, lt, S::R1 * S::R2);
}
