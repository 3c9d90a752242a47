// Warp-per-frame FFT building blocks: every lane holds 32 complex points in registers.
//
//   M = 1024 complex points (n_fft 2048): one frame per warp, M = 32 (lanes, n1) x 32 (registers, n2)
//   M =  512 complex points (n_fft 1024): two frames per warp, M = 16 (lanes of a half-warp, n1) x 32 (registers, n2)
//
// Cooley-Tukey with n = n1 + N1 n2, k = N2 k1 + k2:
//   pass 1: per lane, a 32-point DFT over n2 entirely in registers (radix-2 DIF, constants folded at compile time);
//   twiddle by W_M^{n1 k2} from a [k2][n1] table in shared memory (lanes read consecutive entries: conflict-free);
//   transpose through a padded warp-private shared-memory tile (only __syncwarp, no block barrier);
//   pass 2: per lane, DFTs over n1 (one 32-point or two 16-point) in registers; X[N2 k1 + k2] lands in natural order
//   in shared memory for the epilogues.
// Versus the block-cooperative radix-16/16/4 version this replaces: ~2.4x fewer instructions per frame, no
// __syncthreads in the frame loop, 32 independent loads in flight per thread.
#pragma once
#include "common.cuh"

// cos / sin of 2 pi k / 32, k = 0..31
__device__ constexpr float kCos32[32] = {
    1.0f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f, 0.70710678118654752f, 0.55557023301960218f,
    0.38268343236508977f, 0.19509032201612825f, 0.0f, -0.19509032201612825f, -0.38268343236508977f, -0.55557023301960218f,
    -0.70710678118654752f, -0.83146961230254524f, -0.92387953251128674f, -0.98078528040323043f, -1.0f,
    -0.98078528040323043f, -0.92387953251128674f, -0.83146961230254524f, -0.70710678118654752f, -0.55557023301960218f,
    -0.38268343236508977f, -0.19509032201612825f, 0.0f, 0.19509032201612825f, 0.38268343236508977f, 0.55557023301960218f,
    0.70710678118654752f, 0.83146961230254524f, 0.92387953251128674f, 0.98078528040323043f};
__device__ constexpr float kSin32[32] = {
    0.0f, 0.19509032201612825f, 0.38268343236508977f, 0.55557023301960218f, 0.70710678118654752f, 0.83146961230254524f,
    0.92387953251128674f, 0.98078528040323043f, 1.0f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f,
    0.70710678118654752f, 0.55557023301960218f, 0.38268343236508977f, 0.19509032201612825f, 0.0f,
    -0.19509032201612825f, -0.38268343236508977f, -0.55557023301960218f, -0.70710678118654752f, -0.83146961230254524f,
    -0.92387953251128674f, -0.98078528040323043f, -1.0f, -0.98078528040323043f, -0.92387953251128674f,
    -0.83146961230254524f, -0.70710678118654752f, -0.55557023301960218f, -0.38268343236508977f, -0.19509032201612825f};

__host__ __device__ constexpr int bitrev_const(int i, int bits) {
    int r = 0;
    for (int b = 0; b < bits; ++b) r |= ((i >> b) & 1) << (bits - 1 - b);
    return r;
}
__host__ __device__ constexpr int ilog2_const(int n) { return n <= 1 ? 0 : 1 + ilog2_const(n >> 1); }

// b <- b * exp(-+ 2 pi i E / 32)  (forward: minus sign; INV: plus), E a compile-time constant after unrolling
template <bool INV>
__device__ __forceinline__ float2 mul_w32(float2 b, int E) {
    E &= 31;
    if (E == 0) return b;
    if (E == 8) return INV ? make_float2(-b.y, b.x) : make_float2(b.y, -b.x);    // -+ i
    if (E == 16) return make_float2(-b.x, -b.y);
    if (E == 24) return INV ? make_float2(b.y, -b.x) : make_float2(-b.y, b.x);
    const float c = kCos32[E], s = INV ? kSin32[E] : -kSin32[E];  // w = c + i s
    if (E == 4 || E == 12 || E == 20 || E == 28) {
        // |c| == |s| == sqrt(1/2): two adds and two multiplies
        const float h = 0.70710678118654752f;
        const float sc = c > 0.f ? 1.f : -1.f, ss = s > 0.f ? 1.f : -1.f;
        return make_float2(h * (sc * b.x - ss * b.y), h * (ss * b.x + sc * b.y));
    }
    return make_float2(fmaf(c, b.x, -s * b.y), fmaf(c, b.y, s * b.x));
}

// In-register radix-2 decimation-in-frequency DFT of N (16 or 32) points: natural-order input,
// output X[bitrev(i)] in v[i].  Every loop is fully unrolled, so all indices and twiddles are compile-time constants.
template <int N, bool INV>
__device__ __forceinline__ void fft_dif_regs(float2 (&v)[N]) {
    static_assert(N == 32 || N == 16 || N == 8, "register FFT sizes");
#pragma unroll
    for (int len = N; len >= 2; len >>= 1) {
        const int half = len >> 1;
#pragma unroll
        for (int blk = 0; blk < N; blk += len) {
#pragma unroll
            for (int j = 0; j < half; ++j) {
                // (re, im) pairs add as ONE packed FADD2 (sm_100; IEEE round-to-nearest per component: the same bits as two
                // FADDs, half the issue slots)
                const float2 a = v[blk + j], b = v[blk + j + half];
                v[blk + j] = __fadd2_rn(a, b);
                v[blk + j + half] = mul_w32<INV>(__fadd2_rn(a, make_float2(-b.x, -b.y)), j * (32 / len));
            }
        }
    }
}

// Geometry of the warp-per-frame transform of M complex points.
template <int M>
struct WarpFft {
    static constexpr int N1 = (M == 1024) ? 32 : 16;   // lanes per frame (n1)
    static constexpr int FPW = 32 / N1;                // frames per warp
    static constexpr int TLD = N1 + 1;                 // padded row of the [k2][n1] transpose tile (float2)
    static constexpr int ZF = 32 * TLD;                // float2 per frame buffer (>= M + 1)
    static constexpr int WPB = 8;                      // warps per CTA
    static constexpr int NJ = M / 32 + 1;              // spectrum bins per lane (k = lane + 32 j), last one only lane 0
    static_assert(ZF >= M + 1, "frame buffer must hold M + 1 bins");
};

// twT[k2 * N1 + n1] = W_M^{n1 k2} from the global table tw[k] = exp(-2 pi i k / 2M); call with the whole CTA, then sync.
template <int M>
__device__ __forceinline__ void warp_fft_fill_twiddles(float2* twT, const float2* __restrict__ tw) {
    constexpr int N1 = WarpFft<M>::N1;
    for (int i = threadIdx.x; i < 32 * N1; i += blockDim.x) {
        const int k2 = i / N1, nn = i - k2 * N1;
        twT[i] = __ldg(&tw[2 * nn * k2]);
    }
}

// Forward DFT of the warp's frame(s).  In: v[r] = z[n1 + N1 r] of the frame in slot h = lane / N1 (n1 = lane % N1).
// Out: Z[k], k in [0, M), in natural order at zw[h * ZF + k].  Ends with a __syncwarp (results visible to the warp).
template <int M>
__device__ __forceinline__ void warp_fft_forward(float2 (&v)[32], float2* zw, const float2* twT, int lane) {
    using L = WarpFft<M>;
    constexpr int N1 = L::N1, TLD = L::TLD, ZF = L::ZF;
    const int h = lane / N1, n1 = lane % N1;
    fft_dif_regs<32, false>(v);
    {
        float2* tile = zw + h * ZF;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            const int k2 = bitrev_const(i, 5);
            float2 y = v[i];
            if (k2 != 0) y = cmul(y, twT[k2 * N1 + n1]);
            tile[k2 * TLD + n1] = y;
        }
    }
    __syncwarp();
    if constexpr (M == 1024) {
        float2* tile = zw;
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = tile[lane * TLD + i];
        fft_dif_regs<32, false>(v);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 32; ++i) tile[32 * bitrev_const(i, 5) + lane] = v[i];
    } else {
        float2* tile = zw + h * ZF;
        float2 u0[16], u1[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            u0[i] = tile[n1 * TLD + i];
            u1[i] = tile[(n1 + 16) * TLD + i];
        }
        fft_dif_regs<16, false>(u0);
        fft_dif_regs<16, false>(u1);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            tile[32 * bitrev_const(i, 4) + n1] = u0[i];
            tile[32 * bitrev_const(i, 4) + n1 + 16] = u1[i];
        }
    }
    __syncwarp();
}
