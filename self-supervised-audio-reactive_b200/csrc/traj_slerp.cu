// Spherical interpolation loops written straight into the [size, n_layers, latent_dim] trajectory
// (maua/maua/audiovisual/audioreactive/latent.py:55-81 slerp / slerp_loops, the "slerp" of the north star):
//   knots  = cat([y] * n_loops + [y[0]])                                   (K' = K n_loops + 1 knots, S = K' - 1 segments)
//   t      = linspace(0, 1, n_t),  n_t = round(size / K')
//   row(ti, s) = normalize(a cos(t theta) + c sin(t theta)),  a = y_s / |y_s|, b = y_{s+1} / |y_{s+1}|, theta = acos(a.b),
//                c = normalize(b - (a.b) a)           -- rows are ordered ti-major, exactly as the reference's reshape leaves them
//   out    = F.interpolate(rows, size, mode="linear", align_corners=False) along the row axis.
// Kernel 1 (tiny): per (segment, layer) the unit vectors a, c and the angle.  Kernel 2: one warp per output (frame,
// layer) evaluates the two source rows it blends and writes 4 latent_dim bytes once — HBM-write-bound, no [rows, W, L]
// intermediate (the reference materialises it and reads it back).
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) slerp_prep_kernel(const float* __restrict__ y, int K, int W, int L, int S,
                                                         float* __restrict__ a_hat, float* __restrict__ c_hat,
                                                         float* __restrict__ theta) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= S * W) return;
    const int s = warp / W, w = warp - s * W;
    const float* a = y + ((long long)(s % K) * W + w) * L;
    const float* b = y + ((long long)((s + 1) % K) * W + w) * L;
    float na = 0.f, nb = 0.f;
    for (int i = lane; i < L; i += 32) {
        na = fmaf(a[i], a[i], na);
        nb = fmaf(b[i], b[i], nb);
    }
    na = sqrtf(warp_sum(na));
    nb = sqrtf(warp_sum(nb));
    float d = 0.f;
    for (int i = lane; i < L; i += 32) d = fmaf(a[i] / na, b[i] / nb, d);
    d = warp_sum(d);
    float nc = 0.f;
    for (int i = lane; i < L; i += 32) {
        const float c = b[i] / nb - d * (a[i] / na);
        nc = fmaf(c, c, nc);
    }
    nc = sqrtf(warp_sum(nc));
    float* ao = a_hat + (long long)warp * L;
    float* co = c_hat + (long long)warp * L;
    for (int i = lane; i < L; i += 32) {
        ao[i] = a[i] / na;
        co[i] = (b[i] / nb - d * (a[i] / na)) / nc;
    }
    if (lane == 0) theta[warp] = acosf(d);
}

__device__ __forceinline__ float linspace01(int i, int steps) {
    // torch.linspace(0, 1, steps) in fp32: start + i * step for the first half, end - (steps - 1 - i) * step after
    if (steps <= 1) return 0.f;
    const float step = 1.0f / (float)(steps - 1);
    return i < steps / 2 ? (float)i * step : 1.0f - (float)(steps - 1 - i) * step;
}

template <int VPL>  // floats per lane: latent_dim = 32 * VPL
__global__ void __launch_bounds__(256) slerp_rows_kernel(const float* __restrict__ a_hat, const float* __restrict__ c_hat,
                                                         const float* __restrict__ theta, int W, int S, int n_t,
                                                         int size, float* __restrict__ out) {
    constexpr int L = 32 * VPL;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (warp >= (long long)size * W) return;
    const int n = (int)(warp / W), w = (int)(warp - (long long)n * W);
    const int M = n_t * S;
    // F.interpolate(mode="linear", align_corners=False): src = scale * (dst + 0.5) - 0.5, clamped at 0
    const float scale = (float)M / (float)size;
    float src = scale * ((float)n + 0.5f) - 0.5f;
    if (src < 0.f) src = 0.f;
    const int i0 = (int)src;
    const int i1 = i0 + (i0 < M - 1 ? 1 : 0);
    const float lam1 = src - (float)i0, lam0 = 1.f - lam1;
    float v[2][VPL];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int row = r == 0 ? i0 : i1;
        const int ti = row / S, s = row - ti * S;
        const float p = linspace01(ti, n_t) * __ldg(theta + s * W + w);
        const float cp = cosf(p), sp = sinf(p);
        const float* a = a_hat + ((long long)s * W + w) * L;
        const float* c = c_hat + ((long long)s * W + w) * L;
        float nn = 0.f;
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
            const float x = __ldg(a + lane + 32 * j) * cp + __ldg(c + lane + 32 * j) * sp;
            v[r][j] = x;
            nn = fmaf(x, x, nn);
        }
        nn = sqrtf(warp_sum(nn));
#pragma unroll
        for (int j = 0; j < VPL; ++j) v[r][j] /= nn;
    }
    float* o = out + ((long long)n * W + w) * L;
#pragma unroll
    for (int j = 0; j < VPL; ++j) o[lane + 32 * j] = lam0 * v[0][j] + lam1 * v[1][j];
}

}  // namespace

extern "C" long long ssar_slerp_loops_scratch_floats(int n_latents, int n_loops, int n_layers, int latent_dim) {
    const long long S = (long long)n_latents * n_loops;
    return S * n_layers * (2LL * latent_dim + 1);
}

extern "C" int ssar_slerp_loops(const float* y, int n_latents, int n_layers, int latent_dim, int n_loops, int n_t, int size,
                                float* scratch, float* out, void* stream) {
    if (size == 0) return SSAR_OK;
    SSAR_REQUIRE(y && scratch && out, "slerp_loops: null pointer");
    SSAR_REQUIRE(n_latents > 0 && n_loops > 0 && n_layers > 0 && n_t > 0 && size > 0, "slerp_loops: bad sizes");
    SSAR_REQUIRE(latent_dim == 512 || latent_dim == 256 || latent_dim == 128 || latent_dim == 64 || latent_dim == 32,
                 "slerp_loops: latent_dim must be 32, 64, 128, 256 or 512");
    const int S = n_latents * n_loops;  // segments between the K n_loops + 1 knots
    float* a_hat = scratch;
    float* c_hat = a_hat + (long long)S * n_layers * latent_dim;
    float* theta = c_hat + (long long)S * n_layers * latent_dim;
    cudaStream_t st = (cudaStream_t)stream;
    slerp_prep_kernel<<<ssar_div_up((long long)S * n_layers, 8), 256, 0, st>>>(y, n_latents, n_layers, latent_dim, S, a_hat,
                                                                              c_hat, theta);
    const int grid = ssar_div_up((long long)size * n_layers, 8);
    switch (latent_dim) {
        case 512: slerp_rows_kernel<16><<<grid, 256, 0, st>>>(a_hat, c_hat, theta, n_layers, S, n_t, size, out); break;
        case 256: slerp_rows_kernel<8><<<grid, 256, 0, st>>>(a_hat, c_hat, theta, n_layers, S, n_t, size, out); break;
        case 128: slerp_rows_kernel<4><<<grid, 256, 0, st>>>(a_hat, c_hat, theta, n_layers, S, n_t, size, out); break;
        case 64: slerp_rows_kernel<2><<<grid, 256, 0, st>>>(a_hat, c_hat, theta, n_layers, S, n_t, size, out); break;
        default: slerp_rows_kernel<1><<<grid, 256, 0, st>>>(a_hat, c_hat, theta, n_layers, S, n_t, size, out); break;
    }
    return ssar_launch_status("slerp_loops");
}

// ---------------------------------------------------------------------------------------------------------------
// Cardinal weights of a closed natural-cubic-spline loop (ssar/random/latent.py:7-13): out[t, k] such that
// spline_loop_latents(y, size, n_loops)[t] = sum_k out[t, k] y[k].  The evaluation points linspace(0, n_loops, size) % 1
// (latent.py:10), the segment lookup (torch.bucketize, right = False) and the cubic in fp64 with explicit roundings
// follow ssar_b200/plan.py::spline_loop_weights_cpu; one launch instead of a dozen small torch kernels — a dataset of
// tracks of distinct lengths asks for new weights for every track.
//   coef: fp64 [K1 - 1, 4, K1] = (a, b, two_c, three_d) of the K1 cardinal splines per segment; knots: fp32 [K1]
// ---------------------------------------------------------------------------------------------------------------
namespace {
__global__ void spline_loop_weights_kernel(const double* __restrict__ coef, const float* __restrict__ knots, int K1, int size,
                                           float end, float* __restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= size) return;
    // torch.linspace(0, end, size): step = end / (size - 1); the first half counts up from the start, the second down
    // from the end (ATen RangeFactories)
    float v = 0.f;
    if (size > 1) {
        const float step = __fdiv_rn(end, (float)(size - 1));
        v = t < size / 2 ? __fmul_rn(step, (float)t) : __fsub_rn(end, __fmul_rn(step, (float)(size - 1 - t)));
    }
    v = fmodf(v, 1.f);  // torch.remainder for non-negative operands
    int lo = 0, hi = K1;  // bucketize(right=False): first i with knots[i] >= v
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(knots + mid) < v) lo = mid + 1;
        else hi = mid;
    }
    const int idx = min(max(lo - 1, 0), K1 - 2);
    const double f = (double)__fsub_rn(v, __ldg(knots + idx));
    const double* c = coef + (long long)idx * 4 * K1;
    const int K = K1 - 1;
    float* o = out + (long long)t * K;
    double w0 = 0.0;
    for (int k = 0; k < K1; ++k) {
        const double a = __ldg(c + k), b = __ldg(c + K1 + k), c2 = __ldg(c + 2 * K1 + k), d3 = __ldg(c + 3 * K1 + k);
        // a + (b + (0.5 c2 + d3 f / 3) f) f, one rounding per operation like the host expression
        const double inner = __dadd_rn(__dmul_rn(0.5, c2), __ddiv_rn(__dmul_rn(d3, f), 3.0));
        const double w = __dadd_rn(a, __dmul_rn(__dadd_rn(b, __dmul_rn(inner, f)), f));
        if (k == 0) w0 = w;
        else if (k < K) o[k] = (float)w;
        else w0 = __dadd_rn(w0, w);  // the closing knot is knot 0 again
    }
    o[0] = (float)w0;
}
}  // namespace

extern "C" int ssar_spline_loop_weights(const double* coef, const float* knots, int n_knots_closed, int size, float end,
                                        float* out, void* stream) {
    if (size == 0) return SSAR_OK;
    SSAR_REQUIRE(coef && knots && out && n_knots_closed >= 3 && size > 0, "spline_loop_weights: null pointer or fewer than 3 knots");
    spline_loop_weights_kernel<<<ssar_div_up(size, 128), 128, 0, (cudaStream_t)stream>>>(coef, knots, n_knots_closed, size, end, out);
    return ssar_launch_status("spline_loop_weights");
}
