// Polyphase windowed-sinc resampler (banded taps).
// Replaces torchaudio.functional.resample at ssar/supervised/data.py:63 (44.1 kHz -> fps*1024, sinc_interp_hann)
// and rosa/constantq.py:83 (octave decimation by 2, kaiser_window).  torchaudio builds a dense
// [new, 1, 2*width+orig] kernel and runs conv1d with stride orig; only ~2*width taps per phase are non-zero, so
// the plan keeps those (tap-major [max_taps, new] for coalesced reads) plus the first tap index of every phase.
#include "common.cuh"

namespace {

// Ragged batches: clip c holds clip_n_in[c] <= n_in samples and ZEROS behind them (the convention of every ragged
// producer), so the filter sees exactly torchaudio's zero padding; only the outputs past ceil(newf * clip_n_in[c] / orig)
// — the filter's ringing into the padding, which a per-clip call never produces — have to be forced to zero.
__device__ __forceinline__ long long ragged_n_out(const long long* __restrict__ clip_n_in, int clip, int orig, int newf,
                                                  long long n_out) {
    if (!clip_n_in) return n_out;
    const long long n = __ldg(clip_n_in + clip);
    return min(n_out, (n * newf + orig - 1) / orig);
}

__global__ void resample_kernel(const float* __restrict__ x, long long n_in, long long stride_in,
                                const float* __restrict__ taps, const int* __restrict__ first, int orig, int newf,
                                int width, int max_taps, float post_scale, float* __restrict__ y, long long n_out,
                                long long stride_out, const long long* __restrict__ clip_n_in) {
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * stride_in;
    float* yc = y + (long long)clip * stride_out;
    const long long n_out_c = ragged_n_out(clip_n_in, clip, orig, newf, n_out);
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n_out; j += (long long)gridDim.x * blockDim.x) {
        const long long blk = j / newf;
        const int ph = (int)(j - blk * newf);
        // padded coordinate p = blk*orig + k  <->  input index p - width
        const long long base = blk * orig - width + __ldg(first + ph);
        float acc = 0.f;
        for (int k = 0; k < max_taps; ++k) {
            const long long i = base + k;
            const float v = (i >= 0 && i < n_in) ? __ldg(xc + i) : 0.f;
            acc = fmaf(__ldg(taps + (long long)k * newf + ph), v, acc);
        }
        yc[j] = j < n_out_c ? acc * post_scale : 0.f;
    }
}

// v1: the launch's thread count is a multiple of `newf`, so every output a thread walks to (stride = thread count) has
// the SAME phase: its <= 32 taps are loaded once into registers, and an interior output costs max_taps coalesced loads +
// FMAs with no index arithmetic in the tap loop (same ascending-tap fmaf chain as v0: bit-identical results).
template <int MAXT>
__global__ void __launch_bounds__(256) resample_phase_kernel(const float* __restrict__ x, long long n_in, long long stride_in,
                                                             const float* __restrict__ taps, const int* __restrict__ first,
                                                             int orig, int newf, int width, int max_taps, float post_scale,
                                                             float* __restrict__ y, long long n_out, long long stride_out,
                                                             const long long* __restrict__ clip_n_in) {
    const int clip = blockIdx.y;
    const float* __restrict__ xc = x + (long long)clip * stride_in;
    float* __restrict__ yc = y + (long long)clip * stride_out;
    const long long n_out_c = ragged_n_out(clip_n_in, clip, orig, newf, n_out);
    const long long S = (long long)gridDim.x * blockDim.x;  // multiple of newf
    const long long j0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j0 >= n_out) return;
    const int ph = (int)(j0 % newf);
    float tp[MAXT];
#pragma unroll
    for (int k = 0; k < MAXT; ++k) tp[k] = k < max_taps ? __ldg(taps + (long long)k * newf + ph) : 0.f;
    const long long dbase = (S / newf) * orig;
    long long base = (j0 / newf) * orig - width + __ldg(first + ph);
    for (long long j = j0; j < n_out; j += S, base += dbase) {
        float acc = 0.f;
        if (base >= 0 && base + max_taps <= n_in) {
            const float* __restrict__ p = xc + base;
#pragma unroll
            for (int k = 0; k < MAXT; ++k)
                if (k < max_taps) acc = fmaf(tp[k], __ldg(p + k), acc);
        } else {
#pragma unroll
            for (int k = 0; k < MAXT; ++k) {
                if (k < max_taps) {
                    const long long i = base + k;
                    acc = fmaf(tp[k], (i >= 0 && i < n_in) ? __ldg(xc + i) : 0.f, acc);
                }
            }
        }
        yc[j] = j < n_out_c ? acc * post_scale : 0.f;
    }
}

// Decimation by 2 (orig 2, new 1: the CQT octave chain, rosa/constantq.py:83): every output uses the same taps and the
// windows of neighbouring outputs overlap by max_taps - 2 samples, so a thread makes 4 consecutive outputs from one
// register window of max_taps + 6 samples.  Per output the taps are accumulated in the same ascending fmaf chain as the
// generic kernels.  Read straight from global memory those windows start 8 samples apart across lanes, and every load
// instruction touches 32 different sectors (measured: L1 throughput 97 %, 13 % of HBM), so:
// a CTA stages the
// 2 * 1024 + max_taps - 2 samples its 1024 outputs need with coalesced loads into shared memory, 2 pad floats after
// every 32 samples, and each thread reads its window as 64-bit words: lane t -> bank 8 (t % 4) + 2 (t / 4) + 2 m, all
// 16 lanes of a half-warp on different bank pairs -> conflict-free.  Same ascending fmaf chain per output.
constexpr int D2_OUT = 1024;  // outputs per CTA tile (4 per thread)

template <int MAXT>
__global__ void __launch_bounds__(256) decimate2_staged_kernel(const float* __restrict__ x, long long n_in, long long stride_in,
                                                               const float* __restrict__ taps, const int* __restrict__ first,
                                                               int width, int max_taps, float post_scale,
                                                               float* __restrict__ y, long long n_out, long long stride_out,
                                                               long long n_tiles, const long long* __restrict__ clip_n_in) {
    constexpr int SPAN = 2 * D2_OUT + MAXT + 6;          // samples staged per tile (upper bound)
    constexpr int NW = (MAXT + 6) / 2;                   // 64-bit words per thread window
    __shared__ float2 S2[(SPAN + 2 * (SPAN / 32) + 2 + 1) / 2];
    float* S = reinterpret_cast<float*>(S2);
    const int clip = blockIdx.y;
    const float* __restrict__ xc = x + (long long)clip * stride_in;
    float* __restrict__ yc = y + (long long)clip * stride_out;
    const long long n_out_c = ragged_n_out(clip_n_in, clip, 2, 1, n_out);
    float tp[MAXT];
#pragma unroll
    for (int k = 0; k < MAXT; ++k) tp[k] = k < max_taps ? __ldg(taps + k) : 0.f;
    const int c0 = __ldg(first) - width;
    const int span = 2 * D2_OUT + max_taps + 6;
    const int t = threadIdx.x;
    const int wbase = 8 * t + 2 * (t >> 2);              // padded index of the thread's window start (8 t)
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long j0 = tile * D2_OUT;
        const long long g0 = 2 * j0 + c0;
        if (g0 >= 0 && g0 + span <= n_in) {
            for (int i = t; i < span; i += 256) S[i + 2 * (i >> 5)] = __ldg(xc + g0 + i);
        } else {
            for (int i = t; i < span; i += 256) {
                const long long g = g0 + i;
                S[i + 2 * (i >> 5)] = (g >= 0 && g < n_in) ? __ldg(xc + g) : 0.f;
            }
        }
        __syncthreads();
        float w[MAXT + 6];
#pragma unroll
        for (int m = 0; m < NW; ++m) {
            // sample 8 t + 2 m sits at 8 t + 2 m + 2 ((8 t + 2 m) / 32) = wbase + 2 m + 2 ((8 (t % 4) + 2 m) / 32)
            const int a = 8 * (t & 3) + 2 * m;
            const float2 v = S2[(wbase + 2 * m + 2 * (a >> 5)) >> 1];
            w[2 * m] = v.x;
            w[2 * m + 1] = v.y;
        }
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < MAXT; ++k) {
            if (k < max_taps) {
#pragma unroll
                for (int r = 0; r < 4; ++r) acc[r] = fmaf(tp[k], w[k + 2 * r], acc[r]);
            }
        }
        const long long j = j0 + 4 * t;
        if (clip_n_in) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
                if (j + r >= n_out_c) acc[r] = 0.f;
        }
        if (j + 3 < n_out && ((reinterpret_cast<uintptr_t>(yc + j) & 15) == 0)) {
            *reinterpret_cast<float4*>(yc + j) =
                make_float4(acc[0] * post_scale, acc[1] * post_scale, acc[2] * post_scale, acc[3] * post_scale);
        } else {
#pragma unroll
            for (int r = 0; r < 4; ++r)
                if (j + r < n_out) yc[j + r] = acc[r] * post_scale;
        }
        __syncthreads();
    }
}

}  // namespace

extern "C" int ssar_resample(const float* x, long long n_in, long long stride_in, int n_clips, const float* taps,
                             const int* first, int orig, int newf, int width, int max_taps, float post_scale, float* y,
                             long long n_out, long long stride_out, const long long* clip_n_in, void* stream) {
    if (n_clips == 0 || n_out == 0) return SSAR_OK;
    SSAR_REQUIRE(x && taps && first && y, "resample: null pointer");
    SSAR_REQUIRE(orig > 0 && newf > 0 && max_taps > 0, "resample: bad rates");
    if (orig == 2 && newf == 1 && max_taps <= 28) {
        const long long n_tiles = (n_out + D2_OUT - 1) / D2_OUT;
        long long gq = n_tiles;
        const long long cap = ((long long)SSAR_NUM_SMS * 8 + n_clips - 1) / n_clips;
        if (gq > cap) gq = cap;
        decimate2_staged_kernel<28><<<dim3((unsigned)gq, n_clips), 256, 0, (cudaStream_t)stream>>>(
            x, n_in, stride_in, taps, first, width, max_taps, post_scale, y, n_out, stride_out, n_tiles, clip_n_in);
        return ssar_launch_status("resample");
    }
    if (max_taps <= 32 && (newf <= 256 ? 256 % newf == 0 : newf % 256 == 0)) {
        // thread count = multiple of newf and of 256: whole "phase periods" per grid row
        const long long period = newf <= 256 ? 1 : newf / 256;  // blocks per phase period
        long long gxp = ((n_out + 255) / 256 + period - 1) / period * period;
        const long long cap = (long long)SSAR_NUM_SMS * 16 / period * period;
        if (gxp > cap && cap > 0) gxp = cap;
        if (max_taps <= 24)
            resample_phase_kernel<24><<<dim3((unsigned)gxp, n_clips), 256, 0, (cudaStream_t)stream>>>(
                x, n_in, stride_in, taps, first, orig, newf, width, max_taps, post_scale, y, n_out, stride_out, clip_n_in);
        else
            resample_phase_kernel<32><<<dim3((unsigned)gxp, n_clips), 256, 0, (cudaStream_t)stream>>>(
                x, n_in, stride_in, taps, first, orig, newf, width, max_taps, post_scale, y, n_out, stride_out, clip_n_in);
        return ssar_launch_status("resample");
    }
    int gx = ssar_div_up(n_out, 256);
    if (gx > SSAR_NUM_SMS * 16) gx = SSAR_NUM_SMS * 16;
    resample_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(x, n_in, stride_in, taps, first, orig, newf, width,
                                                                        max_taps, post_scale, y, n_out, stride_out, clip_n_in);
    return ssar_launch_status("resample");
}
