// Noise-map trajectories: the lazy module tree of ssar/random/noise.py (Loop / Blend / Multiply leaves combined by
// Average / Modulate / ScaleBias, :4-86) compiled by the host into a short accumulator program and evaluated per
// output pixel for a batch of frames [i, i + b) -> out [b, H*W] in ONE launch (plus one reduction launch when the
// program contains Loop leaves, which are RMS-normalised over the whole map, :49-53).
//
// Program model (one "acc" and one "cur" register per frame):
//   leaf ops (LOOP, BLEND, MULTIPLY, BUFFER) write cur (or acc when to_acc is set: the base of the chain)
//   AVERAGE   acc = (acc + cur) / 2                      (noise.py:62-63)
//   MODULATE  acc = acc * m[t] + cur * (1 - m[t])        (noise.py:73-75)
//   SCALEBIAS acc = scale * acc + bias                   (noise.py:85-86)
// Each thread owns one pixel and NF consecutive frames, so every noise-basis value is read once per NF frames.
#include "block_utils.cuh"

namespace {

constexpr int NF = 8;  // (16 frames per thread: 128 registers, full-res sweep 620 -> 607 ms, quarter-res slower: the kernel is bound by sinf / cosf, not by the basis reads)

__device__ __forceinline__ float loop_raw(const ssar_noise_op& op, int HW, long long p, float idx) {
    // noise.py:50-51: sin(cos(idx + n0) / (sigma / 50) + n1) * n2
    const float n0 = __ldg(op.noise + p), n1 = __ldg(op.noise + HW + p), n2 = __ldg(op.noise + 2LL * HW + p);
    return sinf(cosf(idx + n0) / op.p0 + n1) * n2;
}

// sum over the map of loop_raw^2 for every (LOOP op, frame) -> rms_scratch[op_slot * b + f]
__global__ void noise_loop_rms_kernel(const ssar_noise_op* __restrict__ ops, int n_ops, int i0, int b, int HW,
                                      float* __restrict__ sumsq) {
    __shared__ float scratch[32];
    const int f = blockIdx.x;
    // blockIdx.y indexes the y-th LOOP op of the program
    int seen = -1, oi = -1;
    for (int o = 0; o < n_ops; ++o)
        if (ops[o].kind == SSAR_NOISE_LOOP && ++seen == (int)blockIdx.y) {
            oi = o;
            break;
        }
    if (oi < 0) return;
    const ssar_noise_op op = ops[oi];
    const float idx = __ldg(op.mod + i0 + f);
    float acc = 0.f;
    for (long long p = (long long)blockIdx.z * blockDim.x + threadIdx.x; p < HW; p += (long long)gridDim.z * blockDim.x) {
        const float v = loop_raw(op, HW, p, idx);
        acc = fmaf(v, v, acc);
    }
    acc = block_sum(acc, scratch);
    if (threadIdx.x == 0) atomicAdd(&sumsq[blockIdx.y * b + f], acc);
}

__global__ void __launch_bounds__(256) noise_eval_kernel(const ssar_noise_op* __restrict__ ops, int n_ops, int i0, int b,
                                                         int HW, const float* __restrict__ loop_sumsq,
                                                         float* __restrict__ out) {
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= HW) return;
    const int f0 = blockIdx.y * NF;
    const int nf = min(NF, b - f0);
    float acc[NF], cur[NF];
#pragma unroll
    for (int f = 0; f < NF; ++f) acc[f] = cur[f] = 0.f;
    int loop_slot = 0;
    for (int o = 0; o < n_ops; ++o) {
        const ssar_noise_op op = ops[o];
        switch (op.kind) {
            case SSAR_NOISE_LOOP: {
                const float n0 = __ldg(op.noise + p), n1 = __ldg(op.noise + HW + p), n2 = __ldg(op.noise + 2LL * HW + p);
#pragma unroll
                for (int f = 0; f < NF; ++f) {
                    if (f < nf) {
                        const float raw = sinf(cosf(__ldg(op.mod + i0 + f0 + f) + n0) / op.p0 + n1) * n2;
                        const float rms = sqrtf(loop_sumsq[loop_slot * b + f0 + f] / (float)HW);
                        cur[f] = raw / (rms + 1.1920928955078125e-07f);  // finfo(float32).eps
                    }
                }
                ++loop_slot;
                break;
            }
            case SSAR_NOISE_BLEND: {
                float l[NF], r[NF];
#pragma unroll
                for (int f = 0; f < NF; ++f) l[f] = r[f] = 0.f;
                for (int m = 0; m < op.M; ++m) {
                    const float a = __ldg(op.noise + (long long)m * HW + p);
                    const float c = __ldg(op.noise + (long long)(op.M + m) * HW + p);
#pragma unroll
                    for (int f = 0; f < NF; ++f) {
                        if (f < nf) {
                            const float w = __ldg(op.mod + (long long)(i0 + f0 + f) * op.M + m);
                            l[f] = fmaf(a, w, l[f]);
                            r[f] = fmaf(c, 1.f - w, r[f]);
                        }
                    }
                }
#pragma unroll
                for (int f = 0; f < NF; ++f) cur[f] = l[f] + r[f];
                break;
            }
            case SSAR_NOISE_MULTIPLY: {
#pragma unroll
                for (int f = 0; f < NF; ++f) cur[f] = 0.f;
                for (int m = 0; m < op.M; ++m) {
                    const float a = __ldg(op.noise + (long long)m * HW + p);
#pragma unroll
                    for (int f = 0; f < NF; ++f)
                        if (f < nf) cur[f] = fmaf(a, __ldg(op.mod + (long long)(i0 + f0 + f) * op.M + m), cur[f]);
                }
                break;
            }
            case SSAR_NOISE_BUFFER: {
#pragma unroll
                for (int f = 0; f < NF; ++f)
                    if (f < nf) cur[f] = __ldg(op.noise + (long long)(f0 + f) * HW + p);
                break;
            }
            case SSAR_NOISE_AVERAGE: {
#pragma unroll
                for (int f = 0; f < NF; ++f) acc[f] = (acc[f] + cur[f]) / 2.f;
                break;
            }
            case SSAR_NOISE_MODULATE: {
#pragma unroll
                for (int f = 0; f < NF; ++f) {
                    if (f < nf) {
                        const float m = __ldg(op.mod + i0 + f0 + f);
                        acc[f] = acc[f] * m + cur[f] * (1.f - m);
                    }
                }
                break;
            }
            case SSAR_NOISE_SCALEBIAS: {
#pragma unroll
                for (int f = 0; f < NF; ++f) acc[f] = op.p0 * acc[f] + op.p1;
                break;
            }
            default:
                break;
        }
        if (op.to_acc) {
#pragma unroll
            for (int f = 0; f < NF; ++f) acc[f] = cur[f];
        }
    }
#pragma unroll
    for (int f = 0; f < NF; ++f)
        if (f < nf) out[(long long)(f0 + f) * HW + p] = acc[f];
}

}  // namespace

extern "C" int ssar_noise_eval(const ssar_noise_op* ops_device, const ssar_noise_op* ops_host, int n_ops, int first_frame,
                               int n_frames, int height, int width, float* loop_scratch, float* out, void* stream) {
    if (n_frames == 0 || height * width == 0) return SSAR_OK;
    SSAR_REQUIRE(ops_device && ops_host && out && n_ops > 0, "noise_eval: null pointer or empty program");
    const int HW = height * width;
    int n_loops = 0;
    for (int o = 0; o < n_ops; ++o) {
        const ssar_noise_op& op = ops_host[o];
        SSAR_REQUIRE(op.kind >= SSAR_NOISE_LOOP && op.kind <= SSAR_NOISE_SCALEBIAS, "noise_eval: bad op kind %d", op.kind);
        if (op.kind == SSAR_NOISE_LOOP) ++n_loops;
        if (op.kind <= SSAR_NOISE_BUFFER) SSAR_REQUIRE(op.noise != nullptr, "noise_eval: leaf %d without data", o);
        if (op.kind == SSAR_NOISE_LOOP || op.kind == SSAR_NOISE_BLEND || op.kind == SSAR_NOISE_MULTIPLY ||
            op.kind == SSAR_NOISE_MODULATE)
            SSAR_REQUIRE(op.mod != nullptr, "noise_eval: op %d without modulator / index", o);
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n_loops > 0) {
        SSAR_REQUIRE(loop_scratch != nullptr, "noise_eval: loop_scratch (n_loops * n_frames floats) is required");
        cudaError_t e = cudaMemsetAsync(loop_scratch, 0, sizeof(float) * n_loops * n_frames, st);
        if (e != cudaSuccess) {
            ssar_set_error("noise_eval: memset: %s", cudaGetErrorString(e));
            return SSAR_ERR_CUDA;
        }
        int gz = ssar_div_up(HW, 256 * 8);
        if (gz > 64) gz = 64;
        if (gz < 1) gz = 1;
        noise_loop_rms_kernel<<<dim3(n_frames, n_loops, gz), 256, 0, st>>>(ops_device, n_ops, first_frame, n_frames, HW,
                                                                            loop_scratch);
    }
    dim3 grid(ssar_div_up(HW, 256), ssar_div_up(n_frames, NF));
    noise_eval_kernel<<<grid, 256, 0, st>>>(ops_device, n_ops, first_frame, n_frames, HW, loop_scratch, out);
    return ssar_launch_status("noise_eval");
}
