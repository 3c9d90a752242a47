// make_noise_pyramid (maua/GAN/wrappers/stylegan2.py:193-210) and the synthesizer's resize of mismatched noise
// (:86-96): one noise sequence [T, 1, H, W] -> one [T, 1, h_l, w_l] tensor per synthesis layer, bicubic
// (align_corners = False, A = -0.75, border indices clamped: ATen's upsample_bicubic2d), every frame divided by its own
// unbiased standard deviation.  ONE launch for all layers: a CTA owns one (frame, layer) pair, stages the input frame in
// shared memory, writes the resized frame, reduces mean and centred sum of squares (two passes over its own writes,
// which sit in L1/L2) and rescales in place.  HBM-bound: 4 B read per input pixel per layer (L2 after the first layer)
// + 4 B written per output pixel.
#include "block_utils.cuh"
#include "common.cuh"

namespace {

constexpr int PYR_MAX_LAYERS = 16;

struct PyramidLayers {
    int n;
    int h[PYR_MAX_LAYERS], w[PYR_MAX_LAYERS];
    float* out[PYR_MAX_LAYERS];
};

__device__ __forceinline__ float cubic1(float x) { return ((-0.75f + 2.f) * x - (-0.75f + 3.f)) * x * x + 1.f; }
__device__ __forceinline__ float cubic2(float x) { return ((-0.75f * x - 5.f * -0.75f) * x + 8.f * -0.75f) * x - 4.f * -0.75f; }

__device__ __forceinline__ void cubic_coeffs(float t, float (&c)[4]) {
    c[0] = cubic2(t + 1.f);
    c[1] = cubic1(t);
    const float u = 1.f - t;
    c[2] = cubic1(u);
    c[3] = cubic2(u + 1.f);
}

__global__ void __launch_bounds__(256) noise_pyramid_kernel(const float* __restrict__ noise, int in_h, int in_w,
                                                            const PyramidLayers layers, int normalize) {
    extern __shared__ float tile[];  // the input frame
    __shared__ float scratch[32];
    __shared__ float s_bcast;
    const int frame = blockIdx.x, l = blockIdx.y;
    const int h = layers.h[l], w = layers.w[l];
    const float* src = noise + (long long)frame * in_h * in_w;
    for (int i = threadIdx.x; i < in_h * in_w; i += blockDim.x) tile[i] = __ldg(src + i);
    __syncthreads();
    float* dst = layers.out[l] + (long long)frame * h * w;
    const float sy = (float)in_h / (float)h, sx = (float)in_w / (float)w;  // area_pixel_compute_scale, align_corners=False
    const int n = h * w;
    float sum = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int oy = i / w, ox = i - oy * w;
        const float ry = sy * ((float)oy + 0.5f) - 0.5f, rx = sx * ((float)ox + 0.5f) - 0.5f;
        const float fy = floorf(ry), fx = floorf(rx);
        const int iy = (int)fy, ix = (int)fx;
        float cy[4], cx[4];
        cubic_coeffs(ry - fy, cy);
        cubic_coeffs(rx - fx, cx);
        float v = 0.f;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const float* row = tile + min(max(iy - 1 + a, 0), in_h - 1) * in_w;
            float r = 0.f;
#pragma unroll
            for (int b = 0; b < 4; ++b) r += row[min(max(ix - 1 + b, 0), in_w - 1)] * cx[b];
            v += r * cy[a];
        }
        dst[i] = v;
        sum += v;
    }
    if (!normalize) return;
    sum = block_sum(sum, scratch);
    if (threadIdx.x == 0) s_bcast = sum / (float)n;
    __syncthreads();
    const float mean = s_bcast;
    float ss = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {  // each thread re-reads exactly what it wrote
        const float d = dst[i] - mean;
        ss = fmaf(d, d, ss);
    }
    __syncthreads();
    ss = block_sum(ss, scratch);
    if (threadIdx.x == 0) s_bcast = sqrtf(ss / (float)(n - 1));  // torch.std: unbiased
    __syncthreads();
    const float sd = s_bcast;
    for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = dst[i] / sd;
}

}  // namespace

extern "C" int ssar_noise_pyramid(const float* noise, int n_frames, int in_h, int in_w, int n_layers, const int* heights,
                                  const int* widths, float* const* outs, int normalize, void* stream) {
    if (n_frames == 0 || n_layers == 0) return SSAR_OK;
    SSAR_REQUIRE(noise && heights && widths && outs, "noise_pyramid: null pointer");
    SSAR_REQUIRE(n_layers > 0 && n_layers <= PYR_MAX_LAYERS, "noise_pyramid: at most %d layers per launch", PYR_MAX_LAYERS);
    SSAR_REQUIRE(in_h > 0 && in_w > 0 && (size_t)in_h * in_w * sizeof(float) <= 200 * 1024,
                 "noise_pyramid: the input frame must fit in shared memory (<= 51 200 pixels)");
    PyramidLayers L;
    L.n = n_layers;
    for (int l = 0; l < n_layers; ++l) {
        SSAR_REQUIRE(heights[l] > 0 && widths[l] > 0 && outs[l], "noise_pyramid: bad layer %d", l);
        SSAR_REQUIRE(!normalize || heights[l] * widths[l] > 1, "noise_pyramid: a 1-pixel layer has no standard deviation");
        L.h[l] = heights[l];
        L.w[l] = widths[l];
        L.out[l] = outs[l];
    }
    const size_t smem = (size_t)in_h * in_w * sizeof(float);
    if (smem > 48 * 1024) cudaFuncSetAttribute(noise_pyramid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    noise_pyramid_kernel<<<dim3(n_frames, n_layers), 256, smem, (cudaStream_t)stream>>>(noise, in_h, in_w, L, normalize);
    return ssar_launch_status("noise_pyramid");
}
