// Time-domain kernels: framed RMS.
#include "block_utils.cuh"

namespace {

// one warp per frame; frame t covers reflect-padded samples [t*hop - L/2, t*hop + L/2)
__global__ void rms_frames_kernel(const float* __restrict__ x, long long n_samples, long long clip_stride, int n_clips,
                                  int frame_length, int hop, int n_frames, float* __restrict__ out) {
    const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, l = threadIdx.x & 31;
    const long long total = (long long)n_clips * n_frames;
    for (long long f = (long long)blockIdx.x * warps + w; f < total; f += (long long)gridDim.x * warps) {
        const int clip = (int)(f / n_frames);
        const int t = (int)(f - (long long)clip * n_frames);
        const float* xc = x + (long long)clip * clip_stride;
        const long long start = (long long)t * hop - frame_length / 2;
        const bool interior = start >= 0 && start + frame_length <= n_samples;
        float acc = 0.f;
        if (interior) {
            for (int i = l; i < frame_length; i += 32) {
                const float v = __ldg(xc + start + i);
                acc = fmaf(v, v, acc);
            }
        } else {
            for (int i = l; i < frame_length; i += 32) {
                const float v = __ldg(xc + ssar_reflect(start + i, n_samples));
                acc = fmaf(v, v, acc);
            }
        }
        acc = warp_sum(acc);
        if (l == 0) out[f] = sqrtf(acc / (float)frame_length);
    }
}

// 16-bit PCM -> fp32 in [-1, 1): x / 32768, exact in fp32 (what torchaudio.load's normalisation does to the PCM files the
// reference reads, supervised/data.py:127, :180): lets a host ship the audio in its on-disk width, half the H2D bytes.
__global__ void pcm16_to_float_kernel(const short* __restrict__ in, long long n, float* __restrict__ out) {
    const long long n8 = n >> 3;
    if (((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
            const int4 v = __ldg(reinterpret_cast<const int4*>(in) + i);
            const short* s = reinterpret_cast<const short*>(&v);
            float4 a, b;
            a.x = (float)s[0] * (1.f / 32768.f); a.y = (float)s[1] * (1.f / 32768.f);
            a.z = (float)s[2] * (1.f / 32768.f); a.w = (float)s[3] * (1.f / 32768.f);
            b.x = (float)s[4] * (1.f / 32768.f); b.y = (float)s[5] * (1.f / 32768.f);
            b.z = (float)s[6] * (1.f / 32768.f); b.w = (float)s[7] * (1.f / 32768.f);
            reinterpret_cast<float4*>(out)[2 * i] = a;
            reinterpret_cast<float4*>(out)[2 * i + 1] = b;
        }
        for (long long i = 8 * n8 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
            out[i] = (float)__ldg(in + i) * (1.f / 32768.f);
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
            out[i] = (float)__ldg(in + i) * (1.f / 32768.f);
    }
}

}  // namespace

extern "C" int ssar_pcm16_to_float(const short* pcm, long long n, float* out, void* stream) {
    if (n == 0) return SSAR_OK;
    SSAR_REQUIRE(pcm && out, "pcm16_to_float: null pointer");
    int grid = ssar_div_up(n, 256 * 8);
    if (grid > SSAR_NUM_SMS * 16) grid = SSAR_NUM_SMS * 16;
    pcm16_to_float_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pcm, n, out);
    return ssar_launch_status("pcm16_to_float");
}

extern "C" int ssar_rms_frames(const float* x, long long n_samples, long long clip_stride, int n_clips,
                               int frame_length, int hop, int n_frames, float* out, void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(x && out, "rms_frames: null pointer");
    SSAR_REQUIRE(frame_length > 0 && hop > 0 && n_samples > frame_length / 2, "rms_frames: bad sizes");
    SSAR_REQUIRE((long long)(n_frames - 1) * hop <= n_samples, "rms_frames: n_frames exceeds 1 + n_samples/hop");
    const long long total = (long long)n_clips * n_frames;
    int grid = ssar_div_up(total, 8);
    if (grid > SSAR_NUM_SMS * 8) grid = SSAR_NUM_SMS * 8;
    rms_frames_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, n_samples, clip_stride, n_clips, frame_length, hop,
                                                             n_frames, out);
    return ssar_launch_status("rms_frames");
}
