// Polyphase windowed-sinc resampler (banded taps).
// Replaces torchaudio.functional.resample at ssar/supervised/data.py:63 (44.1 kHz -> fps*1024, sinc_interp_hann)
// and rosa/constantq.py:83 (octave decimation by 2, kaiser_window).  torchaudio builds a dense
// [new, 1, 2*width+orig] kernel and runs conv1d with stride orig; only ~2*width taps per phase are non-zero, so
// the plan keeps those (tap-major [max_taps, new] for coalesced reads) plus the first tap index of every phase.
#include "common.cuh"

namespace {

__global__ void resample_kernel(const float* __restrict__ x, long long n_in, long long stride_in,
                                const float* __restrict__ taps, const int* __restrict__ first, int orig, int newf,
                                int width, int max_taps, float post_scale, float* __restrict__ y, long long n_out,
                                long long stride_out) {
    const int clip = blockIdx.y;
    const float* xc = x + (long long)clip * stride_in;
    float* yc = y + (long long)clip * stride_out;
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
        yc[j] = acc * post_scale;
    }
}

}  // namespace

extern "C" int ssar_resample(const float* x, long long n_in, long long stride_in, int n_clips, const float* taps,
                             const int* first, int orig, int newf, int width, int max_taps, float post_scale, float* y,
                             long long n_out, long long stride_out, void* stream) {
    if (n_clips == 0 || n_out == 0) return SSAR_OK;
    SSAR_REQUIRE(x && taps && first && y, "resample: null pointer");
    SSAR_REQUIRE(orig > 0 && newf > 0 && max_taps > 0, "resample: bad rates");
    int gx = ssar_div_up(n_out, 256);
    if (gx > SSAR_NUM_SMS * 16) gx = SSAR_NUM_SMS * 16;
    resample_kernel<<<dim3(gx, n_clips), 256, 0, (cudaStream_t)stream>>>(x, n_in, stride_in, taps, first, orig, newf, width,
                                                                        max_taps, post_scale, y, n_out, stride_out);
    return ssar_launch_status("resample");
}
