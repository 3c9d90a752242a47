// Probe: does cuTensorMapEncodeTiled accept a row stride SMALLER than the row extent (overlapping rows), and does the
// TMA unit deliver the expected elements?  View of a contiguous fp16 signal y[n] as frames[t][k] = y[t * hop + k].
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const __grid_constant__ CUtensorMap map, int c0, int c1, int c2, __half* out) {
    __shared__ __align__(1024) __half tile[128 * 64];
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(128 * 64 * 2) : "memory");
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
                smem_u32(tile)),
            "l"(reinterpret_cast<uint64_t>(&map)), "r"(smem_u32(&bar)), "r"(c0), "r"(c1), "r"(c2)
            : "memory");
    }
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n .reg .pred P1;\n mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n selp.b32 %0, 1, 0, P1;\n}"
                     : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    }
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) out[i] = tile[i];
}

int main() {
    EncodeTiledFn enc = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q);
    const int B = 3, Lp = 1024 * 40;
    std::vector<__half> h((size_t)B * Lp);
    for (size_t i = 0; i < h.size(); ++i) h[i] = __float2half((float)(i % 2039));
    __half *d, *o;
    cudaMalloc(&d, h.size() * 2);
    cudaMalloc(&o, 128 * 64 * 2);
    cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
    for (int hop : {1024, 256, 64, 32, 16, 8}) {
        for (int swz = 0; swz < 2; ++swz) {
            const int T = (Lp - 1024) / hop + 1;
            CUtensorMap map;
            const cuuint64_t dims[3] = {1024, (cuuint64_t)T, (cuuint64_t)B};
            const cuuint64_t strides[2] = {(cuuint64_t)hop * 2, (cuuint64_t)Lp * 2};
            const cuuint32_t box[3] = {64, 128, 1};
            const cuuint32_t estr[3] = {1, 1, 1};
            CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) {
                printf("hop %4d swizzle %d: encode FAILED (%d)\n", hop, swz, (int)r);
                continue;
            }
            const int k0 = 192, t0 = 5, b = 1;
            probe<<<1, 128>>>(map, k0, t0, b, o);
            cudaError_t e = cudaDeviceSynchronize();
            std::vector<__half> got(128 * 64);
            cudaMemcpy(got.data(), o, got.size() * 2, cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int row = 0; row < 128; ++row)
                for (int c = 0; c < 64; ++c) {
                    // SWIZZLE_128B: 16-byte chunk index XOR (row % 8)
                    const int chunk = c / 8, cc = swz ? ((chunk ^ (row & 7)) * 8 + c % 8) : c;
                    const float want = __half2float(h[(size_t)b * Lp + (size_t)(t0 + row) * hop + k0 + c]);
                    if (__half2float(got[row * 64 + cc]) != want) ++bad;
                }
            printf("hop %4d swizzle %d: encode ok, run %s, mismatches %d\n", hop, swz, cudaGetErrorString(e), bad);
        }
    }
    return 0;
}
