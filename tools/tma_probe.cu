// Probe (run on the GPU box): semantics of 4-D tiled TMA box loads with negative start coordinates, out-of-bound zero
// fill and elementStrides = 2 — the primitives the implicit-GEMM transposed convolution relies on instead of an
// im2col buffer.   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/tma_probe tools/tma_probe.cu -lcuda
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void probe(const __grid_constant__ CUtensorMap tm, float* out, int nfloats, int c0, int c1, int c2, int c3) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    for (int i = threadIdx.x; i < nfloats; i += blockDim.x) ((float*)smem)[i] = -777.f;
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(nfloats * 4));
        asm volatile(
            "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                smem_u32(smem)),
            "l"(&tm), "r"(smem_u32(&bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
            : "memory");
    }
    asm volatile(
        "{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}\n" ::"r"(
            smem_u32(&bar)));
    __syncthreads();
    for (int i = threadIdx.x; i < nfloats; i += blockDim.x) out[i] = ((float*)smem)[i];
}

int main() {
    const int N = 3, H = 8, W = 8, C = 32;
    std::vector<float> h((size_t)N * H * W * C);
    for (int n = 0; n < N; ++n)
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x)
                for (int c = 0; c < C; ++c) h[(((size_t)n * H + y) * W + x) * C + c] = 1000.f * n + 100.f * y + 10.f * x + c * 0.01f + 1.f;
    float* d;
    cudaMalloc(&d, h.size() * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    auto enc = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    int fails = 0;
    for (int test = 0; test < 4; ++test) {
        const int stride = (test & 1) ? 2 : 1;
        const bool swz = test >= 2;
        const int bw = 4, bh = 4, bn = 2;                           // pixels loaded per dim
        cuuint64_t dims[4] = {C, W, H, N};
        cuuint64_t strides[3] = {C * 4ull, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
        cuuint32_t box[4] = {32, (cuuint32_t)(bw * stride), (cuuint32_t)(bh * stride), (cuuint32_t)bn};
        cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
        CUtensorMap tm;
        CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("test %d: encode failed %d\n", test, (int)r); ++fails; continue; }
        const int nfl = 32 * bw * bh * bn;
        float* o;
        cudaMalloc(&o, nfl * 4);
        const int c1 = -1, c2 = (stride == 2) ? -1 : 6, c3 = 1;      // x from -1; y from -1 (strided) or 6 (runs off the bottom)
        probe<<<1, 128, nfl * 4 + 1024>>>(tm, o, nfl, 0, c1, c2, c3);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("test %d: kernel failed: %s\n", test, cudaGetErrorString(e)); return 1; }
        std::vector<float> got(nfl);
        cudaMemcpy(got.data(), o, nfl * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int b = 0; b < bn; ++b)
            for (int i = 0; i < bh; ++i)
                for (int j = 0; j < bw; ++j)
                    for (int c = 0; c < 32; ++c) {
                        const int n = c3 + b, y = c2 + i * stride, x = c1 + j * stride;
                        const bool in = n >= 0 && n < N && y >= 0 && y < H && x >= 0 && x < W;
                        const float want = in ? h[(((size_t)n * H + y) * W + x) * C + c] : 0.f;
                        const int row = (b * bh + i) * bw + j;
                        int chunk = c / 4;
                        if (swz) chunk ^= (row & 7);
                        const float v = got[row * 32 + chunk * 4 + (c & 3)];
                        if (v != want) {
                            if (bad < 5) printf("  test %d mismatch b=%d i=%d j=%d c=%d got %g want %g\n", test, b, i, j, c, v, want);
                            ++bad;
                        }
                    }
        printf("test %d (elementStride %d, swizzle128 %d): %s (%d mismatches of %d)\n", test, stride, (int)swz, bad ? "FAIL" : "ok", bad, nfl);
        fails += bad != 0;
        cudaFree(o);
    }
    printf(fails ? "PROBE FAILED\n" : "PROBE OK\n");
    return fails != 0;
}
