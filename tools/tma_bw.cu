// Probe (GPU box): sustained TMA load throughput per SM for 16 KB tiles of 128 rows x 128 B fetched as (a) a 2-D box,
// (b) a 4-D box over an NHWC tensor (shifted window), (c) a 4-D box with elementStrides 2 (stride-2 gather).
// 148 CTAs, one issuing thread each, 8 tiles in flight, no compute.   nvcc -arch=sm_100a -o tools/tma_bw tools/tma_bw.cu -lcuda
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t su32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
#define NBUF 8

__global__ void __launch_bounds__(64, 1) bw_kernel(const __grid_constant__ CUtensorMap tm, int mode, int iters, int nimg, int H, int W) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar[NBUF];
    if (threadIdx.x == 0) {
        for (int i = 0; i < NBUF; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(su32(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const int imgs_per_tile = 128 / (H * W);
    for (int it = 0; it < iters + NBUF; ++it) {
        const int b = it % NBUF;
        if (it >= NBUF) {      // wait for the load issued NBUF iterations ago into this buffer
            const int ph = ((it / NBUF) - 1) & 1;
            asm volatile("{\n\t.reg .pred p;\n\tW%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D%=;\n\tbra W%=;\n\tD%=:\n\t}\n" ::"r"(su32(&bar[b])), "r"(ph));
        }
        if (it < iters) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(su32(&bar[b])), "r"(16384));
            const int tile = (blockIdx.x * 977 + it * 13) % (nimg / imgs_per_tile);      // scattered tiles, L2-sized working set
            if (mode == 0) {
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(su32(smem + b * 16384)), "l"(&tm), "r"(su32(&bar[b])), "r"(0), "r"(tile * 128) : "memory");
            } else {
                const int sh = it % 3 - 1;
                asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(su32(smem + b * 16384)), "l"(&tm), "r"(su32(&bar[b])), "r"(0), "r"(sh), "r"(-sh), "r"(tile * imgs_per_tile) : "memory");
            }
        }
    }
}

int main() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    auto enc = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    const int H = 8, W = 8, C = 32;
    for (int big = 0; big < 2; ++big) {
        const int nimg = big ? 7680 * 4 : 2048;                 // 252 MB (HBM-resident) or 16 MB (L2-resident) coarse tensor
        float *d, *d2;
        cudaMalloc(&d, (size_t)nimg * H * W * C * 4);
        cudaMalloc(&d2, (size_t)nimg * 4 * H * W * C * 4);
        cudaMemset(d, 0, (size_t)nimg * H * W * C * 4);
        cudaMemset(d2, 0, (size_t)nimg * 4 * H * W * C * 4);
        for (int mode = 0; mode < 3; ++mode) {
            CUtensorMap tm;
            CUresult r;
            if (mode == 0) {
                cuuint64_t dims[2] = {C, (cuuint64_t)nimg * H * W};
                cuuint64_t st[1] = {C * 4};
                cuuint32_t box[2] = {32, 128}, es[2] = {1, 1};
                r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            } else {
                const int s = mode == 2 ? 2 : 1;
                cuuint64_t dims[4] = {C, (cuuint64_t)W * s, (cuuint64_t)H * s, (cuuint64_t)nimg};
                cuuint64_t st[3] = {C * 4, (cuuint64_t)W * s * C * 4, (cuuint64_t)H * s * W * s * C * 4};
                cuuint32_t box[4] = {32, (cuuint32_t)(W * s), (cuuint32_t)(H * s), 2}, es[4] = {1, (cuuint32_t)s, (cuuint32_t)s, 1};
                r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, s == 2 ? d2 : d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            }
            if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
            cudaFuncSetAttribute(bw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, NBUF * 16384 + 1024);
            const int iters = 4000;
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            bw_kernel<<<148, 64, NBUF * 16384 + 1024>>>(tm, mode, 200, nimg, H, W);
            cudaEventRecord(e0);
            bw_kernel<<<148, 64, NBUF * 16384 + 1024>>>(tm, mode, iters, nimg, H, W);
            cudaEventRecord(e1);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return 1; }
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            const double bytes = 148.0 * iters * 16384;
            printf("%s tensor, %s: %.1f us, %.2f TB/s aggregate, %.1f B/clk/SM (1.965 GHz), %.0f clk per 16 KB tile\n", big ? "252 MB" : "16 MB",
                   mode == 0 ? "2-D box          " : mode == 1 ? "4-D shifted box  " : "4-D stride-2 box ", ms * 1e3, bytes / ms / 1e9,
                   bytes / 148 / (ms * 1e-3 * 1.965e9), ms * 1e-3 * 1.965e9 / iters);
        }
        cudaFree(d); cudaFree(d2);
    }
    return 0;
}
