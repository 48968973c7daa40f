// Microbenchmark (run on the GPU box): cycles per tcgen05.mma kind::tf32, M=128, K=8, as a function of N, operand source
// (A from shared memory vs A from TMEM) and of operand reuse between consecutive instructions.  Operands are whatever is
// in shared memory; only the timing means anything.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/umma_rate tools/umma_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../mmvae_b200/csrc/sm100_ptx.cuh"

__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(acc)
        : "memory");
}

// mode 0: SS, distinct A/B addresses per MMA (4 kk slices of 4 stages);  mode 1: SS, same A every time; mode 2: same B
// every time; mode 3: TS (A from TMEM);  mode 4: SS with the 3xTF32 operand pattern (al,bh)(ah,bl)(ah,bh)
__global__ void __launch_bounds__(128, 1) rate(int N, int mode, int reps, long long* out, int R) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) ((float*)smem)[i] = 1.0f + (i & 7) * 0.125f;
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = umma_idesc_tf32(128, N, 0, 0);
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;       // A tiles: 4 x 16 KB; B tiles: 4 x up to 32 KB
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const int st = r & 3;
            for (int kk = 0; kk < 4; ++kk) {
                const uint32_t aa = a0 + (mode == 1 ? 0 : st * 16384 + kk * 32), bb = b0 + (mode == 2 ? 0 : st * 32768 + kk * 32);
                const uint64_t da = umma_smem_desc(aa, 16, 1024, 2), db = umma_smem_desc(bb, 16, 1024, 2);
                if (mode == 3) umma_tf32_ts(tm, tm + 256 + kk * 8, db, idesc, 1);
                else if (mode == 4) {
                    const uint64_t dal = umma_smem_desc(aa + 8192 * 0 + 65536 * 0 + 16384 * 0 + 4096, 16, 1024, 2);   // another address
                    const uint64_t dbl = umma_smem_desc(bb + 4096, 16, 1024, 2);
                    umma_tf32(tm + 256, dal, db, idesc, 1);
                    umma_tf32(tm + 256, da, dbl, idesc, 1);
                    umma_tf32(tm, da, db, idesc, 1);
                } else if (mode == 5) umma_tf32(tm + ((r * 4 + kk) % R) * N, da, db, idesc, 1);       // R independent accumulators
                else if (mode == 6) umma_tf32(tm + ((r * 4 + kk) % R) * N, umma_smem_desc(a0, 16, 1024, 2), db, idesc, 1);   // + same A
                else umma_tf32(tm, da, db, idesc, 1);
            }
        }
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0) out[0] = t1 - t0;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}

// warp-uniform issue: all 32 lanes run the loop, one elected lane issues; descriptors advance by adds on the low word
__device__ __forceinline__ void umma_tf32_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc)
        : "memory");
}
template <int R>
__global__ void __launch_bounds__(128, 1) rate_u(int N, int reps, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) ((float*)smem)[i] = 1.0f + (i & 7) * 0.125f;
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = slot;
    if (warp == 0) {
        const uint32_t idesc = umma_idesc_tf32(128, N, 0, 0);
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
        const uint64_t da0 = umma_smem_desc(a0, 16, 1024, 2), db0 = umma_smem_desc(b0, 16, 1024, 2);
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const uint64_t da = da0 + (uint64_t)((r & 3) * (16384 >> 4)), db = db0 + (uint64_t)((r & 3) * (32768 >> 4));
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) umma_tf32_elect(tm + (kk % R) * N, da + kk * 2, db + kk * 2, idesc, 1);
        }
        if (threadIdx.x == 0) { umma_commit(&bar); }
        mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}
template <int R> static void run_u(long long* d, int reps) {
    cudaFuncSetAttribute(rate_u<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int N : {32, 64, 128, 256}) {
        if (R * N > 512) continue;
        rate_u<R><<<148, 128, 200 * 1024>>>(N, reps, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("rate_u R %d N %d: %s\n", R, N, cudaGetErrorString(e)); exit(1); }
        long long c;
        cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        printf("uniform issue, N=%3d, %d accumulators round-robin: %7.1f clk/MMA  (floor %d)\n", N, R, (double)c / (reps * 4), N / 2);
    }
}

int main() {
    long long* d;
    cudaMalloc(&d, 8);
    cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int reps = 2000;
    run_u<1>(d, reps); run_u<2>(d, reps); run_u<4>(d, reps);
    const char* names[] = {"SS distinct", "SS same A", "SS same B", "TS (A in TMEM)", "SS 3xTF32 pattern"};
    for (int grid : {148})
        for (int mode = 0; mode < 5; ++mode)
            for (int N : {32, 64, 128, 256}) {
                if (mode == 4 && N == 256) continue;
                rate<<<grid, 128, 200 * 1024>>>(N, mode, reps, d, grid);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("mode %d N %d: %s\n", mode, N, cudaGetErrorString(e)); return 1; }
                long long c;
                cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
                const int n = reps * 4 * (mode == 4 ? 3 : 1);
                printf("grid %3d  %-18s N=%3d: %7.1f clk/MMA  (floor 128*N/256 = %d)\n", grid, names[mode], N, (double)c / n, N / 2);
            }
    for (int mode = 5; mode < 7; ++mode)
        for (int N : {32, 64, 128, 256})
            for (int R : {1, 2, 3, 4, 8, 16}) {
                if (R * N > 512) continue;
                rate<<<148, 128, 200 * 1024>>>(N, mode, reps, d, R);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("mode %d N %d: %s\n", mode, N, cudaGetErrorString(e)); return 1; }
                long long c;
                cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
                printf("%s N=%3d, %2d independent accumulators: %7.1f clk/MMA  (floor %d)\n", mode == 5 ? "SS distinct" : "SS same A  ", N, R, (double)c / (reps * 4), N / 2);
            }
    return 0;
}
