// Probe (run on the GPU box): tcgen05.mma with the A operand in TENSOR MEMORY ("TS" form), kind::f16 (bf16) and kind::tf32.
//   1. numerics: A[128,K] written to TMEM by tcgen05.st.32x32b, B[N,K] in shared memory (K-major, SWIZZLE_128B, written by
//      generic stores), D = A*B^T read back and compared with the host -> settles the TMEM A layout (lane = row, 32-bit
//      column = 1 tf32 / 2 packed bf16, low half = lower k) and the bf16 instruction descriptor;
//   2. timing: clocks per MMA for SS and TS forms, bf16 (K=16) and tf32 (K=8), N = 32..256, warp-uniform issue.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ts_probe tools/ts_probe.cu
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include "../mmvae_b200/csrc/sm100_ptx.cuh"

__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
template <bool bf16> __device__ __forceinline__ void umma_ts_elect(uint32_t d, uint32_t a_tmem, uint64_t db, uint32_t idesc, uint32_t acc) {
    if (bf16)
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
                     "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
                     "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d), "r"(a_tmem), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
template <bool bf16> __device__ __forceinline__ void umma_ss_elect(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    if (bf16)
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
                     "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\telect.sync _|q, 0xffffffff;\n\t"
                     "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st_x16(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
                 "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}

// ---- 1. numerics.  K = 32 (bf16: 2 MMAs of K=16, 16 TMEM columns; tf32: 4 MMAs of K=8, 32 columns), N = 64.
template <bool BF16>
__global__ void __launch_bounds__(128, 1) numerics(const float* A, const float* B, float* D) {
    constexpr int N = 64, K = 32;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = threadIdx.x;
    if (t == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(&slot, 128); tmem_relinquish(); }
    // B[n][k] K-major, 128-byte rows, SWIZZLE_128B: 16-byte chunk c of row n lives at chunk c ^ (n & 7)
    for (int i = t; i < N * K; i += 128) {
        const int n = i / K, k = i % K;
        if (BF16) {
            const int byte = k * 2, c = byte >> 4;
            *(__nv_bfloat16*)(smem + n * 128 + ((c ^ (n & 7)) << 4) + (byte & 15)) = __float2bfloat16(B[i]);
        } else {
            const int byte = k * 4, c = byte >> 4;
            *(float*)(smem + n * 128 + ((c ^ (n & 7)) << 4) + (byte & 15)) = B[i];
        }
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = slot;
    // A row t -> TMEM lane t, columns 64.. : bf16 packs (k even -> low half)
    {
        uint32_t r[32];
        if (BF16) {
            for (int j = 0; j < 16; ++j) {
                const __nv_bfloat16 lo = __float2bfloat16(A[t * K + 2 * j]), hi = __float2bfloat16(A[t * K + 2 * j + 1]);
                r[j] = (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
            }
            tmem_st_x16(tm + ((uint32_t)(warp * 32) << 16) + 64, r);
        } else {
            for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(A[t * K + j]);
            tmem_st_x16(tm + ((uint32_t)(warp * 32) << 16) + 64, r);
            tmem_st_x16(tm + ((uint32_t)(warp * 32) << 16) + 64 + 16, r + 16);
        }
        tmem_st_wait();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    if (warp == 0) {
        const uint64_t db0 = umma_smem_desc(smem_u32(smem), 16, 1024, 2);
        if (BF16) {
            for (int kk = 0; kk < 2; ++kk) umma_ts_elect<true>(tm, tm + 64 + kk * 8, db0 + kk * 2, idesc_bf16(128, N), kk != 0);
        } else {
            for (int kk = 0; kk < 4; ++kk) umma_ts_elect<false>(tm, tm + 64 + kk * 8, db0 + kk * 2, umma_idesc_tf32(128, N, 0, 0), kk != 0);
        }
        umma_commit_elect(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after_sync();
    float v[32];
    for (int c = 0; c < N; c += 32) {
        tmem_ld_32x32(tm + ((uint32_t)(warp * 32) << 16) + c, v);
        for (int j = 0; j < 32; ++j) D[t * N + c + j] = v[j];
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 128);
}

// ---- 2. timing.  mode 0 = SS, 1 = TS; R accumulators round-robin (independent chains)
template <bool BF16, int mode, int R>
__global__ void __launch_bounds__(128, 1) rate(int N, int reps, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(&slot, 512); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3f803f80u;
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = slot;
    if (warp == 0) {
        const uint32_t idesc = BF16 ? idesc_bf16(128, N) : umma_idesc_tf32(128, N, 0, 0);
        const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
        const uint64_t da0 = umma_smem_desc(a0, 16, 1024, 2), db0 = umma_smem_desc(b0, 16, 1024, 2);
        long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            const uint64_t da = da0 + (uint64_t)((r & 3) * (16384 >> 4)), db = db0 + (uint64_t)((r & 3) * (32768 >> 4));
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                if (mode == 1) umma_ts_elect<BF16>(tm + (kk % R) * N, tm + 256 + (r & 3) * 32 + kk * 8, db + kk * 2, idesc, 1);
                else umma_ss_elect<BF16>(tm + (kk % R) * N, da + kk * 2, db + kk * 2, idesc, 1);
            }
        }
        umma_commit_elect(&bar);
        mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 512);
}

template <bool BF16> static int check(const char* name) {
    const int N = 64, K = 32;
    float *hA = (float*)malloc(128 * K * 4), *hB = (float*)malloc(N * K * 4), *hD = (float*)malloc(128 * N * 4);
    srand(1);
    for (int i = 0; i < 128 * K; ++i) hA[i] = (float)((rand() % 17) - 8) / 8.f;       // exact in bf16
    for (int i = 0; i < N * K; ++i) hB[i] = (float)((rand() % 17) - 8) / 4.f;
    float *dA, *dB, *dD;
    cudaMalloc(&dA, 128 * K * 4); cudaMalloc(&dB, N * K * 4); cudaMalloc(&dD, 128 * N * 4);
    cudaMemcpy(dA, hA, 128 * K * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, N * K * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(numerics<BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * 1024);
    numerics<BF16><<<1, 128, 32 * 1024>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s numerics: %s\n", name, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(hD, dD, 128 * N * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)hA[m * K + k] * hB[n * K + k];
            maxerr = fmax(maxerr, fabs(s - hD[m * N + n]));
        }
    printf("%s TS numerics (A in TMEM via tcgen05.st.32x32b, low half = even k): max |err| = %g  %s\n", name, maxerr,
           maxerr == 0 ? "EXACT" : "MISMATCH");
    return maxerr != 0;
}

template <bool BF16, int mode, int R> static void timing(long long* d, const char* name) {
    const int reps = 2000;
    cudaFuncSetAttribute(rate<BF16, mode, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    {
        for (int N : {32, 64, 96, 128, 192, 256}) {
            if (R * N > 256) continue;
            rate<BF16, mode, R><<<148, 128, 200 * 1024>>>(N, reps, d);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s mode %d N %d: %s\n", name, mode, N, cudaGetErrorString(e)); exit(1); }
            long long c;
            cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
            printf("%s %s R=%d N=%3d: %7.1f clk/MMA (K=%d per MMA; N/2 = %d)\n", name, mode ? "TS" : "SS", R, N, (double)c / (reps * 4),
                   BF16 ? 16 : 8, N / 2);
        }
    }
}


// ---- 3. B operand MN-major (rows = k, 64 bf16 of n contiguous = 128 bytes, SWIZZLE_128B), A in TMEM.  K = 16 per MMA, two
// MMAs (32 k rows).  n_off = first n column (0 or 32: a descriptor that starts 64 bytes into the swizzled row), N = 32 or 64.
__global__ void __launch_bounds__(128, 1) numerics_mn(const float* A, const float* B, float* D, int N, int n_off, int sbo, int lbo) {
    constexpr int K = 32;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, t = threadIdx.x;
    if (t == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) { tmem_alloc(&slot, 128); tmem_relinquish(); }
    for (int i = t; i < K * 64; i += 128) {                      // B[k][n], n < 64
        const int k = i / 64, n = i % 64, byte = n * 2, c = byte >> 4;
        *(__nv_bfloat16*)(smem + k * 128 + ((c ^ (k & 7)) << 4) + (byte & 15)) = __float2bfloat16(B[i]);
    }
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm = slot;
    {
        uint32_t r[16];
        for (int j = 0; j < 16; ++j) {
            const __nv_bfloat16 lo = __float2bfloat16(A[t * K + 2 * j]), hi = __float2bfloat16(A[t * K + 2 * j + 1]);
            r[j] = (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
        }
        tmem_st_x16(tm + ((uint32_t)(warp * 32) << 16) + 64, r);
        tmem_st_wait();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    if (warp == 0) {
        const uint32_t idesc = idesc_bf16(128, N) | (1u << 16);                  // B MN-major
        for (int kk = 0; kk < 2; ++kk) {
            const uint64_t db = umma_smem_desc(smem_u32(smem) + kk * 2048 + n_off * 2, lbo, sbo, 2);
            umma_ts_elect<true>(tm, tm + 64 + kk * 8, db, idesc, kk != 0);
        }
        umma_commit_elect(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after_sync();
    float v[32];
    for (int c = 0; c < N; c += 32) {
        tmem_ld_32x32(tm + ((uint32_t)(warp * 32) << 16) + c, v);
        for (int j = 0; j < 32; ++j) D[t * 64 + c + j] = v[j];
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 128);
}
static int check_mn(int N, int n_off, int sbo, int lbo) {
    const int K = 32;
    float *hA = (float*)malloc(128 * K * 4), *hB = (float*)malloc(64 * K * 4), *hD = (float*)malloc(128 * 64 * 4);
    srand(2);
    for (int i = 0; i < 128 * K; ++i) hA[i] = (float)((rand() % 17) - 8) / 8.f;
    for (int i = 0; i < 64 * K; ++i) hB[i] = (float)((rand() % 17) - 8) / 4.f;
    float *dA, *dB, *dD;
    cudaMalloc(&dA, 128 * K * 4); cudaMalloc(&dB, 64 * K * 4); cudaMalloc(&dD, 128 * 64 * 4);
    cudaMemcpy(dA, hA, 128 * K * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, 64 * K * 4, cudaMemcpyHostToDevice);
    cudaMemset(dD, 0, 128 * 64 * 4);
    cudaFuncSetAttribute(numerics_mn, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 * 1024);
    numerics_mn<<<1, 128, 32 * 1024>>>(dA, dB, dD, N, n_off, sbo, lbo);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("MN-major numerics: %s\n", cudaGetErrorString(e)); return 1; }
    cudaMemcpy(hD, dD, 128 * 64 * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)hA[m * K + k] * hB[k * 64 + n_off + n];
            maxerr = fmax(maxerr, fabs(s - hD[m * 64 + n]));
        }
    printf("B MN-major SW128 bf16, N=%d starting at column %d, SBO=%d LBO=%d: max |err| = %g  %s\n", N, n_off, sbo, lbo, maxerr,
           maxerr == 0 ? "EXACT" : "MISMATCH");
    return maxerr != 0;
}

int main() {
    for (int sbo : {1024, 2048}) for (int lbo : {1024, 2048, 4096}) { check_mn(64, 0, sbo, lbo); check_mn(32, 0, sbo, lbo); check_mn(32, 32, sbo, lbo); }
    if (getenv("TS_PROBE_MN_ONLY")) return 0;
    int bad = check<true>("bf16") + check<false>("tf32");
    long long* d;
    cudaMalloc(&d, 8);
    timing<true, 0, 1>(d, "bf16"); timing<true, 1, 1>(d, "bf16");
    timing<true, 0, 2>(d, "bf16"); timing<true, 1, 2>(d, "bf16");
    timing<true, 0, 4>(d, "bf16"); timing<true, 1, 4>(d, "bf16");
    timing<false, 0, 1>(d, "tf32"); timing<false, 1, 1>(d, "tf32");
    timing<false, 0, 4>(d, "tf32"); timing<false, 1, 4>(d, "tf32");
    return bad;
}
