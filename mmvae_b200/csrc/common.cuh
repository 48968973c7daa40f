// Shared device/host helpers for libmmvae_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include "../../include/mmvae_b200.h"

#define MMVAE_LOG_SQRT_2PI 0.91893853320467274178f   // log(sqrt(2*pi))  (torch Normal.log_prob)
#define MMVAE_LOG2         0.69314718055994530942f
#define MMVAE_ETA          1e-6f                     // utils.py:17 Constants.eta / laplace_vae.py:69
#define MMVAE_F32_EPS      1.1920928955078125e-07f   // torch.finfo(float32).eps (clamp_probs)

extern char g_mmvae_err[512];
int mmvae_set_error(int code, const char* fmt, ...);

#define MMVAE_REQUIRE(cond, ...)                                   \
    do {                                                           \
        if (!(cond)) return mmvae_set_error(1, __VA_ARGS__);       \
    } while (0)

#define MMVAE_LAUNCH_CHECK(name)                                                           \
    do {                                                                                   \
        cudaError_t _e = cudaGetLastError();                                               \
        if (_e != cudaSuccess)                                                             \
            return mmvae_set_error(1000 + (int)_e, "%s: %s", name, cudaGetErrorString(_e)); \
    } while (0)

#define MMVAE_NUM_SMS 148

static inline int ceil_div_i(long a, long b) { return (int)((a + b - 1) / b); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum; `red` is >= 32 floats of shared memory.  Result valid in every thread.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : 0.f;
    t = warp_sum(t);
    return t;
}
__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : -INFINITY;
    t = warp_max(t);
    return t;
}

// fp64 variants; `red` is >= 32 doubles of shared memory
__device__ __forceinline__ double block_sum_d(double v, double* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum_d(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    double t = (lane < nw) ? red[lane] : 0.0;
    return warp_sum_d(t);
}
__device__ __forceinline__ double block_max_d(double v, double* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max_d(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    double t = (lane < nw) ? red[lane] : -INFINITY;
    return warp_max_d(t);
}

// streaming 128-bit load that does not allocate in L1 (data touched once)
__device__ __forceinline__ float4 ld_stream4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream4(float4* p, const float4& v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// elementwise log densities of the latent families, as torch.distributions computes them
__device__ __forceinline__ float lat_logprob(int dist, float z, float loc, float scale) {
    if (dist == MMVAE_DIST_LAPLACE) return -logf(2.f * scale) - fabsf(z - loc) / scale;
    const float d = z - loc;
    return -(d * d) / (2.f * scale * scale) - logf(scale) - MMVAE_LOG_SQRT_2PI;
}
// d/dz of the above
__device__ __forceinline__ float lat_dlogprob_dz(int dist, float z, float loc, float scale) {
    if (dist == MMVAE_DIST_LAPLACE) {
        const float d = z - loc;
        return -((d > 0.f) - (d < 0.f)) / scale;
    }
    return -(z - loc) / (scale * scale);
}
// d/dscale
__device__ __forceinline__ float lat_dlogprob_dscale(int dist, float z, float loc, float scale) {
    if (dist == MMVAE_DIST_LAPLACE) return -1.f / scale + fabsf(z - loc) / (scale * scale);
    const float d = z - loc;
    return (d * d) / (scale * scale * scale) - 1.f / scale;
}
