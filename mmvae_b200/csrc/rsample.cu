// Kernel (1): fused reparameterised rsample for all modalities and all K importance samples.
// Reference ops replaced: laplace_vae.py:69-71 (std = softmax(log_var)*L + 1e-6; Laplace.rsample),
// vae_model_adapted.py:84-86,133-137 (std = exp(0.5 log_var); mu + eps*std), joint_encoders.py:81-82,
// executed by the reference on the K-tiled batch (mmvae.py:41-47).  Here mu/raw are (M,B,L) and are broadcast
// over K inside the kernel, so the K-fold tiling never exists.
#include "common.cuh"

#define RS_THREADS 128

__device__ __forceinline__ float noise_to_unit(int dist, float n) {
    // z = mu + std * t(n):  Normal t = n;  Laplace t = -sign(u) * log1p(-|u|)   (torch Laplace.rsample)
    if (dist == MMVAE_DIST_LAPLACE) {
        const float sg = (float)((n > 0.f) - (n < 0.f));
        return -sg * log1pf(-fabsf(n));
    }
    return n;
}

// one block per (m,b) row
__global__ void __launch_bounds__(RS_THREADS)
rsample_fwd_kernel(const float* __restrict__ mu, const float* __restrict__ raw, const float* __restrict__ noise,
                   float* __restrict__ z, float* __restrict__ std_out, int K, int B, int L, int dist, int xform,
                   int bk_layout) {
    extern __shared__ float sm[];          // [L] std, [L] mu, [32] red
    float* s_std = sm;
    float* s_mu = sm + L;
    float* red = sm + 2 * L;
    const int row = blockIdx.x;            // m*B + b
    const int m = row / B, b = row - m * B;
    const float* r = raw + (long)row * L;
    if (xform == MMVAE_XFORM_SOFTMAX) {
        float mx = -INFINITY;
        for (int l = threadIdx.x; l < L; l += blockDim.x) mx = fmaxf(mx, r[l]);
        mx = block_max(mx, red);
        float se = 0.f;
        for (int l = threadIdx.x; l < L; l += blockDim.x) se += expf(r[l] - mx);
        se = block_sum(se, red);
        for (int l = threadIdx.x; l < L; l += blockDim.x)
            s_std[l] = __fadd_rn(__fmul_rn(expf(r[l] - mx) / se, (float)L), MMVAE_ETA);
    } else {
        const float add = (xform == MMVAE_XFORM_EXP_HALF_EPS) ? MMVAE_ETA : 0.f;
        for (int l = threadIdx.x; l < L; l += blockDim.x) s_std[l] = expf(0.5f * r[l]) + add;
    }
    for (int l = threadIdx.x; l < L; l += blockDim.x) s_mu[l] = mu[(long)row * L + l];
    __syncthreads();
    for (int l = threadIdx.x; l < L; l += blockDim.x) std_out[(long)row * L + l] = s_std[l];
    const int n = K * L;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int k = i / L, l = i - k * L;
        const long off = bk_layout ? (((long)row * K + k) * L + l) : ((((long)m * K + k) * B + b) * L + l);
        z[off] = fmaf(s_std[l], noise_to_unit(dist, noise[off]), s_mu[l]);
    }
}

__global__ void __launch_bounds__(RS_THREADS)
rsample_bwd_kernel(const float* __restrict__ gz, const float* __restrict__ gstd_direct,
                   const float* __restrict__ noise, const float* __restrict__ raw, const float* __restrict__ stdv,
                   float* __restrict__ gmu, float* __restrict__ graw, int K, int B, int L, int dist, int xform,
                   int bk_layout) {
    extern __shared__ float sm[];          // [L] gmu, [L] gstd, [2*RS_THREADS] partials, [32] red
    float* s_gmu = sm;
    float* s_gstd = sm + L;
    float* part = sm + 2 * L;
    float* red = part + 2 * RS_THREADS;
    const int row = blockIdx.x;
    const int m = row / B, b = row - m * B;
    const int Lc = min(L, (int)blockDim.x);
    const int nsl = blockDim.x / Lc;
    const int lc = threadIdx.x % Lc, sl = threadIdx.x / Lc;
    for (int l0 = 0; l0 < L; l0 += Lc) {
        const int l = l0 + lc;
        float amu = 0.f, astd = 0.f;
        if (sl < nsl && l < L) {
            for (int k = sl; k < K; k += nsl) {
                const long off = bk_layout ? (((long)row * K + k) * L + l) : ((((long)m * K + k) * B + b) * L + l);
                const float g = gz[off];
                amu += g;
                astd = fmaf(g, noise_to_unit(dist, noise[off]), astd);
            }
        }
        part[threadIdx.x] = amu;
        part[RS_THREADS + threadIdx.x] = astd;
        __syncthreads();
        if (sl == 0 && l < L) {
            for (int s = 1; s < nsl; ++s) {
                amu += part[s * Lc + lc];
                astd += part[RS_THREADS + s * Lc + lc];
            }
            if (gstd_direct) astd += gstd_direct[(long)row * L + l];
            s_gmu[l] = amu;
            s_gstd[l] = astd;
        }
        __syncthreads();
    }
    const float* sd = stdv + (long)row * L;
    if (xform == MMVAE_XFORM_SOFTMAX) {
        // std_l = L*s_l + eta, s = softmax(raw):  d/d raw_l = L * s_l * (g_l - sum_j g_j s_j)
        float dot = 0.f;
        for (int l = threadIdx.x; l < L; l += blockDim.x) dot += s_gstd[l] * ((sd[l] - MMVAE_ETA) / (float)L);
        dot = block_sum(dot, red);
        for (int l = threadIdx.x; l < L; l += blockDim.x) {
            const float s = (sd[l] - MMVAE_ETA) / (float)L;
            graw[(long)row * L + l] = (float)L * s * (s_gstd[l] - dot);
        }
    } else {
        const float sub = (xform == MMVAE_XFORM_EXP_HALF_EPS) ? MMVAE_ETA : 0.f;
        for (int l = threadIdx.x; l < L; l += blockDim.x)
            graw[(long)row * L + l] = s_gstd[l] * 0.5f * (sd[l] - sub);
    }
    for (int l = threadIdx.x; l < L; l += blockDim.x) gmu[(long)row * L + l] = s_gmu[l];
    (void)raw;
}

extern "C" int mmvae_rsample_fwd(const float* mu, const float* raw, const float* noise, float* z, float* std,
                                 int M, int K, int B, int L, int dist, int xform, int bk_layout, void* stream) {
    MMVAE_REQUIRE(mu && raw && noise && z && std, "rsample_fwd: null pointer");
    MMVAE_REQUIRE(M > 0 && K > 0 && B > 0 && L > 0 && L <= 4096, "rsample_fwd: bad sizes M=%d K=%d B=%d L=%d", M, K, B, L);
    MMVAE_REQUIRE(dist == 0 || dist == 1, "rsample_fwd: bad dist %d", dist);
    MMVAE_REQUIRE(xform >= 0 && xform <= 2, "rsample_fwd: bad xform %d", xform);
    const size_t smem = (2 * (size_t)L + 32) * sizeof(float);
    rsample_fwd_kernel<<<M * B, RS_THREADS, smem, (cudaStream_t)stream>>>(mu, raw, noise, z, std, K, B, L, dist,
                                                                         xform, bk_layout);
    MMVAE_LAUNCH_CHECK("rsample_fwd");
    return 0;
}

extern "C" int mmvae_rsample_bwd(const float* grad_z, const float* grad_std_direct, const float* noise,
                                 const float* raw, const float* std, float* grad_mu, float* grad_raw, int M, int K,
                                 int B, int L, int dist, int xform, int bk_layout, void* stream) {
    MMVAE_REQUIRE(grad_z && noise && raw && std && grad_mu && grad_raw, "rsample_bwd: null pointer");
    MMVAE_REQUIRE(M > 0 && K > 0 && B > 0 && L > 0 && L <= 4096, "rsample_bwd: bad sizes");
    MMVAE_REQUIRE(dist == 0 || dist == 1, "rsample_bwd: bad dist %d", dist);
    MMVAE_REQUIRE(xform >= 0 && xform <= 2, "rsample_bwd: bad xform %d", xform);
    const size_t smem = (2 * (size_t)L + 2 * RS_THREADS + 32) * sizeof(float);
    rsample_bwd_kernel<<<M * B, RS_THREADS, smem, (cudaStream_t)stream>>>(grad_z, grad_std_direct, noise, raw, std,
                                                                         grad_mu, grad_raw, K, B, L, dist, xform,
                                                                         bk_layout);
    MMVAE_LAUNCH_CHECK("rsample_bwd");
    return 0;
}
