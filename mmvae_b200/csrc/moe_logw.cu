// Kernel (3): cross-modal log q(z|x_e) mixture + stable two-level logsumexp (experts, then K) + objective,
// with the stop-gradient importance weights produced in the forward and consumed by a fused backward.
// Reference ops replaced (~40 small eager kernels): objectives.py:117-131 (_m_iwae), :343-359 (_m_iwae_looser),
// :372-403 (_m_dreg/m_dreg), :406-438 (_m_dreg_looser/m_dreg_looser), utils.py:143-144 (log_mean_exp).
// Backward semantics follow the reference as it actually runs: the DReG hook never fires,
// so d loss/d lw = softmax weights taken as constants and (mu,std) are detached inside log q in the DReG modes.
#include "common.cuh"

#define MOE_MAX_M 8
#define MOE_THREADS 128

// one block per datapoint b; threads stride over the P = M*K (expert r, sample k) pairs.
// loss_per[b] = per-datapoint term; loss[0] = their sum (moe_loss_reduce_kernel).
// Precision: every elementwise term is fp32 (as in the reference); the SUMS that form lw and the softmax over the
// samples are carried in fp64 (lpx arrives as a (hi, lo) fp32 pair from the log-likelihood kernel).  |lw| is 3000..9000
// at the MNIST-SVHN shapes, where one fp32 ulp (2.4e-4..1e-3) is a 1e-3 relative error on a softmax weight — the
// fp32 reference carries that noise (measured 2e-4..6e-4 on encoder gradients against its own fp64 evaluation,
// DESIGN.md §parity); this kernel does not.
__global__ void __launch_bounds__(MOE_THREADS)
moe_logw_fwd_kernel(const float* __restrict__ z, const float* __restrict__ mu, const float* __restrict__ stdv,
                    const float* __restrict__ lpx, const float* __restrict__ lpx_lo,
                    const float* __restrict__ pz_loc, const float* __restrict__ pz_scale, float* __restrict__ lw,
                    float* __restrict__ wt, float* __restrict__ loss_per, int M, int K, int B, int L, int dist,
                    int mode) {
    extern __shared__ double smd[];               // red[32], lw[P] (fp64), then floats: mu[M*L], std[M*L], ploc[L], pscale[L], z[P*L]
    const int P = M * K;
    double* red = smd;
    double* s_lw = smd + 32;
    float* s_mu = (float*)(s_lw + P);
    float* s_std = s_mu + M * L;
    float* s_pl = s_std + M * L;
    float* s_ps = s_pl + L;
    float* s_z = s_ps + L;                        // the block's P latent rows: every load of the block is in flight at once
    const int b = blockIdx.x;                     // (a thread walking its own row paid one L2 round trip per element)
    for (int i = threadIdx.x; i < P * L; i += blockDim.x) {
        const int p = i / L, l = i - p * L;
        s_z[i] = z[((long)p * B + b) * L + l];
    }
    for (int i = threadIdx.x; i < M * L; i += blockDim.x) {
        const int e = i / L, l = i - e * L;
        s_mu[i] = mu[((long)e * B + b) * L + l];
        s_std[i] = stdv[((long)e * B + b) * L + l];
    }
    for (int l = threadIdx.x; l < L; l += blockDim.x) { s_pl[l] = pz_loc[l]; s_ps[l] = pz_scale[l]; }
    __syncthreads();
    const double log_M = log((double)M);
    for (int p = threadIdx.x; p < P; p += blockDim.x) {
        const int r = p / K, k = p - r * K;
        const float* zr = s_z + p * L;
        double lpz = 0.0, lq[MOE_MAX_M];
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) lq[e] = 0.0;
        for (int l = 0; l < L; ++l) {
            const float zz = zr[l];
            lpz += (double)lat_logprob(dist, zz, s_pl[l], s_ps[l]);
#pragma unroll
            for (int e = 0; e < MOE_MAX_M; ++e)
                if (e < M) lq[e] += (double)lat_logprob(dist, zz, s_mu[e * L + l], s_std[e * L + l]);
        }
        double mx = -INFINITY;
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) if (e < M) mx = fmax(mx, lq[e]);
        double se = 0.0;
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) if (e < M) se += exp(lq[e] - mx);
        const double lqmix = mx + log(se) - log_M;
        double lpxs = 0.0;
        for (int d = 0; d < M; ++d) {
            const long o = (((long)r * M + d) * K + k) * B + b;
            lpxs += (double)lpx[o];
            if (lpx_lo) lpxs += (double)lpx_lo[o];
        }
        const double v = (lpz + lpxs) - lqmix;
        s_lw[p] = v;
        lw[((long)r * K + k) * B + b] = (float)v;
    }
    __syncthreads();
    const bool looser = (mode == MMVAE_OBJ_IWAE_LOOSER || mode == MMVAE_OBJ_DREG_LOOSER);
    const bool dreg = (mode == MMVAE_OBJ_DREG || mode == MMVAE_OBJ_DREG_LOOSER);
    const int G = looser ? M : 1;                  // softmax groups: per expert (looser) or over all M*K (tight)
    const int Pg = P / G;
    double obj = 0.0;
    for (int g = 0; g < G; ++g) {
        double mx = -INFINITY;
        for (int q = threadIdx.x; q < Pg; q += blockDim.x) mx = fmax(mx, s_lw[g * Pg + q]);
        mx = block_max_d(mx, red);
        double se = 0.0;
        for (int q = threadIdx.x; q < Pg; q += blockDim.x) se += exp(s_lw[g * Pg + q] - mx);
        se = block_sum_d(se, red);
        const double lse = mx + log(se);
        const double inv_se = 1.0 / se;
        const double gscale = looser ? 1.0 / (double)M : 1.0;
        double dot = 0.0;
        for (int q = threadIdx.x; q < Pg; q += blockDim.x) {
            const int p = g * Pg + q;
            const double v = s_lw[p];
            const double w = exp(v - mx) * inv_se;        // (lw - logsumexp(lw)).exp() (objectives.py:399,435)
            wt[(long)p * B + b] = (float)(w * gscale);
            dot = fma(w, v, dot);
        }
        if (dreg) {
            dot = block_sum_d(dot, red);
            obj += dot * gscale;                   // (grad_wt * lw).sum() / .mean(0).sum()
        } else {
            obj += (lse - log((double)Pg)) * gscale;   // log_mean_exp
        }
    }
    if (threadIdx.x == 0) loss_per[b] = (float)obj;
}

__global__ void moe_loss_reduce_kernel(float* __restrict__ loss, const float* __restrict__ loss_per, int B) {
    __shared__ double red[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < B; i += blockDim.x) s += (double)loss_per[i];
    s = block_sum_d(s, red);
    if (threadIdx.x == 0) loss[0] = (float)s;
}

// backward.  dynamic smem: mu[M*L], std[M*L], ploc[L], pscale[L], cpi[P*M], z[P*L]
__global__ void __launch_bounds__(MOE_THREADS)
moe_logw_bwd_kernel(const float* __restrict__ gloss, const float* __restrict__ z, const float* __restrict__ mu,
                    const float* __restrict__ stdv, const float* __restrict__ wt, const float* __restrict__ pz_loc,
                    const float* __restrict__ pz_scale, float* __restrict__ gz, float* __restrict__ gmu,
                    float* __restrict__ gstd, float* __restrict__ glpx, int M, int K, int B, int L, int dist,
                    int mode) {
    extern __shared__ float sm[];
    float* s_mu = sm;
    float* s_std = sm + M * L;
    float* s_pl = sm + 2 * M * L;
    float* s_ps = s_pl + L;
    float* s_cpi = s_ps + L;                      // [P][M]: c_p * pi_e
    const int b = blockIdx.x;
    const int P = M * K;
    float* s_z = s_cpi + P * M;                   // the block's P latent rows (see the forward kernel)
    const float g = gloss[0];
    for (int i = threadIdx.x; i < P * L; i += blockDim.x) {
        const int p = i / L, l = i - p * L;
        s_z[i] = z[((long)p * B + b) * L + l];
    }
    for (int i = threadIdx.x; i < M * L; i += blockDim.x) {
        const int e = i / L, l = i - e * L;
        s_mu[i] = mu[((long)e * B + b) * L + l];
        s_std[i] = stdv[((long)e * B + b) * L + l];
    }
    for (int l = threadIdx.x; l < L; l += blockDim.x) { s_pl[l] = pz_loc[l]; s_ps[l] = pz_scale[l]; }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += blockDim.x) {
        const int r = p / K, k = p - r * K;
        const long zo = (((long)r * K + k) * B + b) * L;
        const float c = g * wt[(long)p * B + b];
        float lq[MOE_MAX_M];
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) lq[e] = 0.f;
        for (int l = 0; l < L; ++l) {
            const float zz = s_z[p * L + l];
#pragma unroll
            for (int e = 0; e < MOE_MAX_M; ++e)
                if (e < M) lq[e] += lat_logprob(dist, zz, s_mu[e * L + l], s_std[e * L + l]);
        }
        float mx = -INFINITY;
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) if (e < M) mx = fmaxf(mx, lq[e]);
        float se = 0.f;
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) if (e < M) { lq[e] = expf(lq[e] - mx); se += lq[e]; }
        const float inv = 1.f / se;
#pragma unroll
        for (int e = 0; e < MOE_MAX_M; ++e) if (e < M) { lq[e] *= inv; s_cpi[p * M + e] = c * lq[e]; }
        for (int l = 0; l < L; ++l) {
            const float zz = s_z[p * L + l];
            float d = lat_dlogprob_dz(dist, zz, s_pl[l], s_ps[l]);
#pragma unroll
            for (int e = 0; e < MOE_MAX_M; ++e)
                if (e < M) d -= lq[e] * lat_dlogprob_dz(dist, zz, s_mu[e * L + l], s_std[e * L + l]);
            gz[zo + l] = c * d;
        }
        for (int d = 0; d < M; ++d) glpx[(((long)r * M + d) * K + k) * B + b] = c;
    }
    __syncthreads();
    const bool dreg = (mode == MMVAE_OBJ_DREG || mode == MMVAE_OBJ_DREG_LOOSER);
    // gradient through log q_e w.r.t. (mu_e, std_e): deterministic loop over pairs, one thread per (e,l)
    for (int i = threadIdx.x; i < M * L; i += blockDim.x) {
        const int e = i / L, l = i - e * L;
        float am = 0.f, as = 0.f;
        if (!dreg) {
            const float m_ = s_mu[i], s_ = s_std[i];
            for (int p = 0; p < P; ++p) {
                const float zz = s_z[p * L + l];
                const float w = s_cpi[p * M + e];
                am += w * lat_dlogprob_dz(dist, zz, m_, s_);        // -(-d/dz) : lw has -lqmix, d/dmu = -d/dz
                as -= w * lat_dlogprob_dscale(dist, zz, m_, s_);
            }
        }
        gmu[((long)e * B + b) * L + l] = am;
        gstd[((long)e * B + b) * L + l] = as;
    }
}

extern "C" int mmvae_moe_logw_fwd(const float* z, const float* mu, const float* std, const float* lpx,
                                  const float* lpx_lo, const float* pz_loc, const float* pz_scale, float* lw, float* wt, float* loss,
                                  float* loss_per, int M, int K, int B, int L, int dist, int mode, void* stream) {
    MMVAE_REQUIRE(z && mu && std && lpx && pz_loc && pz_scale && lw && wt && loss && loss_per, "moe_logw_fwd: null pointer");
    MMVAE_REQUIRE(M >= 1 && M <= MOE_MAX_M && K > 0 && B > 0 && L > 0, "moe_logw_fwd: bad sizes M=%d K=%d B=%d L=%d", M, K, B, L);
    MMVAE_REQUIRE(dist == 0 || dist == 1, "moe_logw_fwd: bad dist");
    MMVAE_REQUIRE(mode >= 0 && mode <= 3, "moe_logw_fwd: bad mode %d", mode);
    const size_t smem = ((size_t)32 + (size_t)M * K) * sizeof(double) + ((size_t)2 * M * L + 2 * L + (size_t)M * K * L) * sizeof(float);
    MMVAE_REQUIRE(smem <= 200 * 1024, "moe_logw_fwd: M*K (+ M*L) too large for shared memory");
    cudaStream_t st = (cudaStream_t)stream;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(moe_logw_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "moe_logw_fwd: %s", cudaGetErrorString(e));
    }
    moe_logw_fwd_kernel<<<B, MOE_THREADS, smem, st>>>(z, mu, std, lpx, lpx_lo, pz_loc, pz_scale, lw, wt, loss_per, M, K, B, L,
                                                      dist, mode);
    MMVAE_LAUNCH_CHECK("moe_logw_fwd");
    moe_loss_reduce_kernel<<<1, 256, 0, st>>>(loss, loss_per, B);
    MMVAE_LAUNCH_CHECK("moe_loss_reduce");
    return 0;
}

extern "C" int mmvae_moe_logw_bwd(const float* grad_loss, const float* z, const float* mu, const float* std,
                                  const float* wt, const float* pz_loc, const float* pz_scale, float* grad_z,
                                  float* grad_mu, float* grad_std, float* grad_lpx, int M, int K, int B, int L,
                                  int dist, int mode, void* stream) {
    MMVAE_REQUIRE(grad_loss && z && mu && std && wt && pz_loc && pz_scale && grad_z && grad_mu && grad_std && grad_lpx,
                  "moe_logw_bwd: null pointer");
    MMVAE_REQUIRE(M >= 1 && M <= MOE_MAX_M && K > 0 && B > 0 && L > 0, "moe_logw_bwd: bad sizes");
    MMVAE_REQUIRE(mode >= 0 && mode <= 3, "moe_logw_bwd: bad mode %d", mode);
    const size_t smem = ((size_t)2 * M * L + 2 * L + (size_t)M * K * M + (size_t)M * K * L) * sizeof(float);
    MMVAE_REQUIRE(smem <= 200 * 1024, "moe_logw_bwd: M*K*M too large for shared memory");
    cudaStream_t st = (cudaStream_t)stream;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(moe_logw_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "moe_logw_bwd: %s", cudaGetErrorString(e));
    }
    moe_logw_bwd_kernel<<<B, MOE_THREADS, smem, st>>>(grad_loss, z, mu, std, wt, pz_loc, pz_scale, grad_z, grad_mu,
                                                      grad_std, grad_lpx, M, K, B, L, dist, mode);
    MMVAE_LAUNCH_CHECK("moe_logw_bwd");
    return 0;
}
