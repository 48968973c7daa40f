// m_elbo (objectives.py:87-111) and the importance-sampled likelihood evaluators' log-weight / logsumexp pieces
// (mmvae.py:139-175,198-232; multi_vaes.py:290-316; jmvae_nf.py:116-141,231-268).
#include "common.cuh"

#define EL_MAX_M 8
#define EL_THREADS 128

__device__ __forceinline__ float lat_kl(int dist, float loc, float scale, float ploc, float pscale) {
    if (dist == MMVAE_DIST_LAPLACE) {         // torch _kl_laplace_laplace
        const float r = scale / pscale, d = fabsf(loc - ploc);
        return -logf(r) + d / pscale + r * expf(-d / scale) - 1.f;
    }
    const float vr = (scale / pscale) * (scale / pscale);   // torch _kl_normal_normal
    const float t1 = ((loc - ploc) / pscale) * ((loc - ploc) / pscale);
    return 0.5f * (vr + t1 - 1.f - logf(vr));
}
__device__ __forceinline__ void lat_kl_grad(int dist, float loc, float scale, float ploc, float pscale, float& dloc,
                                            float& dscale) {
    if (dist == MMVAE_DIST_LAPLACE) {
        const float dd = loc - ploc, d = fabsf(dd), ex = expf(-d / scale);
        const float sg = (float)((dd > 0.f) - (dd < 0.f));
        dloc = sg * (1.f / pscale - ex / pscale);
        dscale = -1.f / scale + ex / pscale * (1.f + d / scale);
        return;
    }
    dloc = (loc - ploc) / (pscale * pscale);
    dscale = scale / (pscale * pscale) - 1.f / scale;
}

// one block per datapoint.  smem: mu[M*L], std[M*L], lwt[M*M*K] (log importance ratio per (r,d,k)), red[32]
// BWD=false: loss[b] (per-datapoint) and terms;  BWD=true: gradients.
template <bool BWD>
__global__ void __launch_bounds__(EL_THREADS)
elbo_kernel(const float* __restrict__ gloss, const float* __restrict__ z, const float* __restrict__ mu,
            const float* __restrict__ stdv, const float* __restrict__ lpx, const float* __restrict__ pz_loc,
            const float* __restrict__ pz_scale, float* __restrict__ loss, float* __restrict__ terms,
            float* __restrict__ gmu, float* __restrict__ gstd, float* __restrict__ glpx, int M, int K, int B, int L,
            int dist) {
    extern __shared__ float sm[];
    float* s_mu = sm;
    float* s_std = sm + M * L;
    float* s_w = sm + 2 * M * L;               // [M(r)][M(d)][K]
    float* red = s_w + M * M * K;
    const int b = blockIdx.x;
    for (int i = threadIdx.x; i < M * L; i += blockDim.x) {
        const int e = i / L, l = i - e * L;
        s_mu[i] = mu[((long)e * B + b) * L + l];
        s_std[i] = stdv[((long)e * B + b) * L + l];
    }
    __syncthreads();
    // importance ratio w[r,d,k] = exp(sum_l log q_r(z_d) - log q_d(z_d)), 1 on the diagonal (objectives.py:100-106)
    for (int i = threadIdx.x; i < M * M * K; i += blockDim.x) {
        const int r = i / (M * K), d = (i / K) % M, k = i % K;
        float w = 1.f;
        if (r != d) {
            const float* zd = z + (((long)d * K + k) * B + b) * L;
            float acc = 0.f;
            for (int l = 0; l < L; ++l)
                acc += lat_logprob(dist, zd[l], s_mu[r * L + l], s_std[r * L + l]) -
                       lat_logprob(dist, zd[l], s_mu[d * L + l], s_std[d * L + l]);
            w = expf(acc);
        }
        s_w[i] = w;
    }
    __syncthreads();
    const float invM = 1.f / (float)M, invK = 1.f / (float)K;
    if (!BWD) {
        float obj = 0.f;
        for (int rd = 0; rd < M * M; ++rd) {
            const int d = rd % M;
            float t = 0.f;
            for (int k = threadIdx.x; k < K; k += blockDim.x) t += s_w[rd * K + k] * lpx[((long)d * K + k) * B + b];
            t = block_sum(t, red);
            if (threadIdx.x == 0) atomicAdd(terms + rd, t);
            obj += t * invK;
        }
        float kl = 0.f;
        for (int i = threadIdx.x; i < M * L; i += blockDim.x) {
            const int l = i % L;
            kl += lat_kl(dist, s_mu[i], s_std[i], pz_loc[l], pz_scale[l]);
        }
        kl = block_sum(kl, red);
        if (threadIdx.x == 0) loss[b] = invM * (obj - kl);
    } else {
        const float g = gloss[0] * invM;
        for (int i = threadIdx.x; i < M * K; i += blockDim.x) {           // d lpx_diag[d,k,b]
            const int d = i / K, k = i - d * K;
            float s = 0.f;
            for (int r = 0; r < M; ++r) s += s_w[(r * M + d) * K + k];
            glpx[((long)d * K + k) * B + b] = g * invK * s;
        }
        for (int i = threadIdx.x; i < M * L; i += blockDim.x) {           // d (mu_r, std_r)[l]
            const int r = i / L, l = i - r * L;
            const float m_ = s_mu[i], s_ = s_std[i];
            float dl, ds;
            lat_kl_grad(dist, m_, s_, pz_loc[l], pz_scale[l], dl, ds);
            float am = -g * dl, as = -g * ds;
            for (int d = 0; d < M; ++d) {
                if (d == r) continue;
                for (int k = 0; k < K; ++k) {
                    const float c = g * invK * s_w[(r * M + d) * K + k] * lpx[((long)d * K + k) * B + b];
                    const float zz = z[(((long)d * K + k) * B + b) * L + l];
                    am -= c * lat_dlogprob_dz(dist, zz, m_, s_);          // d/dmu = -d/dz
                    as += c * lat_dlogprob_dscale(dist, zz, m_, s_);
                }
            }
            gmu[((long)r * B + b) * L + l] = am;
            gstd[((long)r * B + b) * L + l] = as;
        }
    }
}

__global__ void sum_tail_kernel(float* __restrict__ loss, const float* __restrict__ per, int B) {
    __shared__ float red[32];
    float s = 0.f;
    for (int i = threadIdx.x; i < B; i += blockDim.x) s += per[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) loss[0] = s;
}

static int elbo_check(int M, int K, int B, int L, int dist, size_t* smem) {
    MMVAE_REQUIRE(M >= 1 && M <= EL_MAX_M && K > 0 && B > 0 && L > 0, "elbo: bad sizes M=%d K=%d B=%d L=%d", M, K, B, L);
    MMVAE_REQUIRE(dist == 0 || dist == 1, "elbo: bad dist");
    *smem = ((size_t)2 * M * L + (size_t)M * M * K + 32) * sizeof(float);
    MMVAE_REQUIRE(*smem <= 200 * 1024, "elbo: M*M*K too large for shared memory");
    return 0;
}

extern "C" int mmvae_elbo_fwd(const float* z, const float* mu, const float* std, const float* lpx_diag,
                              const float* pz_loc, const float* pz_scale, float* loss, float* loss_per, float* terms,
                              int M, int K, int B, int L, int dist, void* stream) {
    MMVAE_REQUIRE(z && mu && std && lpx_diag && pz_loc && pz_scale && loss && loss_per && terms, "elbo_fwd: null pointer");
    size_t smem;
    if (int rc = elbo_check(M, K, B, L, dist, &smem)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (smem > 48 * 1024) cudaFuncSetAttribute(elbo_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaMemsetAsync(terms, 0, sizeof(float) * M * M, st);
    elbo_kernel<false><<<B, EL_THREADS, smem, st>>>(nullptr, z, mu, std, lpx_diag, pz_loc, pz_scale, loss_per, terms,
                                                    nullptr, nullptr, nullptr, M, K, B, L, dist);
    MMVAE_LAUNCH_CHECK("elbo_fwd");
    sum_tail_kernel<<<1, 256, 0, st>>>(loss, loss_per, B);
    MMVAE_LAUNCH_CHECK("elbo_fwd_reduce");
    return 0;
}

extern "C" int mmvae_elbo_bwd(const float* grad_loss, const float* z, const float* mu, const float* std,
                              const float* lpx_diag, const float* pz_loc, const float* pz_scale, float* grad_mu,
                              float* grad_std, float* grad_lpx_diag, int M, int K, int B, int L, int dist,
                              void* stream) {
    MMVAE_REQUIRE(grad_loss && z && mu && std && lpx_diag && pz_loc && pz_scale && grad_mu && grad_std && grad_lpx_diag,
                  "elbo_bwd: null pointer");
    size_t smem;
    if (int rc = elbo_check(M, K, B, L, dist, &smem)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (smem > 48 * 1024) cudaFuncSetAttribute(elbo_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    elbo_kernel<true><<<B, EL_THREADS, smem, st>>>(grad_loss, z, mu, std, lpx_diag, pz_loc, pz_scale, nullptr, nullptr,
                                                   grad_mu, grad_std, grad_lpx_diag, M, K, B, L, dist);
    MMVAE_LAUNCH_CHECK("elbo_bwd");
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// evaluator pieces
// a[b,k] = sum_d lpx[d,b,k] + log p(z[b,k]) - qterm      (one thread per (b,k))
__global__ void __launch_bounds__(256)
eval_logw_kernel(const float* __restrict__ z, const float* __restrict__ mu, const float* __restrict__ stdv,
                 const float* __restrict__ lpx, const float* __restrict__ lq_ext, const float* __restrict__ pz_loc,
                 const float* __restrict__ pz_scale, float* __restrict__ a, int E, int D, int B, int K, int L,
                 int dist, int qmode) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)B * K) return;
    const int b = (int)(i / K);
    const float* zr = z + i * L;
    float lpz = 0.f, lq[EL_MAX_M];
#pragma unroll
    for (int e = 0; e < EL_MAX_M; ++e) lq[e] = 0.f;
    for (int l = 0; l < L; ++l) {
        const float zz = zr[l];
        lpz += lat_logprob(dist, zz, pz_loc[l], pz_scale[l]);
        if (qmode <= 2) {
#pragma unroll
            for (int e = 0; e < EL_MAX_M; ++e)
                if (e < E) lq[e] += lat_logprob(dist, zz, mu[((long)e * B + b) * L + l], stdv[((long)e * B + b) * L + l]);
        }
    }
    float q = 0.f;
    if (qmode == 0 || qmode == 1) {
        float mx = -INFINITY;
#pragma unroll
        for (int e = 0; e < EL_MAX_M; ++e) if (e < E) mx = fmaxf(mx, lq[e]);
        float se = 0.f;
#pragma unroll
        for (int e = 0; e < EL_MAX_M; ++e) if (e < E) se += expf(lq[e] - mx);
        const float lse = mx + logf(se);
        q = (qmode == 0) ? lse * 0.5f : lse - logf((float)E);       // mmvae.py:166 divides by 2
    } else if (qmode == 2) {
        q = lq[0];
    } else if (qmode == 4) {
        q = lq_ext[i];
    }
    float s = 0.f;
    for (int d = 0; d < D; ++d) s += lpx[(long)d * B * K + i];
    a[i] = (s + lpz) - q;
}

// per-datapoint running (max, sum exp): one block per b
__global__ void __launch_bounds__(256)
lse_partial_kernel(const float* __restrict__ a, float* __restrict__ mx, float* __restrict__ se, int Kc,
                   int accumulate) {
    __shared__ float red[32];
    const int b = blockIdx.x;
    const float* ar = a + (long)b * Kc;
    float m = -INFINITY;
    for (int k = threadIdx.x; k < Kc; k += blockDim.x) m = fmaxf(m, ar[k]);
    m = block_max(m, red);
    float s = 0.f;
    for (int k = threadIdx.x; k < Kc; k += blockDim.x) s += expf(ar[k] - m);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        if (accumulate) {
            const float m0 = mx[b], s0 = se[b];
            const float mm = fmaxf(m, m0);
            s = s * expf(m - mm) + s0 * expf(m0 - mm);
            m = mm;
        }
        mx[b] = m;
        se[b] = s;
    }
}
__global__ void lse_finish_kernel(const float* __restrict__ mx, const float* __restrict__ se, float* __restrict__ ll,
                                  int B, float log_K) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) ll[b] = mx[b] + logf(se[b]) - log_K;
}

extern "C" int mmvae_eval_logw(const float* z, const float* mu, const float* std, const float* lpx,
                               const float* lq_ext, const float* pz_loc, const float* pz_scale, float* a, int E,
                               int D, int B, int K, int L, int dist, int qmode, void* stream) {
    MMVAE_REQUIRE(z && lpx && pz_loc && pz_scale && a, "eval_logw: null pointer");
    MMVAE_REQUIRE(qmode >= 0 && qmode <= 4, "eval_logw: bad qmode %d", qmode);
    MMVAE_REQUIRE(qmode > 2 || (mu && std && E >= 1 && E <= EL_MAX_M), "eval_logw: proposal parameters missing");
    MMVAE_REQUIRE(qmode != 4 || lq_ext, "eval_logw: lq_ext missing");
    MMVAE_REQUIRE(D >= 1 && B > 0 && K > 0 && L > 0, "eval_logw: bad sizes");
    const long n = (long)B * K;
    eval_logw_kernel<<<ceil_div_i(n, 256), 256, 0, (cudaStream_t)stream>>>(z, mu, std, lpx, lq_ext, pz_loc, pz_scale,
                                                                          a, E, D, B, K, L, dist, qmode);
    MMVAE_LAUNCH_CHECK("eval_logw");
    return 0;
}
extern "C" int mmvae_lse_partial(const float* a, float* mx, float* se, int B, int Kc, int accumulate, void* stream) {
    MMVAE_REQUIRE(a && mx && se && B > 0 && Kc > 0, "lse_partial: bad arguments");
    lse_partial_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(a, mx, se, Kc, accumulate);
    MMVAE_LAUNCH_CHECK("lse_partial");
    return 0;
}
extern "C" int mmvae_lse_finish(const float* mx, const float* se, float* ll, int B, float log_K_total, void* stream) {
    MMVAE_REQUIRE(mx && se && ll && B > 0, "lse_finish: bad arguments");
    lse_finish_kernel<<<ceil_div_i(B, 128), 128, 0, (cudaStream_t)stream>>>(mx, se, ll, B, log_K_total);
    MMVAE_LAUNCH_CHECK("lse_finish");
    return 0;
}
