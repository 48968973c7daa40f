// Kernel (2): decoder log-likelihood reduced over pixels, one streaming pass over recon (R,D).
// Reference ops replaced: px_z(recon, scale=1).log_prob(x).view(K,B,-1).mul(lik_scaling).sum(-1)
// (objectives.py:381-385,416-419; mmvae.py:154-156; multi_vaes.py:240,306-311) — six elementwise kernels, each a
// full HBM round trip of the (K,B,D) tensor, plus view/mul/sum.  HBM-bound: algorithmic bytes = R*D*4 read (fwd),
// R*D*4 read + R*D*4 written (bwd); x (B,D) stays in L2.
#include "common.cuh"

#define LL_THREADS 256

template <int LIK>
__device__ __forceinline__ float ll_elem(float x, float r) {
    if (LIK == MMVAE_LIK_NORMAL) {            // -(x-r)^2/2 - log sqrt(2pi); the constant is added once per row
        const float d = x - r;
        return -0.5f * d * d;
    } else if (LIK == MMVAE_LIK_LAPLACE) {    // -log 2 - |x-r|
        return -fabsf(x - r);
    } else if (LIK == MMVAE_LIK_BERNOULLI) {  // Bernoulli(probs=clamp(r, eps, 1-eps))
        const float p = fminf(fmaxf(r, MMVAE_F32_EPS), 1.f - MMVAE_F32_EPS);
        return x * logf(p) + (1.f - x) * log1pf(-p);
    } else if (LIK == MMVAE_LIK_NEG_MSE) {    // -F.mse_loss(reduction='none'): no 1/2, no constant (objectives.py:177)
        const float d = r - x;
        return -(d * d);
    } else if (LIK == MMVAE_LIK_NEG_L1) {     // -F.l1_loss
        return -fabsf(r - x);
    } else {                                  // -F.binary_cross_entropy: logs clamped at -100 (torch semantics)
        return x * fmaxf(logf(r), -100.f) + (1.f - x) * fmaxf(log1pf(-r), -100.f);
    }
}

template <int LIK>
__device__ __forceinline__ float ll_grad(float x, float r) {
    if (LIK == MMVAE_LIK_NORMAL) return x - r;
    if (LIK == MMVAE_LIK_LAPLACE) {
        const float d = x - r;
        return (float)((d > 0.f) - (d < 0.f));
    }
    if (LIK == MMVAE_LIK_BERNOULLI) {
        if (r < MMVAE_F32_EPS || r > 1.f - MMVAE_F32_EPS) return 0.f;  // clamp passes no gradient outside
        return (x - r) / (r * (1.f - r));
    }
    if (LIK == MMVAE_LIK_NEG_MSE) return 2.f * (x - r);
    if (LIK == MMVAE_LIK_NEG_L1) {
        const float d = x - r;
        return (float)((d > 0.f) - (d < 0.f));
    }
    return (x - r) / fmaxf(r * (1.f - r), 1e-12f);                     // torch binary_cross_entropy_backward
}

__device__ __forceinline__ void ll_store(float* lp, float* lp_lo, long row, double v) {
    const float hi = (float)v;
    lp[row] = hi;
    if (lp_lo) lp_lo[row] = (float)(v - (double)hi);
}

// TPR threads cooperate on one row; LL_THREADS/TPR rows per block.  D4 = D/4.
template <int LIK, int TPR>
__global__ void __launch_bounds__(LL_THREADS)
loglik_fwd_kernel(const float4* __restrict__ recon, const float4* __restrict__ x, float* __restrict__ lp,
                  float* __restrict__ lp_lo, int R, int B, int D4, int x_div, float scale, double row_const) {
    __shared__ double red[LL_THREADS / 32];
    const int rows_per_block = LL_THREADS / TPR;
    const int sub = threadIdx.x / TPR, t = threadIdx.x % TPR;
    const long row = (long)blockIdx.x * rows_per_block + sub;
    // per-pixel terms are fp32 (as the reference's); their SUM is carried in fp64 so that lp (|lp| ~ 3000..9000, fp32
    // ulp 2.4e-4..1e-3) can be handed to the softmax over K as a (hi, lo) pair instead of losing those digits
    double acc = 0.0;
    if (row < R) {
        const float4* rr = recon + row * D4;
        const float4* xr = x + (long)((row / x_div) % B) * D4;
        int i = t;
        for (; i + 3 * TPR < D4; i += 4 * TPR) {       // 4 independent 128-bit loads in flight per thread
            const float4 a0 = ld_stream4(rr + i), a1 = ld_stream4(rr + i + TPR);
            const float4 a2 = ld_stream4(rr + i + 2 * TPR), a3 = ld_stream4(rr + i + 3 * TPR);
            const float4 x0 = __ldg(xr + i), x1 = __ldg(xr + i + TPR);
            const float4 x2 = __ldg(xr + i + 2 * TPR), x3 = __ldg(xr + i + 3 * TPR);
            acc += (double)(ll_elem<LIK>(x0.x, a0.x) + ll_elem<LIK>(x0.y, a0.y) + ll_elem<LIK>(x0.z, a0.z) + ll_elem<LIK>(x0.w, a0.w));
            acc += (double)(ll_elem<LIK>(x1.x, a1.x) + ll_elem<LIK>(x1.y, a1.y) + ll_elem<LIK>(x1.z, a1.z) + ll_elem<LIK>(x1.w, a1.w));
            acc += (double)(ll_elem<LIK>(x2.x, a2.x) + ll_elem<LIK>(x2.y, a2.y) + ll_elem<LIK>(x2.z, a2.z) + ll_elem<LIK>(x2.w, a2.w));
            acc += (double)(ll_elem<LIK>(x3.x, a3.x) + ll_elem<LIK>(x3.y, a3.y) + ll_elem<LIK>(x3.z, a3.z) + ll_elem<LIK>(x3.w, a3.w));
        }
        for (; i < D4; i += TPR) {
            const float4 a = ld_stream4(rr + i);
            const float4 v = __ldg(xr + i);
            acc += (double)(ll_elem<LIK>(v.x, a.x) + ll_elem<LIK>(v.y, a.y) + ll_elem<LIK>(v.z, a.z) + ll_elem<LIK>(v.w, a.w));
        }
    }
    acc = warp_sum_d(acc);
    if (TPR == 32) {
        if (t == 0 && row < R) ll_store(lp, lp_lo, row, (double)scale * (acc + row_const));
    } else {                                            // TPR == LL_THREADS: one row per block
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x < 32) {
            double v = (threadIdx.x < LL_THREADS / 32) ? red[threadIdx.x] : 0.0;
            v = warp_sum_d(v);
            if (threadIdx.x == 0 && row < R) ll_store(lp, lp_lo, row, (double)scale * (v + row_const));
        }
    }
}

// elementwise backward: grad[r,j] = g[r] * c * dl/dr (* r(1-r) through the sigmoid).  Flat grid-stride over float4.
template <int LIK>
__global__ void __launch_bounds__(LL_THREADS)
loglik_bwd_kernel(const float* __restrict__ glp, const float4* __restrict__ recon, const float4* __restrict__ x,
                  float4* __restrict__ grad, long total4, int B, int D4, int x_div, float scale,
                  int through_sigmoid) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long)gridDim.x * blockDim.x) {
        const long row = i / D4;
        const int j = (int)(i - row * D4);
        const float g = __ldg(glp + row) * scale;
        const float4 r = ld_stream4(recon + i);
        const float4 v = __ldg(x + (long)((row / x_div) % B) * D4 + j);
        float4 o;
        o.x = g * ll_grad<LIK>(v.x, r.x);
        o.y = g * ll_grad<LIK>(v.y, r.y);
        o.z = g * ll_grad<LIK>(v.z, r.z);
        o.w = g * ll_grad<LIK>(v.w, r.w);
        if (through_sigmoid) {
            o.x *= r.x * (1.f - r.x);
            o.y *= r.y * (1.f - r.y);
            o.z *= r.z * (1.f - r.z);
            o.w *= r.w * (1.f - r.w);
        }
        st_stream4(grad + i, o);
    }
}

template <int LIK>
static int launch_fwd(const float* recon, const float* x, float* lp, float* lp_lo, int R, int B, int D, int x_div,
                      float c, cudaStream_t st) {
    // the reference adds the fp32 constant to every pixel term; D * fp32(const) in fp64 is that sum without rounding
    const double row_const = (LIK == MMVAE_LIK_NORMAL) ? -(double)D * (double)MMVAE_LOG_SQRT_2PI
                           : (LIK == MMVAE_LIK_LAPLACE) ? -(double)D * (double)MMVAE_LOG2 : 0.0;
    const int D4 = D / 4;
    // one warp per row keeps 4 x 128-bit loads per thread in flight and needs no block barrier; one block per row only pays when
    // there are too few rows to fill the machine with warps (measured at R = 7680, D = 3072: 27 us per block-row, see below)
    if (D4 <= 512 || (long)R >= 8L * MMVAE_NUM_SMS) {
        const int rpb = LL_THREADS / 32;
        loglik_fwd_kernel<LIK, 32><<<ceil_div_i(R, rpb), LL_THREADS, 0, st>>>(
            (const float4*)recon, (const float4*)x, lp, lp_lo, R, B, D4, x_div, c, row_const);
    } else {
        loglik_fwd_kernel<LIK, LL_THREADS><<<R, LL_THREADS, 0, st>>>(
            (const float4*)recon, (const float4*)x, lp, lp_lo, R, B, D4, x_div, c, row_const);
    }
    return 0;
}

extern "C" int mmvae_loglik_fwd(const float* recon, const float* x, float* lp, float* lp_lo, int R, int B, int D,
                                int x_div, int lik, float lik_scaling, void* stream) {
    MMVAE_REQUIRE(recon && x && lp, "loglik_fwd: null pointer");
    MMVAE_REQUIRE(R > 0 && B > 0 && D > 0 && D % 4 == 0 && x_div > 0, "loglik_fwd: bad sizes R=%d B=%d D=%d", R, B, D);
    MMVAE_REQUIRE(((uintptr_t)recon & 15) == 0 && ((uintptr_t)x & 15) == 0, "loglik_fwd: pointers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    if (lik == MMVAE_LIK_NORMAL) launch_fwd<MMVAE_LIK_NORMAL>(recon, x, lp, lp_lo, R, B, D, x_div, lik_scaling, st);
    else if (lik == MMVAE_LIK_LAPLACE) launch_fwd<MMVAE_LIK_LAPLACE>(recon, x, lp, lp_lo, R, B, D, x_div, lik_scaling, st);
    else if (lik == MMVAE_LIK_BERNOULLI) launch_fwd<MMVAE_LIK_BERNOULLI>(recon, x, lp, lp_lo, R, B, D, x_div, lik_scaling, st);
    else if (lik == MMVAE_LIK_NEG_MSE) launch_fwd<MMVAE_LIK_NEG_MSE>(recon, x, lp, lp_lo, R, B, D, x_div, lik_scaling, st);
    else if (lik == MMVAE_LIK_NEG_L1) launch_fwd<MMVAE_LIK_NEG_L1>(recon, x, lp, lp_lo, R, B, D, x_div, lik_scaling, st);
    else if (lik == MMVAE_LIK_NEG_BCE) launch_fwd<MMVAE_LIK_NEG_BCE>(recon, x, lp, lp_lo, R, B, D, x_div, lik_scaling, st);
    else return mmvae_set_error(1, "loglik_fwd: bad likelihood %d", lik);
    MMVAE_LAUNCH_CHECK("loglik_fwd");
    return 0;
}

extern "C" int mmvae_loglik_bwd(const float* grad_lp, const float* recon, const float* x, float* grad, int R, int B,
                                int D, int x_div, int lik, float lik_scaling, int through_sigmoid, void* stream) {
    MMVAE_REQUIRE(grad_lp && recon && x && grad, "loglik_bwd: null pointer");
    MMVAE_REQUIRE(R > 0 && B > 0 && D > 0 && D % 4 == 0 && x_div > 0, "loglik_bwd: bad sizes");
    MMVAE_REQUIRE(((uintptr_t)recon & 15) == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)grad & 15) == 0,
                  "loglik_bwd: pointers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const long total4 = (long)R * (D / 4);
    const int grid = (int)min((long)MMVAE_NUM_SMS * 8, (total4 + LL_THREADS - 1) / LL_THREADS);
    const int D4 = D / 4;
#define LL_BWD(LIKV)                                                                                              \
    loglik_bwd_kernel<LIKV><<<grid, LL_THREADS, 0, st>>>(grad_lp, (const float4*)recon, (const float4*)x, (float4*)grad, \
                                                         total4, B, D4, x_div, lik_scaling, through_sigmoid)
    if (lik == MMVAE_LIK_NORMAL) LL_BWD(MMVAE_LIK_NORMAL);
    else if (lik == MMVAE_LIK_LAPLACE) LL_BWD(MMVAE_LIK_LAPLACE);
    else if (lik == MMVAE_LIK_BERNOULLI) LL_BWD(MMVAE_LIK_BERNOULLI);
    else if (lik == MMVAE_LIK_NEG_MSE) LL_BWD(MMVAE_LIK_NEG_MSE);
    else if (lik == MMVAE_LIK_NEG_L1) LL_BWD(MMVAE_LIK_NEG_L1);
    else if (lik == MMVAE_LIK_NEG_BCE) LL_BWD(MMVAE_LIK_NEG_BCE);
#undef LL_BWD
    else return mmvae_set_error(1, "loglik_bwd: bad likelihood %d", lik);
    MMVAE_LAUNCH_CHECK("loglik_bwd");
    return 0;
}
