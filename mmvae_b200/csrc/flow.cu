// Elementwise pieces of the MADE / MAF / IAF flows and of the JMVAE-NF regulariser (kernel (4)'s non-GEMM half).
// The masked-linear contractions themselves run through mmvae_gemm (tcgen05 / SIMT); these kernels are the affine
// couplings between them, fused with the log|det J| accumulation, and the Gaussian log-density row reductions.
// Reference ops replaced: maf_model.py:62-74 (x=(x-mu)*exp(-lv); ladj-=sum lv; flip), :91-104 (x[:,i]=y[:,i]*exp(lv[:,i])
// +mu[:,i]; ladj+=lv[:,i]); iaf_model.py:60-78,94-100 (same pair, roles swapped); jmvae_nf.py:66-69,131-134
// (log N(z0;mu,exp(lv))), objectives.py:209-211 (closed-form KL to the prior), qz_xy.log_prob(z).sum() (jmvae_nf.py:69).
#include "common.cuh"

#define FLOW_MAX_L 64
#define MMVAE_LOG_2PI 1.8378770664093453f

// one thread per row r.  h = [mu | lv] (R x 2L).  mode 0 "normalise": u = (x-mu)*exp(-lv), ladj -= lv;
// mode 1 "generate": u = x*exp(lv)+mu, ladj += lv.  col < 0: all columns (out[r, flip?L-1-l:l] = u_l);
// col >= 0: out = carry with column `col` replaced by u_col (one step of the sequential direction).
__global__ void flow_affine_fwd_kernel(const float* __restrict__ x, const float* __restrict__ h,
                                       const float* __restrict__ carry, const float* __restrict__ ladj_in,
                                       float* __restrict__ out, float* __restrict__ ladj_out, int R, int L, int mode,
                                       int col, int flip_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float* xr = x + (long)r * L;
    const float* hr = h + (long)r * 2 * L;
    float* orow = out + (long)r * L;
    float acc = 0.f;
    if (col < 0) {
        for (int l = 0; l < L; ++l) {
            const float mu = hr[l], lv = hr[L + l];
            const float u = mode == 0 ? (xr[l] - mu) * expf(-lv) : xr[l] * expf(lv) + mu;
            orow[flip_out ? L - 1 - l : l] = u;
            acc += lv;
        }
    } else {
        const float* cr = carry + (long)r * L;
        for (int l = 0; l < L; ++l) orow[l] = cr[l];
        const float mu = hr[col], lv = hr[L + col];
        orow[col] = mode == 0 ? (xr[col] - mu) * expf(-lv) : xr[col] * expf(lv) + mu;
        acc = lv;
    }
    ladj_out[r] = ladj_in[r] + (mode == 0 ? -acc : acc);
}

__global__ void flow_affine_bwd_kernel(const float* __restrict__ g_out, const float* __restrict__ g_ladj,
                                       const float* __restrict__ x, const float* __restrict__ h,
                                       float* __restrict__ g_x, float* __restrict__ g_h, float* __restrict__ g_carry,
                                       int R, int L, int mode, int col, int flip_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float* xr = x + (long)r * L;
    const float* hr = h + (long)r * 2 * L;
    const float* go = g_out + (long)r * L;
    float* gx = g_x + (long)r * L;
    float* gh = g_h + (long)r * 2 * L;
    const float gl = g_ladj ? g_ladj[r] : 0.f;
    const int l0 = col < 0 ? 0 : col, l1 = col < 0 ? L : col + 1;
    if (col >= 0) {
        float* gc = g_carry + (long)r * L;
        for (int l = 0; l < L; ++l) { gc[l] = (l == col) ? 0.f : go[l]; gx[l] = 0.f; gh[l] = 0.f; gh[L + l] = 0.f; }
    }
    for (int l = l0; l < l1; ++l) {
        const float mu = hr[l], lv = hr[L + l];
        const float gu = go[(col < 0 && flip_out) ? L - 1 - l : l];
        if (mode == 0) {
            const float e = expf(-lv);
            gx[l] = gu * e;
            gh[l] = -gu * e;
            gh[L + l] = -gu * (xr[l] - mu) * e - gl;
        } else {
            const float e = expf(lv);
            gx[l] = gu * e;
            gh[l] = gu;
            gh[L + l] = gu * xr[l] * e + gl;
        }
    }
}

__global__ void flip_cols_kernel(const float* __restrict__ x, float* __restrict__ out, long n, int L) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long r = i / L;
        const int l = (int)(i - r * L);
        out[r * L + (L - 1 - l)] = x[i];
    }
}

// Gaussian row reductions.  kind 0: sum_l log N(z; mu, s) with s = std; kind 1: s = log_var, the reference's
// -0.5*(lv + log 2pi + (z-mu)^2/exp(lv)) (jmvae_nf.py:67); kind 2: closed-form KL(N(mu,s^2)||N(0,1)) written as the
// reference does, -0.5*sum(1 + lv - mu^2 - exp(lv)) with lv = 2 log s (objectives.py:208-211; z unused).
// mu/s rows are shared by `div` consecutive z rows (evaluator: z [B,K,L] against mu [B,L]).
__global__ void gauss_rows_fwd_kernel(const float* __restrict__ z, const float* __restrict__ mu,
                                      const float* __restrict__ s, float* __restrict__ out, int R, int L, int kind,
                                      int div) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const long pr = (long)(r / div) * L;
    float acc = 0.f;
    for (int l = 0; l < L; ++l) {
        const float m = mu[pr + l], sv = s[pr + l];
        if (kind == 0) {
            const float d = z[(long)r * L + l] - m;
            acc += -(d * d) / (2.f * sv * sv) - logf(sv) - MMVAE_LOG_SQRT_2PI;
        } else if (kind == 1) {
            const float d = z[(long)r * L + l] - m;
            acc += -0.5f * (sv + MMVAE_LOG_2PI + d * d / expf(sv));
        } else {
            const float lv = 2.f * logf(sv);
            acc += -0.5f * (1.f + lv - m * m - expf(lv));
        }
    }
    out[r] = acc;
}

// g: [R].  grads are per z row; when div > 1 the caller reduces g_mu/g_s over the shared rows (evaluator paths carry
// no gradient, so only div == 1 is differentiated).
__global__ void gauss_rows_bwd_kernel(const float* __restrict__ g, const float* __restrict__ z,
                                      const float* __restrict__ mu, const float* __restrict__ s,
                                      float* __restrict__ g_z, float* __restrict__ g_mu, float* __restrict__ g_s, int R,
                                      int L, int kind) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= R) return;
    const float gr = g[r];
    for (int l = 0; l < L; ++l) {
        const long o = (long)r * L + l;
        const float m = mu[o], sv = s[o];
        if (kind == 0) {
            const float d = z[o] - m;
            const float dz = -d / (sv * sv);
            g_z[o] = gr * dz;
            g_mu[o] = -gr * dz;
            g_s[o] = gr * (d * d / (sv * sv * sv) - 1.f / sv);
        } else if (kind == 1) {
            const float d = z[o] - m, e = expf(-sv);
            g_z[o] = -gr * d * e;
            g_mu[o] = gr * d * e;
            g_s[o] = gr * (-0.5f + 0.5f * d * d * e);
        } else {
            if (g_z) g_z[o] = 0.f;
            g_mu[o] = gr * m;
            g_s[o] = gr * (-1.f / sv + sv);              // d/ds of -0.5*(1 + 2 log s - mu^2 - s^2)
        }
    }
}

extern "C" int mmvae_flow_affine_fwd(const float* x, const float* h, const float* carry, const float* ladj_in,
                                     float* out, float* ladj_out, int R, int L, int mode, int col, int flip_out,
                                     void* stream) {
    MMVAE_REQUIRE(x && h && ladj_in && out && ladj_out, "flow_affine_fwd: null pointer");
    MMVAE_REQUIRE(R > 0 && L > 0 && L <= FLOW_MAX_L && (mode == 0 || mode == 1) && col < L, "flow_affine_fwd: bad sizes R=%d L=%d col=%d", R, L, col);
    MMVAE_REQUIRE(col < 0 || carry, "flow_affine_fwd: a sequential step needs `carry`");
    flow_affine_fwd_kernel<<<ceil_div_i(R, 128), 128, 0, (cudaStream_t)stream>>>(x, h, carry, ladj_in, out, ladj_out, R, L,
                                                                                 mode, col, flip_out);
    MMVAE_LAUNCH_CHECK("flow_affine_fwd");
    return 0;
}

extern "C" int mmvae_flow_affine_bwd(const float* g_out, const float* g_ladj, const float* x, const float* h,
                                     float* g_x, float* g_h, float* g_carry, int R, int L, int mode, int col,
                                     int flip_out, void* stream) {
    MMVAE_REQUIRE(g_out && x && h && g_x && g_h, "flow_affine_bwd: null pointer");
    MMVAE_REQUIRE(R > 0 && L > 0 && L <= FLOW_MAX_L && (mode == 0 || mode == 1) && col < L, "flow_affine_bwd: bad sizes");
    MMVAE_REQUIRE(col < 0 || g_carry, "flow_affine_bwd: a sequential step needs `g_carry`");
    flow_affine_bwd_kernel<<<ceil_div_i(R, 128), 128, 0, (cudaStream_t)stream>>>(g_out, g_ladj, x, h, g_x, g_h, g_carry, R,
                                                                                 L, mode, col, flip_out);
    MMVAE_LAUNCH_CHECK("flow_affine_bwd");
    return 0;
}

extern "C" int mmvae_flip_cols(const float* x, float* out, int R, int L, void* stream) {
    MMVAE_REQUIRE(x && out && R > 0 && L > 0 && x != out, "flip_cols: bad arguments");
    const long n = (long)R * L;
    const int grid = (int)min((long)MMVAE_NUM_SMS * 8, (n + 255) / 256);
    flip_cols_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, out, n, L);
    MMVAE_LAUNCH_CHECK("flip_cols");
    return 0;
}

extern "C" int mmvae_gauss_rows_fwd(const float* z, const float* mu, const float* s, float* out, int R, int L,
                                    int kind, int div, void* stream) {
    MMVAE_REQUIRE(mu && s && out && (z || kind == 2), "gauss_rows_fwd: null pointer");
    MMVAE_REQUIRE(R > 0 && L > 0 && kind >= 0 && kind <= 2 && div >= 1, "gauss_rows_fwd: bad arguments");
    gauss_rows_fwd_kernel<<<ceil_div_i(R, 128), 128, 0, (cudaStream_t)stream>>>(z, mu, s, out, R, L, kind, div);
    MMVAE_LAUNCH_CHECK("gauss_rows_fwd");
    return 0;
}

extern "C" int mmvae_gauss_rows_bwd(const float* g, const float* z, const float* mu, const float* s, float* g_z,
                                    float* g_mu, float* g_s, int R, int L, int kind, void* stream) {
    MMVAE_REQUIRE(g && mu && s && g_mu && g_s && ((z && g_z) || kind == 2), "gauss_rows_bwd: null pointer");
    MMVAE_REQUIRE(R > 0 && L > 0 && kind >= 0 && kind <= 2, "gauss_rows_bwd: bad arguments");
    gauss_rows_bwd_kernel<<<ceil_div_i(R, 128), 128, 0, (cudaStream_t)stream>>>(g, z, mu, s, g_z, g_mu, g_s, R, L, kind);
    MMVAE_LAUNCH_CHECK("gauss_rows_bwd");
    return 0;
}
