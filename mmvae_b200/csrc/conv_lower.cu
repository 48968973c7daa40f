// Conv2d / ConvTranspose2d lowering around the GEMM (encoders.py:80-136: 4x4 stride-2 pad-1 stacks, plus the
// 1x1<->4x4 ends).  Activations are NHWC internally so every lowered matrix is K-contiguous; generic element
// strides (sn,sc,sy,sx) describe the tensor so NCHW images at the boundary need no separate transpose.
//   conv  fwd : col = im2col(x)            ; y = col * Wp^T            (Wp[Cout,(ky,kx,ci)])
//   conv  bwd : dcol = dy * Wp             ; dx = col2im(dcol)         ; dWp = dy^T * col
//   convT fwd : col = x * Wp               ; y = col2im(col)+bias,act  (Wp[Cin,(ky,kx,co)])
//   convT bwd : dcol = im2col(dy)          ; dx = dcol * Wp^T          ; dWp = x^T * dcol
// col2im is written in gather form (each output element sums its <= ceil(k/s)^2 taps): deterministic, no atomics.
#include "common.cuh"

__device__ __forceinline__ float act_fwd(float v, int act) {
    if (act == MMVAE_ACT_RELU) return fmaxf(v, 0.f);
    if (act == MMVAE_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
    return v;
}

// w: [C0][C1][kk] (torch layout: conv (Cout,Cin,kh*kw), convT (Cin,Cout,kh*kw)) <-> wp: [C0][kk][C1]
__global__ void permute_weight_kernel(const float* __restrict__ src, float* __restrict__ dst, int C0, int C1, int kk,
                                      int to_permuted) {
    const long n = (long)C0 * C1 * kk;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        // i indexes dst
        if (to_permuted) {
            const int c1 = (int)(i % C1), t = (int)((i / C1) % kk), c0 = (int)(i / ((long)C1 * kk));
            dst[i] = src[((long)c0 * C1 + c1) * kk + t];
        } else {
            const int t = (int)(i % kk), c1 = (int)((i / kk) % C1), c0 = (int)(i / ((long)C1 * kk));
            dst[i] = src[((long)c0 * kk + t) * C1 + c1];
        }
    }
}

// col[(n,oy,ox)][(ky,kx,c)] = x[n, c, oy*s-p+ky, ox*s-p+kx]  (0 outside).  VEC: 4 channels per thread (needs sc==1, C%4==0).
template <bool VEC>
__global__ void __launch_bounds__(256)
im2col_kernel(const float* __restrict__ x, float* __restrict__ col, int C, int H, int W, int Ho, int Wo, int k, int s,
              int p, long sn, long sc, long sy, long sx, long total) {
    const int CV = VEC ? C / 4 : C;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % CV);
        long r = i / CV;
        const int kx = (int)(r % k); r /= k;
        const int ky = (int)(r % k); r /= k;
        const int ox = (int)(r % Wo); r /= Wo;
        const int oy = (int)(r % Ho);
        const long n = r / Ho;
        const int iy = oy * s - p + ky, ix = ox * s - p + kx;
        const bool in = (iy >= 0 && iy < H && ix >= 0 && ix < W);
        if (VEC) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (in) v = *reinterpret_cast<const float4*>(x + n * sn + iy * sy + ix * sx + c * 4);
            reinterpret_cast<float4*>(col)[i] = v;
        } else {
            col[i] = in ? x[n * sn + (long)c * sc + iy * sy + ix * sx] : 0.f;
        }
    }
}

// y[n,c,oy,ox] = act(bias[c] + sum_{ky,kx} col[(n,iy,ix)][(ky,kx,c)]),  iy = (oy+p-ky)/s when divisible and in range.
// C_FAST: thread index runs over c fastest (NHWC output, coalesced both sides); else over ox fastest (NCHW output).
template <bool VEC, bool C_FAST>
__global__ void __launch_bounds__(256)
col2im_kernel(const float* __restrict__ col, const float* __restrict__ bias, float* __restrict__ y, int C, int H,
              int W, int Ho, int Wo, int k, int s, int p, long sn, long sc, long sy, long sx, int act, long total) {
    const int CV = VEC ? C / 4 : C;
    const long rowlen = (long)k * k * C;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        int c, ox, oy;
        long n;
        if (C_FAST) {
            c = (int)(i % CV);
            long r = i / CV;
            ox = (int)(r % Wo); r /= Wo;
            oy = (int)(r % Ho);
            n = r / Ho;
        } else {
            ox = (int)(i % Wo);
            long r = i / Wo;
            oy = (int)(r % Ho); r /= Ho;
            c = (int)(r % CV);
            n = r / CV;
        }
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int ky = 0; ky < k; ++ky) {
            const int ty = oy + p - ky;
            if (ty < 0 || ty % s) continue;
            const int iy = ty / s;
            if (iy >= H) continue;
            for (int kx = 0; kx < k; ++kx) {
                const int tx = ox + p - kx;
                if (tx < 0 || tx % s) continue;
                const int ix = tx / s;
                if (ix >= W) continue;
                const float* src = col + ((n * H + iy) * W + ix) * rowlen + (long)(ky * k + kx) * C;
                if (VEC) {
                    const float4 v = *reinterpret_cast<const float4*>(src + c * 4);
                    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                } else {
                    acc.x += src[c];
                }
            }
        }
        if (VEC) {
            if (bias) {
                const float4 bv = *reinterpret_cast<const float4*>(bias + c * 4);
                acc.x += bv.x; acc.y += bv.y; acc.z += bv.z; acc.w += bv.w;
            }
            acc.x = act_fwd(acc.x, act); acc.y = act_fwd(acc.y, act);
            acc.z = act_fwd(acc.z, act); acc.w = act_fwd(acc.w, act);
            *reinterpret_cast<float4*>(y + n * sn + oy * sy + ox * sx + c * 4) = acc;
        } else {
            const float v = acc.x + (bias ? bias[c] : 0.f);
            y[n * sn + (long)c * sc + oy * sy + ox * sx] = act_fwd(v, act);
        }
    }
}

static inline int grid_for(long total) {
    long g = (total + 255) / 256;
    const long cap = (long)MMVAE_NUM_SMS * 16;
    return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

extern "C" int mmvae_permute_weight(const float* w, float* wp, int C0, int C1, int kk, int to_permuted, void* stream) {
    MMVAE_REQUIRE(w && wp && C0 > 0 && C1 > 0 && kk > 0, "permute_weight: bad arguments");
    const long n = (long)C0 * C1 * kk;
    permute_weight_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(w, wp, C0, C1, kk, to_permuted);
    MMVAE_LAUNCH_CHECK("permute_weight");
    return 0;
}

extern "C" int mmvae_im2col(const float* x, float* col, int N, int C, int H, int W, int k, int s, int p, long sn,
                            long sc, long sy, long sx, void* stream) {
    MMVAE_REQUIRE(x && col && N > 0 && C > 0 && H > 0 && W > 0 && k > 0 && s > 0 && p >= 0, "im2col: bad arguments");
    MMVAE_REQUIRE(H + 2 * p >= k && W + 2 * p >= k, "im2col: kernel larger than padded input");
    const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
    const bool vec = (sc == 1) && (C % 4 == 0) && (sn % 4 == 0) && (sy % 4 == 0) && (sx % 4 == 0) &&
                     (((uintptr_t)x & 15) == 0) && (((uintptr_t)col & 15) == 0);
    cudaStream_t st = (cudaStream_t)stream;
    if (vec) {
        const long total = (long)N * Ho * Wo * k * k * (C / 4);
        im2col_kernel<true><<<grid_for(total), 256, 0, st>>>(x, col, C, H, W, Ho, Wo, k, s, p, sn, sc, sy, sx, total);
    } else {
        const long total = (long)N * Ho * Wo * k * k * C;
        im2col_kernel<false><<<grid_for(total), 256, 0, st>>>(x, col, C, H, W, Ho, Wo, k, s, p, sn, sc, sy, sx, total);
    }
    MMVAE_LAUNCH_CHECK("im2col");
    return 0;
}

extern "C" int mmvae_col2im(const float* col, const float* bias, float* y, int N, int C, int H, int W, int k, int s,
                            int p, long sn, long sc, long sy, long sx, int act, void* stream) {
    MMVAE_REQUIRE(col && y && N > 0 && C > 0 && H > 0 && W > 0 && k > 0 && s > 0 && p >= 0, "col2im: bad arguments");
    const int Ho = (H - 1) * s - 2 * p + k, Wo = (W - 1) * s - 2 * p + k;
    MMVAE_REQUIRE(Ho > 0 && Wo > 0, "col2im: empty output");
    const bool vec = (sc == 1) && (C % 4 == 0) && (sn % 4 == 0) && (sy % 4 == 0) && (sx % 4 == 0) &&
                     (((uintptr_t)y & 15) == 0) && (((uintptr_t)col & 15) == 0) &&
                     (!bias || ((uintptr_t)bias & 15) == 0);
    cudaStream_t st = (cudaStream_t)stream;
    const long total = (long)N * Ho * Wo * (vec ? C / 4 : C);
    if (vec) col2im_kernel<true, true><<<grid_for(total), 256, 0, st>>>(col, bias, y, C, H, W, Ho, Wo, k, s, p, sn, sc, sy, sx, act, total);
    else if (sc == 1) col2im_kernel<false, true><<<grid_for(total), 256, 0, st>>>(col, bias, y, C, H, W, Ho, Wo, k, s, p, sn, sc, sy, sx, act, total);
    else col2im_kernel<false, false><<<grid_for(total), 256, 0, st>>>(col, bias, y, C, H, W, Ho, Wo, k, s, p, sn, sc, sy, sx, act, total);
    MMVAE_LAUNCH_CHECK("col2im");
    return 0;
}
