// ConvTranspose2d(Cin -> Cout <= 4, kernel 4, stride 2, pad 1): the image-producing last layer of Decoder_VAE_SVHN
// (encoders.py:125-126: ConvT(32 -> 3) + Sigmoid, 16x16 -> 32x32) as DIRECT kernels, forward / dgrad / wgrad.
// With Cout = 3 this layer is 8.5 % of the decoder's FLOPs but, lowered to a GEMM, the widest memory hog of the step
// (a 377 MB col buffer written and re-read in each direction, N = 3 padded to a tensor-core tile): it is HBM-bound
// integer-stride gather work, so it runs on the FP32 pipes with the image staged in shared memory and touches HBM once:
//   fwd  : reads x (N,H,W,Cin) NHWC once, writes y (N,Cout,2H,2W) NCHW once, bias + activation fused;
//   dgrad: reads g once (+ y when the activation derivative is fused here), writes gx once, ReLU mask of the layer
//          below fused (gx is then the gradient w.r.t. that layer's PRE-activation);
//   wgrad: reads x and g once, per-CTA register accumulators across images, deterministic two-stage reduction, the
//          bias gradient folded in.
// Output pixel (2i+py, 2j+px) receives taps ky in {1,3} (py=0) / {0,2} (py=1) from input rows i, i-1 / i+1, i.
#include "common.cuh"
#include "sm100_ptx.cuh"

#define CS_THREADS 256
#define CS_PAD 4                      // floats of padding per input pixel row in smem: 128-bit loads stay conflict-free

__device__ __forceinline__ float cs_act(float v, int act) {
    if (act == MMVAE_ACT_RELU) return fmaxf(v, 0.f);
    if (act == MMVAE_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
    return v;
}
// derivative of the activation expressed with its OUTPUT y
__device__ __forceinline__ float cs_dact(float y, int act) {
    if (act == MMVAE_ACT_RELU) return y > 0.f ? 1.f : 0.f;
    if (act == MMVAE_ACT_SIGMOID) return y * (1.f - y);
    return 1.f;
}

// ------------------------------------------------------------------------------------------------------ forward
// Persistent CTAs of 128 threads (two per SM) over images.  The NHWC image lands by one bulk async copy; a short
// transform pass spreads it into a zero-bordered copy with a 36-float pixel pitch (conflict-free 128-bit reads), after
// which the raw buffer is free and the NEXT image's copy is issued, to land while this image is computed.
// One thread = two vertically adjacent input pixels = a 4 x 2 block of output pixels x COUT: every weight vector is
// fetched once for both pixels and the channel loop runs on packed FFMA2 (pairs of channels).
// smem: xs[(H+2)*(W+2)][CIN+PAD], raw[H*W][CIN], ws[16][COUT][CIN], bs[COUT], one mbarrier.
#define CSF_THREADS 128
__host__ __device__ inline size_t csf_smem_bytes(int H, int W, int Cin, int Cout) {
    return ((size_t)(H + 2) * (W + 2) * (Cin + CS_PAD) + (size_t)H * W * Cin + 16 * (size_t)Cout * Cin + 4) * 4 + 16;
}
template <int COUT, int CIN, int HT, int WT>
__global__ void __launch_bounds__(CSF_THREADS)
convt_small_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                       float* __restrict__ y, int N, int Hrt, int Wrt, int act) {
    extern __shared__ float4 smem4[];
    const int H = HT ? HT : Hrt, W = WT ? WT : Wrt, HW = H * W, Wp = W + 2;
    constexpr int xstride = CIN + CS_PAD;
    float* xs = (float*)smem4;                                        // [(H+2)*(W+2)][xstride], zero border
    float* raw = xs + (H + 2) * Wp * xstride;                         // [HW][CIN]
    float* ws = raw + HW * CIN;                                       // [16][COUT][CIN]
    float* bs = ws + 16 * COUT * CIN;
    uint64_t* full = (uint64_t*)(bs + 4);
    if (threadIdx.x == 0) {
        mbar_init(full, 1);
        mbar_fence_init();
        if ((int)blockIdx.x < N) {
            mbar_arrive_expect_tx(full, HW * CIN * 4);
            bulk_load_1d(raw, x + (long)blockIdx.x * HW * CIN, HW * CIN * 4, full);
        }
    }
    for (int i = threadIdx.x; i < (H + 2) * Wp * xstride; i += CSF_THREADS) xs[i] = 0.f;       // the border stays zero
    for (int i = threadIdx.x; i < 16 * COUT * CIN; i += CSF_THREADS) {     // torch layout w[ci][co][tap] -> ws[tap][co][ci]
        const int ci = i % CIN, co = (i / CIN) % COUT, tap = i / (CIN * COUT);
        ws[i] = w[((long)ci * COUT + co) * 16 + tap];
    }
    if (threadIdx.x < COUT) bs[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
    const int H2 = 2 * H, W2 = 2 * W;
    int it = 0;
    for (int n = blockIdx.x; n < N; n += gridDim.x, ++it) {
        __syncthreads();                                              // xs of image n-1 fully consumed
        mbar_wait(full, it & 1);
        for (int i = threadIdx.x; i < HW * (CIN / 4); i += CSF_THREADS) {
            const int p = i / (CIN / 4), c = i - p * (CIN / 4), pi = p / W, pj = p - pi * W;
            *(float4*)(xs + ((pi + 1) * Wp + pj + 1) * xstride + 4 * c) = *(const float4*)(raw + 4 * i);
        }
        __syncthreads();                                              // xs ready, raw free
        if (threadIdx.x == 0 && n + (int)gridDim.x < N) {
            mbar_arrive_expect_tx(full, HW * CIN * 4);
            bulk_load_1d(raw, x + (long)(n + gridDim.x) * HW * CIN, HW * CIN * 4, full);
        }
        for (int q = threadIdx.x; q < HW / 2; q += CSF_THREADS) {
            const int r = q / W, j = q - r * W, i0 = 2 * r;           // input pixels (i0, j) and (i0+1, j)
            float2 acc[2][2][2][COUT];                                // [pixel][py][px][co], channel-pair partial sums
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b)
#pragma unroll
                        for (int c = 0; c < COUT; ++c) acc[u][a][b][c] = make_float2(0.f, 0.f);
            const float* xw = xs + (i0 * Wp + j) * xstride;           // bordered (i0-1+1, j-1+1): top-left of the 4 x 3 window
#pragma unroll 1
            for (int c = 0; c < CIN; c += 4) {
                float4 xv[4][3];
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 3; ++b) xv[a][b] = *(const float4*)(xw + (a * Wp + b) * xstride + c);
#pragma unroll
                for (int ky = 0; ky < 4; ++ky) {
                    const int py = (ky + 1) & 1, dy = (py + 1 - ky) / 2;          // compile-time after unrolling
                    // the 4*COUT weight vectors of this kernel row are fetched as one batch: with two warps per scheduler
                    // a load issued right before its use costs a full shared-memory round trip each time
                    float4 wv[4][COUT];
#pragma unroll
                    for (int kx = 0; kx < 4; ++kx)
#pragma unroll
                        for (int co = 0; co < COUT; ++co)
                            wv[kx][co] = *(const float4*)(ws + ((ky * 4 + kx) * COUT + co) * CIN + c);
#pragma unroll
                    for (int kx = 0; kx < 4; ++kx) {
                        const int px = (kx + 1) & 1, dx = (px + 1 - kx) / 2;
#pragma unroll
                        for (int co = 0; co < COUT; ++co) {
                            const float2 w0 = make_float2(wv[kx][co].x, wv[kx][co].y), w1 = make_float2(wv[kx][co].z, wv[kx][co].w);
#pragma unroll
                            for (int u = 0; u < 2; ++u) {
                                const float4 v = xv[u + dy + 1][dx + 1];
                                float2 t = acc[u][py][px][co];
                                t = __ffma2_rn(make_float2(v.x, v.y), w0, t);
                                t = __ffma2_rn(make_float2(v.z, v.w), w1, t);
                                acc[u][py][px][co] = t;
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int co = 0; co < COUT; ++co)
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int py = 0; py < 2; ++py) {
                        const float2 a0 = acc[u][py][0][co], a1 = acc[u][py][1][co];
                        const float2 o = make_float2(cs_act(a0.x + a0.y + bs[co], act), cs_act(a1.x + a1.y + bs[co], act));
                        *(float2*)(y + (((long)n * COUT + co) * H2 + 2 * (i0 + u) + py) * W2 + 2 * j) = o;
                    }
        }
    }
}

// stage g (N,Cout,2H,2W) of image n into gs[COUT][2H+2][2W+2] with a zero border, optionally times act'(y)
template <int COUT, int NT = CS_THREADS>
__device__ __forceinline__ void cs_stage_g(float* gs, const float* __restrict__ g, const float* __restrict__ yact,
                                           int n, int H2, int W2, int act, float* bias_acc) {
    const int gw = W2 + 2, plane = (H2 + 2) * gw;
    for (int i = threadIdx.x; i < COUT * plane; i += NT) gs[i] = 0.f;
    __syncthreads();
    const long base = (long)n * COUT * H2 * W2;
    for (int i = threadIdx.x; i < COUT * H2 * W2; i += NT) {
        const int xx = i % W2, yy = (i / W2) % H2, co = i / (W2 * H2);
        float v = g[base + i];
        if (yact) v *= cs_dact(yact[base + i], act);
        gs[co * plane + (yy + 1) * gw + xx + 1] = v;
        if (bias_acc) {
#pragma unroll
            for (int c = 0; c < COUT; ++c) bias_acc[c] += (c == co) ? v : 0.f;
        }
    }
}

// -------------------------------------------------------------------------------------------------------- dgrad
// gx[n,i,j,ci] = sum_{ky,kx,co} g[n,co,2i-1+ky,2j-1+kx] * w[ci,co,ky,kx]   (* relu'(x_mask))
// Persistent, one CTA of 512 threads per SM.  Each lane owns TWO input channels and keeps their 2 x COUT x 16 weights
// in registers; the two half-warps work on two horizontally adjacent pixels, so per pixel the COUT*16 gradient values
// come from COUT*4 broadcast 128-bit shared loads and feed 2*COUT*16 FFMAs per lane, and the mask read / gx write of
// one pixel is one contiguous 128-byte row segment per half-warp.  The inputs of image n+1 (g, y for the fused
// activation derivative, the ReLU mask source x) are fetched by bulk asynchronous copies into a second set of
// shared-memory buffers while image n is computed: with the loads issued at their point of use (first and second
// versions) every warp sat out a DRAM round trip per pixel and the FMA pipe was 27 % busy.
// The staged gradient image is kept twice, the second copy shifted by two floats, so that the 4-float window of both
// even and odd columns starts on a 16-byte boundary.
#define CSD_THREADS 512
#define CSD_PITCH(W2) (((W2) + 2 + 2 + 3) & ~3)
__host__ __device__ inline size_t csd_smem_bytes(int H, int W, int Cin, int Cout, bool has_y, bool has_mask) {
    const size_t gs = 2 * (size_t)Cout * (2 * H + 2) * CSD_PITCH(2 * W) * 4;
    const size_t raw = (size_t)Cout * 4 * H * W * 4;
    return gs + 2 * raw * (has_y ? 2 : 1) + (has_mask ? 2 * (size_t)H * W * Cin * 4 : 0) + 64;
}
// HT, WT > 0: compile-time image size (every shared-memory offset an immediate, divisions become shifts); 0: run time
template <int COUT, int CIN, int HT, int WT>
__global__ void __launch_bounds__(CSD_THREADS, 1)
convt_small_dgrad_kernel(const float* __restrict__ g, const float* __restrict__ yact, const float* __restrict__ w,
                         const float* __restrict__ x_mask, float* __restrict__ gx, int N, int Hrt, int Wrt, int act) {
    static_assert(CIN == 32, "one lane per channel pair");
    extern __shared__ float4 smem4[];
    const int H = HT ? HT : Hrt, W = WT ? WT : Wrt;
    const int H2 = 2 * H, W2 = 2 * W, pitch = CSD_PITCH(W2), plane = (H2 + 2) * pitch, HW = H * W;
    const int nraw = COUT * H2 * W2;                                 // floats of one image of g (and of y)
    float* gs = (float*)smem4;
    float* graw = gs + 2 * COUT * plane;                              // [2][nraw]
    float* yraw = graw + 2 * nraw;                                    // [2][nraw] when yact
    float* mraw = yraw + (yact ? 2 * nraw : 0);                       // [2][HW*CIN] when x_mask
    uint64_t* full = (uint64_t*)(mraw + (x_mask ? 2 * HW * CIN : 0)); // [2]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, half = lane >> 4, cp = lane & 15;
    const uint32_t tx_bytes = (uint32_t)(nraw * 4 * (yact ? 2 : 1) + (x_mask ? HW * CIN * 4 : 0));
    auto issue = [&](int n, int buf) {                                // one thread
        mbar_arrive_expect_tx(full + buf, tx_bytes);
        bulk_load_1d(graw + buf * nraw, g + (long)n * nraw, nraw * 4, full + buf);
        if (yact) bulk_load_1d(yraw + buf * nraw, yact + (long)n * nraw, nraw * 4, full + buf);
        if (x_mask) bulk_load_1d(mraw + buf * HW * CIN, x_mask + (long)n * HW * CIN, HW * CIN * 4, full + buf);
    };
    if (threadIdx.x == 0) {
        mbar_init(full, 1);
        mbar_init(full + 1, 1);
        mbar_fence_init();
        if ((int)blockIdx.x < N) issue(blockIdx.x, 0);
    }
    // weights as (kx, kx+1) pairs: the inner product runs on packed FFMA2 (two fp32 FMAs per issue slot; scalar FFMA
    // issues every other cycle per scheduler on sm_100, so FFMA2 is what reaches the fp32 peak)
    float2 wr[2][COUT * 8];
#pragma unroll
    for (int q = 0; q < 2; ++q)
#pragma unroll
        for (int t = 0; t < COUT * 4; ++t) {
            const float4 v = __ldg((const float4*)(w + (long)(2 * cp + q) * COUT * 16) + t);      // [ci][co][tap]
            wr[q][2 * t] = make_float2(v.x, v.y);
            wr[q][2 * t + 1] = make_float2(v.z, v.w);
        }
    for (int i = threadIdx.x; i < 2 * COUT * plane; i += CSD_THREADS) gs[i] = 0.f;      // the border stays zero for good
    // copy `half`: column xx of the zero-bordered image (x = xx - 1) stored at xx + 2*half
    const float* gmine = gs + half * COUT * plane + 2 * half;
    int it = 0;
    for (int n = blockIdx.x; n < N; n += gridDim.x, ++it) {
        const int buf = it & 1;
        __syncthreads();                          // image n-1 fully consumed: its buffers (the other set) and gs are free
        if (threadIdx.x == 0 && n + (int)gridDim.x < N) issue(n + gridDim.x, buf ^ 1);
        mbar_wait(full + buf, (it >> 1) & 1);
        const float4* g4 = (const float4*)(graw + buf * nraw);
        const float4* y4 = (const float4*)(yraw + buf * nraw);
        for (int i = threadIdx.x; i < nraw / 4; i += CSD_THREADS) {
            const int e = 4 * i, xx = e % W2, yy = (e / W2) % H2, co = e / (W2 * H2);
            float4 v = g4[i];
            if (yact) {
                const float4 yv = y4[i];
                v.x *= cs_dact(yv.x, act); v.y *= cs_dact(yv.y, act); v.z *= cs_dact(yv.z, act); v.w *= cs_dact(yv.w, act);
            }
            float* r0 = gs + co * plane + (yy + 1) * pitch + xx + 1;
            r0[0] = v.x; r0[1] = v.y; r0[2] = v.z; r0[3] = v.w;
            float* r1 = r0 + COUT * plane + 2;
            r1[0] = v.x; r1[1] = v.y; r1[2] = v.z; r1[3] = v.w;
        }
        __syncthreads();
        const float* mk = mraw + buf * HW * CIN;
        for (int pp = warp; pp < HW / 2; pp += CSD_THREADS / 32) {
            const int p = 2 * pp + half, i = p / W, j = p - i * W;   // W even: the pair shares a row and j's parity == half
            float2 s0 = make_float2(0.f, 0.f), s1 = s0, t0 = s0, t1 = s0;      // two packed chains per channel
            const float* gp = gmine + (2 * i) * pitch + 2 * j;        // row 2i-1+ky -> bordered row 2i+ky; column 2j-1 -> 2j
#pragma unroll
            for (int co = 0; co < COUT; ++co)
#pragma unroll
                for (int ky = 0; ky < 4; ++ky) {
                    const float4 v = *(const float4*)(gp + co * plane + ky * pitch);
                    const int t = co * 8 + ky * 2;
                    const float2 va = make_float2(v.x, v.y), vb = make_float2(v.z, v.w);
                    s0 = __ffma2_rn(va, wr[0][t], s0);     t0 = __ffma2_rn(va, wr[1][t], t0);
                    s1 = __ffma2_rn(vb, wr[0][t + 1], s1); t1 = __ffma2_rn(vb, wr[1][t + 1], t1);
                }
            float2 o = make_float2((s0.x + s0.y) + (s1.x + s1.y), (t0.x + t0.y) + (t1.x + t1.y));
            if (x_mask) {
                const float2 m = *(const float2*)(mk + p * CIN + 2 * cp);
                o.x = m.x > 0.f ? o.x : 0.f;
                o.y = m.y > 0.f ? o.y : 0.f;
            }
            *(float2*)(gx + ((long)n * HW + p) * CIN + 2 * cp) = o;
        }
    }
}

// -------------------------------------------------------------------------------------------------------- wgrad
// gw[ci,co,ky,kx] = sum_{n,i,j} x[n,i,j,ci] * g[n,co,2i-1+ky,2j-1+kx];  gb[co] = sum g.   Cin == 32.
// Persistent, one CTA of 512 threads per SM, inputs of the next image by bulk async copies (as dgrad above).
// warp = one replica of the 32 x (COUT*16) accumulator tile, lane -> (ci group of 4, ky), thread tile 4 ci x COUT x 4 kx
// held as (kx, kx+1) pairs for packed FFMA2; the warps split the pixels.  part[cta][Cin*COUT*16 + COUT]
#define CSW_THREADS 512
__host__ __device__ inline size_t csw_smem_bytes(int H, int W, int Cin, int Cout, bool has_y) {
    const size_t gs = (size_t)Cout * (2 * H + 2) * CSD_PITCH(2 * W) * 4;
    const size_t raw = (size_t)Cout * 4 * H * W * 4;
    size_t tot = gs + 2 * raw * (has_y ? 2 : 1) + 2 * (size_t)H * W * Cin * 4 + 64;
    const size_t red = (size_t)(CSW_THREADS / 32) * Cin * Cout * 16 * 4 + 64;
    return tot > red ? tot : red;
}
template <int COUT, int HT, int WT>
__global__ void __launch_bounds__(CSW_THREADS, 1)
convt_small_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ yact,
                         float* __restrict__ part, int N, int Hrt, int Wrt, int act) {
    constexpr int CIN = 32;
    extern __shared__ float4 smem4[];
    const int H = HT ? HT : Hrt, W = WT ? WT : Wrt;
    const int H2 = 2 * H, W2 = 2 * W, pitch = CSD_PITCH(W2), plane = (H2 + 2) * pitch, HW = H * W;
    const int nraw = COUT * H2 * W2;
    float* gs = (float*)smem4;                                        // [COUT][H2+2][pitch], zero border
    float* graw = gs + COUT * plane;                                  // [2][nraw]
    float* yraw = graw + 2 * nraw;                                    // [2][nraw] when yact
    float* xraw = yraw + (yact ? 2 * nraw : 0);                       // [2][HW*CIN]
    uint64_t* full = (uint64_t*)(xraw + 2 * HW * CIN);                // [2]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cg = lane & 7, ky = lane >> 3;
    const uint32_t tx_bytes = (uint32_t)(nraw * 4 * (yact ? 2 : 1) + HW * CIN * 4);
    auto issue = [&](int n, int buf) {                                // one thread
        mbar_arrive_expect_tx(full + buf, tx_bytes);
        bulk_load_1d(graw + buf * nraw, g + (long)n * nraw, nraw * 4, full + buf);
        if (yact) bulk_load_1d(yraw + buf * nraw, yact + (long)n * nraw, nraw * 4, full + buf);
        bulk_load_1d(xraw + buf * HW * CIN, x + (long)n * HW * CIN, HW * CIN * 4, full + buf);
    };
    if (threadIdx.x == 0) {
        mbar_init(full, 1);
        mbar_init(full + 1, 1);
        mbar_fence_init();
        if ((int)blockIdx.x < N) issue(blockIdx.x, 0);
    }
    float2 acc[4][COUT][2];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < COUT; ++b) acc[a][b][0] = acc[a][b][1] = make_float2(0.f, 0.f);
    float bacc[COUT];
#pragma unroll
    for (int c = 0; c < COUT; ++c) bacc[c] = 0.f;
    for (int i = threadIdx.x; i < COUT * plane; i += CSW_THREADS) gs[i] = 0.f;          // the border stays zero for good
    int it = 0;
    for (int n = blockIdx.x; n < N; n += gridDim.x, ++it) {
        const int buf = it & 1;
        __syncthreads();                          // image n-1 fully consumed: its buffers (the other set) and gs are free
        if (threadIdx.x == 0 && n + (int)gridDim.x < N) issue(n + gridDim.x, buf ^ 1);
        mbar_wait(full + buf, (it >> 1) & 1);
        const float4* g4 = (const float4*)(graw + buf * nraw);
        const float4* y4 = (const float4*)(yraw + buf * nraw);
        for (int i = threadIdx.x; i < nraw / 4; i += CSW_THREADS) {
            const int e = 4 * i, xx = e % W2, yy = (e / W2) % H2, co = e / (W2 * H2);
            float4 v = g4[i];
            if (yact) {
                const float4 yv = y4[i];
                v.x *= cs_dact(yv.x, act); v.y *= cs_dact(yv.y, act); v.z *= cs_dact(yv.z, act); v.w *= cs_dact(yv.w, act);
            }
            float* r0 = gs + co * plane + (yy + 1) * pitch + xx + 1;
            r0[0] = v.x; r0[1] = v.y; r0[2] = v.z; r0[3] = v.w;
            const float sv = (v.x + v.y) + (v.z + v.w);
#pragma unroll
            for (int c = 0; c < COUT; ++c) bacc[c] += (c == co) ? sv : 0.f;
        }
        __syncthreads();
        const float* xs = xraw + buf * HW * CIN;
        const float* gk = gs + ky * pitch;                            // row 2i-1+ky -> bordered row 2i+ky
        for (int p = warp; p < HW; p += CSW_THREADS / 32) {
            const int i = p / W, j = p - i * W;
            const float4 xv = *(const float4*)(xs + p * CIN + 4 * cg);
            const float2 x0 = make_float2(xv.x, xv.x), x1 = make_float2(xv.y, xv.y), x2 = make_float2(xv.z, xv.z),
                         x3 = make_float2(xv.w, xv.w);
#pragma unroll
            for (int co = 0; co < COUT; ++co) {
                const float* r = gk + co * plane + (2 * i) * pitch + 2 * j;
                const float2 a = *(const float2*)r, b = *(const float2*)(r + 2);
                acc[0][co][0] = __ffma2_rn(x0, a, acc[0][co][0]); acc[0][co][1] = __ffma2_rn(x0, b, acc[0][co][1]);
                acc[1][co][0] = __ffma2_rn(x1, a, acc[1][co][0]); acc[1][co][1] = __ffma2_rn(x1, b, acc[1][co][1]);
                acc[2][co][0] = __ffma2_rn(x2, a, acc[2][co][0]); acc[2][co][1] = __ffma2_rn(x2, b, acc[2][co][1]);
                acc[3][co][0] = __ffma2_rn(x3, a, acc[3][co][0]); acc[3][co][1] = __ffma2_rn(x3, b, acc[3][co][1]);
            }
        }
    }
    // cross-warp reduction through shared memory (reuses the staging area), fixed order -> deterministic
    __syncthreads();
    float* red = (float*)smem4;                                      // [warps][CIN*COUT*16]
    constexpr int TOT = CIN * COUT * 16;
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int co = 0; co < COUT; ++co) {
            float* o = red + warp * TOT + ((4 * cg + a) * COUT + co) * 16 + ky * 4;
            o[0] = acc[a][co][0].x; o[1] = acc[a][co][0].y; o[2] = acc[a][co][1].x; o[3] = acc[a][co][1].y;
        }
    __syncthreads();
    float* out = part + (long)blockIdx.x * (TOT + COUT);
    for (int i = threadIdx.x; i < TOT; i += CSW_THREADS) {
        float s = 0.f;
#pragma unroll
        for (int wv = 0; wv < CSW_THREADS / 32; ++wv) s += red[wv * TOT + i];
        out[i] = s;
    }
    __syncthreads();
    float* bred = red;                                               // [COUT][threads]
#pragma unroll
    for (int c = 0; c < COUT; ++c) bred[c * CSW_THREADS + threadIdx.x] = bacc[c];
    __syncthreads();
    if (threadIdx.x < COUT) {
        float s = 0.f;
        for (int t = 0; t < CSW_THREADS; ++t) s += bred[threadIdx.x * CSW_THREADS + t];
        out[TOT + threadIdx.x] = s;
    }
}

__global__ void convt_small_wgrad_reduce(const float* __restrict__ part, float* __restrict__ gw, float* __restrict__ gb,
                                         int P, int tot, int cout) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= tot + cout) return;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;                     // fixed association -> deterministic
    int p = 0;
    for (; p + 4 <= P; p += 4) {
        s0 += part[(long)p * (tot + cout) + i];
        s1 += part[(long)(p + 1) * (tot + cout) + i];
        s2 += part[(long)(p + 2) * (tot + cout) + i];
        s3 += part[(long)(p + 3) * (tot + cout) + i];
    }
    for (; p < P; ++p) s0 += part[(long)p * (tot + cout) + i];
    const float s = (s0 + s1) + (s2 + s3);
    if (i < tot) gw[i] = s;
    else if (gb) gb[i - tot] = s;
}

// ---------------------------------------------------------------------------------------------------- host side
static int cs_wgrad_ctas(int N) { return N < MMVAE_NUM_SMS ? N : MMVAE_NUM_SMS; }

extern "C" int mmvae_convt_small_supported(int H, int W, int Cin, int Cout) {
    if (Cout < 1 || Cout > 4 || Cin != 32 || H < 2 || W < 2 || (W & 1) || (H & 1)) return 0;   // Cin is a template constant (fBase = 32)
    const size_t fwd = csf_smem_bytes(H, W, Cin, Cout);
    const size_t gsz = (((size_t)Cout * (2 * H + 2) * (2 * W + 2) + 3) & ~(size_t)3) * 4;
    const size_t dg = csd_smem_bytes(H, W, Cin, Cout, true, true);
    const size_t wg = csw_smem_bytes(H, W, Cin, Cout, true);
    const size_t lim = 200 * 1024;
    if (fwd > lim || dg > lim) return 0;
    return (Cin == 32 && wg <= lim) ? 2 : 1;       // 2: forward, dgrad and wgrad; 1: forward and dgrad only
}
extern "C" long mmvae_convt_small_wgrad_workspace_bytes(int N, int Cin, int Cout) {
    return (long)cs_wgrad_ctas(N) * ((long)Cin * Cout * 16 + Cout) * (long)sizeof(float);
}

int mmvae_convt_l4_fwd_tc(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int Cin,
                          int Cout, int act, cudaStream_t st);                                   // convt_l4_tc.cu

int mmvae_convt_l4_dgrad_tc(const float* g, const float* y_act, const float* w, const float* x_mask, float* gx, int N,
                            int H, int W, int Cin, int Cout, int act, cudaStream_t st);          // convt_l4_tc.cu

int mmvae_convt_l4_wgrad_ts(const float* x, const float* g, const float* y_act, float* part, int P, int N, int H, int W, int Cin,
                            int Cout, int act, cudaStream_t st);
int mmvae_convt_l4_wgrad_tc(const float* x, const float* g, const float* y_act, float* part, int P, int N, int H, int W,
                            int Cin, int Cout, int act, cudaStream_t st);                        // convt_l4_tc.cu
int mmvae_tc_dbg_bits();                                                                         // gemm_tc.cu
static inline int mmvae_tc_dbg_bits_() { return mmvae_tc_dbg_bits(); }

#define CS_SWITCH(COUTV, FN, ...)                         \
    ((COUTV) == 1 ? FN<1>(__VA_ARGS__) : (COUTV) == 2 ? FN<2>(__VA_ARGS__) : (COUTV) == 3 ? FN<3>(__VA_ARGS__) : FN<4>(__VA_ARGS__))

template <typename K>
static int cs_smem_attr(K kern, size_t smem, const char* name) {
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "%s: %s", name, cudaGetErrorString(e));
    }
    return 0;
}

template <int CO>
static int cs_fwd(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int Cin, int act,
                  size_t smem, cudaStream_t st) {
    const int ctas = N < 2 * MMVAE_NUM_SMS ? N : 2 * MMVAE_NUM_SMS;
    if (H == 16 && W == 16) {
        const int rc = cs_smem_attr(convt_small_fwd_kernel<CO, 32, 16, 16>, smem, "convt_small_fwd");
        if (rc) return rc;
        convt_small_fwd_kernel<CO, 32, 16, 16><<<ctas, CSF_THREADS, smem, st>>>(x, w, bias, y, N, H, W, act);
        return 0;
    }
    const int rc = cs_smem_attr(convt_small_fwd_kernel<CO, 32, 0, 0>, smem, "convt_small_fwd");
    if (rc) return rc;
    convt_small_fwd_kernel<CO, 32, 0, 0><<<ctas, CSF_THREADS, smem, st>>>(x, w, bias, y, N, H, W, act);
    return 0;
}
template <int CO>
static int cs_dgrad(const float* g, const float* y_act, const float* w, const float* x_mask, float* gx, int N, int H,
                    int W, int Cin, int act, size_t smem, cudaStream_t st) {
    const int ctas = N < MMVAE_NUM_SMS ? N : MMVAE_NUM_SMS;
    if (H == 16 && W == 16) {
        const int rc = cs_smem_attr(convt_small_dgrad_kernel<CO, 32, 16, 16>, smem, "convt_small_dgrad");
        if (rc) return rc;
        convt_small_dgrad_kernel<CO, 32, 16, 16><<<ctas, CSD_THREADS, smem, st>>>(g, y_act, w, x_mask, gx, N, H, W, act);
        return 0;
    }
    const int rc = cs_smem_attr(convt_small_dgrad_kernel<CO, 32, 0, 0>, smem, "convt_small_dgrad");
    if (rc) return rc;
    convt_small_dgrad_kernel<CO, 32, 0, 0><<<ctas, CSD_THREADS, smem, st>>>(g, y_act, w, x_mask, gx, N, H, W, act);
    return 0;
}
template <int CO>
static int cs_wgrad(const float* x, const float* g, const float* y_act, float* ws, int P, int N, int H, int W, int act,
                    size_t smem, cudaStream_t st) {
    if (H == 16 && W == 16) {
        const int rc = cs_smem_attr(convt_small_wgrad_kernel<CO, 16, 16>, smem, "convt_small_wgrad");
        if (rc) return rc;
        convt_small_wgrad_kernel<CO, 16, 16><<<P, CSW_THREADS, smem, st>>>(x, g, y_act, ws, N, H, W, act);
        return 0;
    }
    const int rc = cs_smem_attr(convt_small_wgrad_kernel<CO, 0, 0>, smem, "convt_small_wgrad");
    if (rc) return rc;
    convt_small_wgrad_kernel<CO, 0, 0><<<P, CSW_THREADS, smem, st>>>(x, g, y_act, ws, N, H, W, act);
    return 0;
}

extern "C" int mmvae_convt_small_fwd(const float* x, const float* w, const float* bias, float* y, int N, int H, int W,
                                     int Cin, int Cout, int act, void* stream) {
    MMVAE_REQUIRE(x && w && y && N > 0, "convt_small_fwd: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_small_supported(H, W, Cin, Cout), "convt_small_fwd: unsupported shape H=%d W=%d Cin=%d Cout=%d", H, W, Cin, Cout);
    MMVAE_REQUIRE((H & 1) == 0, "convt_small_fwd: H must be even");
    {   // 16x16 inputs (the SVHN decoder's last layer): tcgen05 lowered-GEMM kernel with the col block kept on the SM
        const int rc = mmvae_convt_l4_fwd_tc(x, w, bias, y, N, H, W, Cin, Cout, act, (cudaStream_t)stream);
        if (rc >= 0) return rc;
    }
    const size_t smem = csf_smem_bytes(H, W, Cin, Cout);
    const int rc = CS_SWITCH(Cout, cs_fwd, x, w, bias, y, N, H, W, Cin, act, smem, (cudaStream_t)stream);
    if (rc) return rc;
    MMVAE_LAUNCH_CHECK("convt_small_fwd");
    return 0;
}

extern "C" int mmvae_convt_small_dgrad(const float* g, const float* y_act, const float* w, const float* x_mask,
                                       float* gx, int N, int H, int W, int Cin, int Cout, int act, void* stream) {
    MMVAE_REQUIRE(g && w && gx && N > 0, "convt_small_dgrad: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_small_supported(H, W, Cin, Cout), "convt_small_dgrad: unsupported shape");
    MMVAE_REQUIRE((W & 1) == 0 && ((H * W) & 1) == 0 && ((4L * H * W * Cout) % 4) == 0, "convt_small_dgrad: W must be even");
    {   // 16x16 inputs: tcgen05 kernel with the gathered operand built in shared memory (convt_l4_tc.cu)
        const int rc = mmvae_convt_l4_dgrad_tc(g, y_act, w, x_mask, gx, N, H, W, Cin, Cout, act, (cudaStream_t)stream);
        if (rc >= 0) return rc;
    }
    const size_t smem = csd_smem_bytes(H, W, Cin, Cout, y_act != nullptr, x_mask != nullptr);
    const int rc = CS_SWITCH(Cout, cs_dgrad, g, y_act, w, x_mask, gx, N, H, W, Cin, act, smem, (cudaStream_t)stream);
    if (rc) return rc;
    MMVAE_LAUNCH_CHECK("convt_small_dgrad");
    return 0;
}

extern "C" int mmvae_convt_small_wgrad(const float* x, const float* g, const float* y_act, float* gw, float* gb,
                                       float* workspace, int N, int H, int W, int Cin, int Cout, int act, void* stream) {
    MMVAE_REQUIRE(x && g && gw && workspace && N > 0, "convt_small_wgrad: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_small_supported(H, W, Cin, Cout) == 2, "convt_small_wgrad: unsupported shape (needs Cin == 32)");
    const size_t smem = csw_smem_bytes(H, W, Cin, Cout, y_act != nullptr);
    const int P = cs_wgrad_ctas(N);
    // the tensor-memory kernel of convt_ts.cu (bf16x3) unless debug bit 32 asks for the fp32-FMA kernel; bit 16384: the
    // first-generation shared-memory-operand tcgen05 kernel (slower than either, kept for comparison)
    int rc = (mmvae_tc_dbg_bits_() & 16384) ? mmvae_convt_l4_wgrad_tc(x, g, y_act, workspace, P, N, H, W, Cin, Cout, act, (cudaStream_t)stream)
             : ((mmvae_tc_dbg_bits_() & 32) ? -1 : mmvae_convt_l4_wgrad_ts(x, g, y_act, workspace, P, N, H, W, Cin, Cout, act, (cudaStream_t)stream));
    if (rc > 0) return rc;
    if (rc < 0) {
        rc = CS_SWITCH(Cout, cs_wgrad, x, g, y_act, workspace, P, N, H, W, act, smem, (cudaStream_t)stream);
        if (rc) return rc;
        MMVAE_LAUNCH_CHECK("convt_small_wgrad");
    }
    const int tot = Cin * Cout * 16;
    convt_small_wgrad_reduce<<<ceil_div_i(tot + Cout, 128), 128, 0, (cudaStream_t)stream>>>(workspace, gw, gb, P, tot, Cout);
    MMVAE_LAUNCH_CHECK("convt_small_wgrad_reduce");
    return 0;
}
