// Fused MADE for training (the parallel direction of a flow block: MAF.forward maf_model.py:62-74, IAF.inverse
// iaf_model.py:94-100; network made_model.py:21-111, masked layers layers.py:14-25).  A MADE on B <= 1024 rows is four masked
// Linear layers of 10^6 MACs each: as separate launches (mask, GEMM, bias gradient, activation backward, ...) one block costs 27
// latency-bound kernels per step.  Three kernels replace them:
//   made_fwd_kernel          8 rows per CTA through every layer; the masked weight (W * mask) of a layer is formed while it is
//                            staged in shared memory; hidden activations are saved for the backward pass
//   made_bwd_data_kernel     8 rows per CTA back through every layer: the gradient w.r.t. the input and the pre-activation
//                            gradients of every layer
//   made_bwd_weights_kernel  one CTA per (layer, 16 output units): masked weight gradient and bias gradient over all rows, in a
//                            fixed order (deterministic); it feeds only the optimizer, so the host side issues it on the
//                            weight-gradient stream
// Exact fp32 FMA arithmetic.  Every layer width must be a multiple of 4 and at most 128.
#include "common.cuh"

#define MD_ROWS 8
#define MD_RPT (MD_ROWS / 2)      // rows per thread (two row groups of 128 threads)
#define MD_THREADS 256
#define MD_MAXW 128
#define MD_MAXL 5
#define MD_ACT_LD (MD_MAXW + 4)

struct MadeArgs {
    const float* W[MD_MAXL];
    const float* M[MD_MAXL];
    const float* b[MD_MAXL];
    float* gW[MD_MAXL];
    float* gb[MD_MAXL];
    int in[MD_MAXL], out[MD_MAXL];
    int n_layers, R;
    const float* x;          // [R, in[0]]
    float* h;                // [n_layers-1][R][H] post-ReLU hidden activations (H = out[0])
    float* y;                // [R, out[last]]
    const float* gy;         // [R, out[last]]
    float* gpre;             // [n_layers-1][R][H] gradients w.r.t. the hidden pre-activations
    float* gx;               // [R, in[0]] or NULL
};

// leading dimension of a staged weight [out][in]: a multiple of 4 that is 4 mod 32, so that 32 consecutive rows read as float4
// cover every bank exactly four times (the minimum for 512 bytes)
__host__ __device__ inline int md_ld(int in) { return ((in + 27) / 32) * 32 + 4; }

__device__ __forceinline__ float4 md_ld4(const float* __restrict__ p, bool vec) {
    if (vec) return __ldg(reinterpret_cast<const float4*>(p));
    return make_float4(__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3));          // a view into a flat parameter buffer
}

// Ws[o][i] = W[o][i] * mask[o][i].  Eight 16-byte loads per thread are in flight before the first store: the staging of a layer is
// one L2 round trip, not one per element.
__device__ __forceinline__ void md_stage_weight(float* Ws, const float* __restrict__ W, const float* __restrict__ M, int out, int in) {
    const int ld = md_ld(in), q = in / 4, total = out * q;
    const bool wv = (((uintptr_t)W) & 15) == 0, mv = (((uintptr_t)M) & 15) == 0;
    constexpr int U = 4;
    for (int e0 = threadIdx.x; e0 < total; e0 += MD_THREADS * U) {
        float4 w[U], m[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int e = e0 + u * MD_THREADS;
            if (e < total) {
                w[u] = md_ld4(W + (long)e * 4, wv);
                m[u] = md_ld4(M + (long)e * 4, mv);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int e = e0 + u * MD_THREADS;
            if (e < total) {
                const int o = e / q, i4 = e - o * q;
                *reinterpret_cast<float4*>(Ws + o * ld + i4 * 4) = make_float4(w[u].x * m[u].x, w[u].y * m[u].y, w[u].z * m[u].z, w[u].w * m[u].w);
            }
        }
    }
}

__global__ void __launch_bounds__(MD_THREADS)
made_fwd_kernel(const MadeArgs a) {
    extern __shared__ __align__(16) float md_smem[];
    float* Ws = md_smem;                                   // [128][132]
    float* act = md_smem + MD_MAXW * MD_ACT_LD;            // [2][MD_ROWS][132]
    const int r0 = blockIdx.x * MD_ROWS, c = threadIdx.x % MD_MAXW, rg = threadIdx.x / MD_MAXW;
    const int H = a.out[0];
    for (int e = threadIdx.x; e < MD_ROWS * a.in[0]; e += MD_THREADS) {
        const int r = e / a.in[0], i = e - r * a.in[0];
        act[r * MD_ACT_LD + i] = (r0 + r < a.R) ? a.x[(long)(r0 + r) * a.in[0] + i] : 0.f;
    }
    int cur = 0;
    for (int l = 0; l < a.n_layers; ++l) {
        const int K = a.in[l], N = a.out[l], ld = md_ld(K);
        md_stage_weight(Ws, a.W[l], a.M[l], N, K);
        __syncthreads();
        const float* ain = act + cur * MD_ROWS * MD_ACT_LD + rg * MD_RPT * MD_ACT_LD;
        float* aout = act + (cur ^ 1) * MD_ROWS * MD_ACT_LD + rg * MD_RPT * MD_ACT_LD;
        if (c < N) {
            const float bias = a.b[l] ? a.b[l][c] : 0.f;
            float acc[MD_RPT];
#pragma unroll
            for (int r = 0; r < MD_RPT; ++r) acc[r] = bias;
            const float* wrow = Ws + c * ld;
            for (int k = 0; k < K; k += 4) {
                const float4 w = *reinterpret_cast<const float4*>(wrow + k);
#pragma unroll
                for (int r = 0; r < MD_RPT; ++r) {
                    const float4 v = *reinterpret_cast<const float4*>(ain + r * MD_ACT_LD + k);     // same address in every lane
                    acc[r] = fmaf(v.x, w.x, acc[r]); acc[r] = fmaf(v.y, w.y, acc[r]);
                    acc[r] = fmaf(v.z, w.z, acc[r]); acc[r] = fmaf(v.w, w.w, acc[r]);
                }
            }
            const bool last = (l == a.n_layers - 1);
#pragma unroll
            for (int r = 0; r < MD_RPT; ++r) {
                const float v = last ? acc[r] : fmaxf(acc[r], 0.f);
                aout[r * MD_ACT_LD + c] = v;
                const int gr = r0 + rg * MD_RPT + r;
                if (gr < a.R) {
                    if (last) a.y[(long)gr * N + c] = v;
                    else a.h[((long)l * a.R + gr) * H + c] = v;
                }
            }
        }
        __syncthreads();
        cur ^= 1;
    }
}

__global__ void __launch_bounds__(MD_THREADS)
made_bwd_data_kernel(const MadeArgs a) {
    extern __shared__ __align__(16) float md_smem[];
    float* Ws = md_smem;
    float* gbuf = md_smem + MD_MAXW * MD_ACT_LD;           // [2][MD_ROWS][132]
    const int r0 = blockIdx.x * MD_ROWS, c = threadIdx.x % MD_MAXW, rg = threadIdx.x / MD_MAXW;
    const int H = a.out[0], last = a.n_layers - 1;
    {
        const int N = a.out[last];
        for (int e = threadIdx.x; e < MD_ROWS * N; e += MD_THREADS) {
            const int r = e / N, o = e - r * N;
            gbuf[r * MD_ACT_LD + o] = (r0 + r < a.R) ? a.gy[(long)(r0 + r) * N + o] : 0.f;
        }
    }
    int cur = 0;
    for (int l = last; l >= 0; --l) {
        if (l == 0 && !a.gx) break;
        const int K = a.in[l], N = a.out[l], ld = md_ld(K);
        md_stage_weight(Ws, a.W[l], a.M[l], N, K);
        __syncthreads();
        const float* gin = gbuf + cur * MD_ROWS * MD_ACT_LD + rg * MD_RPT * MD_ACT_LD;
        float* gout = gbuf + (cur ^ 1) * MD_ROWS * MD_ACT_LD + rg * MD_RPT * MD_ACT_LD;
        if (c < K) {
            float acc[MD_RPT];
#pragma unroll
            for (int r = 0; r < MD_RPT; ++r) acc[r] = 0.f;
            for (int o = 0; o < N; o += 4) {
                const float w0 = Ws[(o + 0) * ld + c], w1 = Ws[(o + 1) * ld + c], w2 = Ws[(o + 2) * ld + c], w3 = Ws[(o + 3) * ld + c];
#pragma unroll
                for (int r = 0; r < MD_RPT; ++r) {
                    const float4 g = *reinterpret_cast<const float4*>(gin + r * MD_ACT_LD + o);       // broadcast
                    acc[r] = fmaf(g.x, w0, acc[r]); acc[r] = fmaf(g.y, w1, acc[r]);
                    acc[r] = fmaf(g.z, w2, acc[r]); acc[r] = fmaf(g.w, w3, acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < MD_RPT; ++r) {
                const int gr = r0 + rg * MD_RPT + r;
                float v = acc[r];
                if (l > 0) {                                                    // through the ReLU that produced h_{l-1}
                    const float hv = (gr < a.R) ? a.h[((long)(l - 1) * a.R + gr) * H + c] : 0.f;
                    v = hv > 0.f ? v : 0.f;
                    if (gr < a.R) a.gpre[((long)(l - 1) * a.R + gr) * H + c] = v;
                } else if (gr < a.R) {
                    a.gx[(long)gr * K + c] = v;
                }
                gout[r * MD_ACT_LD + c] = v;
            }
        }
        __syncthreads();
        cur ^= 1;
    }
}

// grid: sum over layers of ceil(out/16) CTAs; blockIdx.x -> (layer, output block) through the prefix table in `first`
struct MadeWgradMap { int first[MD_MAXL + 1]; };

__global__ void __launch_bounds__(MD_THREADS)
made_bwd_weights_kernel(const MadeArgs a, const MadeWgradMap map) {
    extern __shared__ __align__(16) float md_smem[];
    float* hs = md_smem;                                   // [128 rows][K]   inputs of the layer
    float* gs = md_smem + MD_MAXW * MD_MAXW;               // [128 rows][16]  pre-activation gradients of the CTA's 16 outputs
    int l = 0;
    while (l + 1 < a.n_layers && (int)blockIdx.x >= map.first[l + 1]) ++l;
    const int ob = ((int)blockIdx.x - map.first[l]) * 16;
    const int K = a.in[l], N = a.out[l], H = a.out[0], last = a.n_layers - 1;
    const float* in = (l == 0) ? a.x : a.h + (long)(l - 1) * a.R * H;       // [R][K]
    const float* g = (l == last) ? a.gy : a.gpre + (long)l * a.R * H;       // [R][N]
    const int c = threadIdx.x % MD_MAXW, og = threadIdx.x / MD_MAXW;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    float bsum = 0.f;
    const bool in_vec = (((uintptr_t)in) & 15) == 0;
    for (int rc = 0; rc < a.R; rc += 128) {
        const int rows = min(128, a.R - rc);
        {                                                   // rows of the layer input are contiguous: [rows*K] floats, 4 loads in flight
            const int total = rows * (K / 4);
            const float* src = in + (long)rc * K;
            for (int e0 = threadIdx.x; e0 < total; e0 += MD_THREADS * 4) {
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (e0 + u * MD_THREADS < total) v[u] = md_ld4(src + (long)(e0 + u * MD_THREADS) * 4, in_vec);
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (e0 + u * MD_THREADS < total) *reinterpret_cast<float4*>(hs + (e0 + u * MD_THREADS) * 4) = v[u];
            }
        }
        for (int e = threadIdx.x; e < rows * 16; e += MD_THREADS) {
            const int r = e / 16, o = e % 16;
            gs[r * 16 + o] = (ob + o < N) ? g[(long)(rc + r) * N + ob + o] : 0.f;
        }
        __syncthreads();
        if (c < K) {
            for (int r = 0; r < rows; ++r) {
                const float hv = hs[r * K + c];
                const float4 g0 = *reinterpret_cast<const float4*>(gs + r * 16 + og * 8);           // broadcast
                const float4 g1 = *reinterpret_cast<const float4*>(gs + r * 16 + og * 8 + 4);
                acc[0] = fmaf(g0.x, hv, acc[0]); acc[1] = fmaf(g0.y, hv, acc[1]); acc[2] = fmaf(g0.z, hv, acc[2]); acc[3] = fmaf(g0.w, hv, acc[3]);
                acc[4] = fmaf(g1.x, hv, acc[4]); acc[5] = fmaf(g1.y, hv, acc[5]); acc[6] = fmaf(g1.z, hv, acc[6]); acc[7] = fmaf(g1.w, hv, acc[7]);
            }
        }
        if (threadIdx.x < 16)
            for (int r = 0; r < rows; ++r) bsum += gs[r * 16 + threadIdx.x];
        __syncthreads();
    }
    if (c < K && a.gW[l]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int o = ob + og * 8 + j;
            if (o < N) a.gW[l][(long)o * K + c] = acc[j] * a.M[l][(long)o * K + c];
        }
    }
    if (threadIdx.x < 16 && ob + threadIdx.x < N && a.gb[l]) a.gb[l][ob + threadIdx.x] = bsum;
}

#define MD_FWD_SMEM ((MD_MAXW * MD_ACT_LD + 2 * MD_ROWS * MD_ACT_LD) * (int)sizeof(float))
#define MD_WG_SMEM ((MD_MAXW * MD_MAXW + 128 * 16) * (int)sizeof(float))

static int md_fill(MadeArgs* a, const void* const* ptrs, int n_layers, const int* dims, int R) {
    if (n_layers < 2 || n_layers > MD_MAXL) return mmvae_set_error(1, "made: %d layers (2..%d supported)", n_layers, MD_MAXL);
    for (int l = 0; l < n_layers; ++l) {
        a->in[l] = dims[l]; a->out[l] = dims[l + 1];
        if (dims[l] % 4 || dims[l + 1] % 4 || dims[l] > MD_MAXW || dims[l + 1] > MD_MAXW || dims[l] < 4)
            return mmvae_set_error(1, "made: layer %d is %d -> %d (widths must be multiples of 4, at most %d)", l, dims[l], dims[l + 1], MD_MAXW);
        if (l > 0 && l < n_layers - 1 && dims[l + 1] != dims[1]) return mmvae_set_error(1, "made: hidden widths differ");
        a->W[l] = (const float*)ptrs[3 * l]; a->M[l] = (const float*)ptrs[3 * l + 1]; a->b[l] = (const float*)ptrs[3 * l + 2];
        if (!a->W[l] || !a->M[l]) return mmvae_set_error(1, "made: null weight / mask of layer %d", l);
    }
    a->n_layers = n_layers; a->R = R;
    return 0;
}

static int md_attr(const void* kern, int bytes, bool* done) {
    if (*done) return 0;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "made: smem attribute: %s", cudaGetErrorString(e));
    *done = true;
    return 0;
}

extern "C" int mmvae_made_supported(int n_layers, const int* dims) {
    if (n_layers < 2 || n_layers > MD_MAXL) return 0;
    for (int l = 0; l <= n_layers; ++l)
        if (dims[l] % 4 || dims[l] > MD_MAXW || dims[l] < 4) return 0;
    for (int l = 2; l < n_layers; ++l)
        if (dims[l] != dims[1]) return 0;
    return 1;
}

// ptrs: host array of 3*n_layers device pointers {W_l [out,in], mask_l [out,in], bias_l [out] or NULL}; dims: host array of
// n_layers+1 widths.  x [R,dims[0]] -> y [R,dims[n_layers]]; h: [n_layers-1][R][dims[1]] (saved for the backward pass).
extern "C" int mmvae_made_fwd(const float* x, const void* ptrs, const int* dims, int n_layers, float* h, float* y, int R, void* stream) {
    MMVAE_REQUIRE(x && ptrs && dims && h && y && R > 0, "made_fwd: bad arguments");
    MadeArgs a{};
    if (int rc = md_fill(&a, (const void* const*)ptrs, n_layers, dims, R)) return rc;
    a.x = x; a.h = h; a.y = y;
    static bool done = false;
    if (int rc = md_attr((const void*)made_fwd_kernel, MD_FWD_SMEM, &done)) return rc;
    made_fwd_kernel<<<(R + MD_ROWS - 1) / MD_ROWS, MD_THREADS, MD_FWD_SMEM, (cudaStream_t)stream>>>(a);
    MMVAE_LAUNCH_CHECK("made_fwd");
    return 0;
}

// gy [R,out] -> gx [R,in] (or NULL) and gpre [n_layers-1][R][H], the gradients w.r.t. the hidden pre-activations
extern "C" int mmvae_made_bwd_data(const float* gy, const void* ptrs, const int* dims, int n_layers, const float* h, float* gpre,
                                   float* gx, int R, void* stream) {
    MMVAE_REQUIRE(gy && ptrs && dims && h && gpre && R > 0, "made_bwd_data: bad arguments");
    MadeArgs a{};
    if (int rc = md_fill(&a, (const void* const*)ptrs, n_layers, dims, R)) return rc;
    a.gy = gy; a.h = const_cast<float*>(h); a.gpre = gpre; a.gx = gx;
    static bool done = false;
    if (int rc = md_attr((const void*)made_bwd_data_kernel, MD_FWD_SMEM, &done)) return rc;
    made_bwd_data_kernel<<<(R + MD_ROWS - 1) / MD_ROWS, MD_THREADS, MD_FWD_SMEM, (cudaStream_t)stream>>>(a);
    MMVAE_LAUNCH_CHECK("made_bwd_data");
    return 0;
}

// grads: host array of 2*n_layers device pointers {gW_l [out,in], gb_l [out]} (NULL = not wanted); both are OVERWRITTEN with
// mask * (gpre_l^T input_l) and the column sums of gpre_l over the R rows
extern "C" int mmvae_made_bwd_weights(const float* x, const float* gy, const void* ptrs, const void* grads, const int* dims,
                                      int n_layers, const float* h, const float* gpre, int R, void* stream) {
    MMVAE_REQUIRE(x && gy && ptrs && grads && dims && h && gpre && R > 0, "made_bwd_weights: bad arguments");
    MadeArgs a{};
    if (int rc = md_fill(&a, (const void* const*)ptrs, n_layers, dims, R)) return rc;
    a.x = x; a.gy = gy; a.h = const_cast<float*>(h); a.gpre = const_cast<float*>(gpre);
    MadeWgradMap map{};
    int total = 0;
    for (int l = 0; l < n_layers; ++l) {
        a.gW[l] = (float*)((void* const*)grads)[2 * l]; a.gb[l] = (float*)((void* const*)grads)[2 * l + 1];
        map.first[l] = total;
        total += (a.out[l] + 15) / 16;
    }
    map.first[n_layers] = total;
    static bool done = false;
    if (int rc = md_attr((const void*)made_bwd_weights_kernel, MD_WG_SMEM, &done)) return rc;
    made_bwd_weights_kernel<<<total, MD_THREADS, MD_WG_SMEM, (cudaStream_t)stream>>>(a, map);
    MMVAE_LAUNCH_CHECK("made_bwd_weights");
    return 0;
}
