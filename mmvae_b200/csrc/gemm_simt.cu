// Kernel (4), exact-fp32 path: tiled SIMT GEMM with fused bias + activation epilogue and deterministic split-K.
// C[M,N] = act(A*B + bias).  Used for every shape the tcgen05 path does not take (skinny K = latent dim, tiny M,
// unaligned sizes) and as the bit-honest fp32 cross-check of the tensor-core path.
// Reference ops replaced: torch.nn.Linear / F.linear (pythae Encoder_VAE_MLP / Decoder_AE_MLP; joint_encoders.py:61-82;
// encoders.py:181-210; layers.py:24-25) and, through conv_lower.cu, Conv2d / ConvTranspose2d (encoders.py:80-136),
// plus their autograd backward (dgrad: a_kmajor=1,b_kmajor=0; wgrad: a_kmajor=0,b_kmajor=0).
#include "common.cuh"

#define G_BM 128
#define G_BK 16
#define G_PAD 4
#define G_THREADS 256

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == MMVAE_ACT_RELU) return fmaxf(v, 0.f);
    if (act == MMVAE_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
    return v;
}

// Load one (ROWS x G_BK) operand tile into registers.  KMAJOR: src[row*ld + k] ; else src[k*ld + row].
// Each thread owns NV float4 (vector path) — zero-filled outside [0,rows_total) x [k_begin,k_end).
template <int ROWS, bool KMAJOR>
struct TileLoader {
    static constexpr int NV = ROWS * G_BK / 4 / G_THREADS;      // float4 per thread
    float4 v[NV];
    __device__ __forceinline__ void load(const float* __restrict__ src, int ld, int row0, int rows_total, int k0,
                                         int k_end, bool vec_ok) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int f = threadIdx.x + i * G_THREADS;
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (KMAJOR) {
                const int row = f / (G_BK / 4), kq = (f % (G_BK / 4)) * 4;
                const int gr = row0 + row, gk = k0 + kq;
                if (gr < rows_total) {
                    const float* p = src + (long)gr * ld + gk;
                    if (vec_ok && gk + 3 < k_end) t = *reinterpret_cast<const float4*>(p);
                    else {
                        if (gk < k_end) t.x = p[0];
                        if (gk + 1 < k_end) t.y = p[1];
                        if (gk + 2 < k_end) t.z = p[2];
                        if (gk + 3 < k_end) t.w = p[3];
                    }
                }
            } else {
                const int k = f / (ROWS / 4), rq = (f % (ROWS / 4)) * 4;
                const int gk = k0 + k, gr = row0 + rq;
                if (gk < k_end) {
                    const float* p = src + (long)gk * ld + gr;
                    if (vec_ok && gr + 3 < rows_total) t = *reinterpret_cast<const float4*>(p);
                    else {
                        if (gr < rows_total) t.x = p[0];
                        if (gr + 1 < rows_total) t.y = p[1];
                        if (gr + 2 < rows_total) t.z = p[2];
                        if (gr + 3 < rows_total) t.w = p[3];
                    }
                }
            }
            v[i] = t;
        }
    }
    // smem tile layout: s[k][row], row stride ROWS + G_PAD
    __device__ __forceinline__ void store(float* __restrict__ s) const {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int f = threadIdx.x + i * G_THREADS;
            if (KMAJOR) {
                const int row = f / (G_BK / 4), kq = (f % (G_BK / 4)) * 4;
                s[(kq + 0) * (ROWS + G_PAD) + row] = v[i].x;
                s[(kq + 1) * (ROWS + G_PAD) + row] = v[i].y;
                s[(kq + 2) * (ROWS + G_PAD) + row] = v[i].z;
                s[(kq + 3) * (ROWS + G_PAD) + row] = v[i].w;
            } else {
                const int k = f / (ROWS / 4), rq = (f % (ROWS / 4)) * 4;
                *reinterpret_cast<float4*>(&s[k * (ROWS + G_PAD) + rq]) = v[i];
            }
        }
    }
};

// grid: (ceil(N/BN), ceil(M/BM), S).  S == 1: writes act(acc+bias) to C; S > 1: raw partials to ws[z][M][N].
template <int BN, bool A_KMAJOR, bool B_KMAJOR>
__global__ void __launch_bounds__(G_THREADS, 2)
gemm_simt_kernel(const float* __restrict__ A, const float* __restrict__ Bm, const float* __restrict__ bias,
                 float* __restrict__ C, float* __restrict__ ws, int M, int N, int K, int act, int tiles_per_split) {
    constexpr int TN = BN / 16;                                   // columns per thread (8 or 4)
    __shared__ __align__(16) float As[2][G_BK][G_BM + G_PAD];
    __shared__ __align__(16) float Bs[2][G_BK][BN + G_PAD];
    const int m0 = blockIdx.y * G_BM, n0 = blockIdx.x * BN;
    const int kt_begin = blockIdx.z * tiles_per_split;
    const int nk_total = (K + G_BK - 1) / G_BK;
    const int kt_end = min(nk_total, kt_begin + tiles_per_split);
    const int lda = A_KMAJOR ? K : M, ldb = B_KMAJOR ? K : N;
    const bool a_vec = (((uintptr_t)A & 15) == 0) && (lda % 4 == 0);
    const bool b_vec = (((uintptr_t)Bm & 15) == 0) && (ldb % 4 == 0);
    const int ty = threadIdx.x / 16, tx = threadIdx.x % 16;

    float acc[8][TN];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    TileLoader<G_BM, A_KMAJOR> la;
    TileLoader<BN, B_KMAJOR> lb;
    if (kt_begin < kt_end) {
        la.load(A, lda, m0, M, kt_begin * G_BK, K, a_vec);
        lb.load(Bm, ldb, n0, N, kt_begin * G_BK, K, b_vec);
        la.store(&As[0][0][0]);
        lb.store(&Bs[0][0][0]);
    }
    __syncthreads();
    int cur = 0;
    for (int kt = kt_begin; kt < kt_end; ++kt) {
        const bool more = (kt + 1 < kt_end);
        if (more) {
            la.load(A, lda, m0, M, (kt + 1) * G_BK, K, a_vec);
            lb.load(Bm, ldb, n0, N, (kt + 1) * G_BK, K, b_vec);
        }
#pragma unroll
        for (int kk = 0; kk < G_BK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[cur][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[cur][kk][64 + ty * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float bv[TN];
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[cur][kk][tx * 4]);
            bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
            if (TN == 8) {
                const float4 b1 = *reinterpret_cast<const float4*>(&Bs[cur][kk][BN / 2 + tx * 4]);
                bv[TN - 4] = b1.x; bv[TN - 3] = b1.y; bv[TN - 2] = b1.z; bv[TN - 1] = b1.w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (more) {
            la.store(&As[cur ^ 1][0][0]);
            lb.store(&Bs[cur ^ 1][0][0]);
        }
        __syncthreads();
        cur ^= 1;
    }

    const bool split = (gridDim.z > 1);
    float* out = split ? ws + (long)blockIdx.z * M * N : C;
    const bool c_vec = (((uintptr_t)out & 15) == 0) && (N % 4 == 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (gm >= M) continue;
#pragma unroll
        for (int h = 0; h < TN / 4; ++h) {
            const int gn = n0 + (h == 0 ? tx * 4 : BN / 2 + tx * 4);
            float r[4] = {acc[i][h * 4 + 0], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]};
            if (!split) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (gn + j < N) r[j] = apply_act(r[j] + (bias ? bias[gn + j] : 0.f), act);
            }
            float* p = out + (long)gm * N + gn;
            if (c_vec && gn + 3 < N) *reinterpret_cast<float4*>(p) = make_float4(r[0], r[1], r[2], r[3]);
            else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (gn + j < N) p[j] = r[j];
            }
        }
    }
}

// C = act(sum_z ws[z] + bias), partials added in the order z = 0..S-1 (deterministic).  Eight loads are in flight before the
// first add: the S loads of an element are independent, only the adds are ordered (the plain loop paid one L2 round trip per
// partial: 9 us per launch, 11 launches per C2 step).  VEC: 4 consecutive elements per thread (MN, N multiples of 4).
template <bool VEC>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ ws, const float* __restrict__ bias, float* __restrict__ C, long MN,
                     int N, int S, int act) {
    if (VEC) {
        const long MN4 = MN >> 2;
        const float4* w4 = reinterpret_cast<const float4*>(ws);
        for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < MN4; i += (long)gridDim.x * blockDim.x) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            int z = 0;
            for (; z + 8 <= S; z += 8) {
                float4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __ldcs(w4 + (long)(z + u) * MN4 + i);
#pragma unroll
                for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
            }
            for (; z < S; ++z) {
                const float4 v = __ldcs(w4 + (long)z * MN4 + i);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            if (bias) {
                const int n = (int)((i * 4) % N);
                s.x += bias[n]; s.y += bias[n + 1]; s.z += bias[n + 2]; s.w += bias[n + 3];
            }
            reinterpret_cast<float4*>(C)[i] = make_float4(apply_act(s.x, act), apply_act(s.y, act), apply_act(s.z, act), apply_act(s.w, act));
        }
    } else {
        for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < MN; i += (long)gridDim.x * blockDim.x) {
            float s = 0.f;
            int z = 0;
            for (; z + 8 <= S; z += 8) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __ldcs(ws + (long)(z + u) * MN + i);
#pragma unroll
                for (int u = 0; u < 8; ++u) s += v[u];
            }
            for (; z < S; ++z) s += __ldcs(ws + (long)z * MN + i);
            if (bias) s += bias[i % N];
            C[i] = apply_act(s, act);
        }
    }
}

// Small outputs with many partials (the weight gradient of the encoder's first convolution: 32 x 48 outputs, K = 32768): the
// kernel above would put all S partials of an element on one thread of a handful of blocks (44 us measured).  Here a block
// owns 32 float4 of the output and its 8 warps take the partials z = w, w + 8, ... (coalesced 512-byte rows); the eight sums are
// added in warp order through shared memory -- a fixed association, so still deterministic.
__global__ void __launch_bounds__(256)
splitk_reduce_small_kernel(const float* __restrict__ ws, const float* __restrict__ bias, float* __restrict__ C, long MN, int N,
                           int S, int act) {
    __shared__ float4 red[8][32];
    const long MN4 = MN >> 2;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const long i = (long)blockIdx.x * 32 + lane;
    const float4* w4 = reinterpret_cast<const float4*>(ws);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < MN4) {
        int z = w;
        for (; z + 24 < S; z += 32) {
            const float4 v0 = __ldcs(w4 + (long)z * MN4 + i), v1 = __ldcs(w4 + (long)(z + 8) * MN4 + i);
            const float4 v2 = __ldcs(w4 + (long)(z + 16) * MN4 + i), v3 = __ldcs(w4 + (long)(z + 24) * MN4 + i);
            s.x += v0.x; s.y += v0.y; s.z += v0.z; s.w += v0.w;
            s.x += v1.x; s.y += v1.y; s.z += v1.z; s.w += v1.w;
            s.x += v2.x; s.y += v2.y; s.z += v2.z; s.w += v2.w;
            s.x += v3.x; s.y += v3.y; s.z += v3.z; s.w += v3.w;
        }
        for (; z < S; z += 8) {
            const float4 v = __ldcs(w4 + (long)z * MN4 + i);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    }
    red[w][lane] = s;
    __syncthreads();
    if (w == 0 && i < MN4) {
        float4 t = red[0][lane];
#pragma unroll
        for (int j = 1; j < 8; ++j) { const float4 v = red[j][lane]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
        if (bias) {
            const int n = (int)((i * 4) % N);
            t.x += bias[n]; t.y += bias[n + 1]; t.z += bias[n + 2]; t.w += bias[n + 3];
        }
        reinterpret_cast<float4*>(C)[i] = make_float4(apply_act(t.x, act), apply_act(t.y, act), apply_act(t.z, act), apply_act(t.w, act));
    }
}

static void splitk_reduce_launch(const float* ws, const float* bias, float* C, long MN, int N, int S, int act, cudaStream_t st) {
    const bool vec = (N % 4 == 0) && ((((uintptr_t)ws | (uintptr_t)C) & 15) == 0);
    const long work = vec ? MN / 4 : MN;
    if (vec && S >= 16 && work <= 32L * MMVAE_NUM_SMS * 2) {
        splitk_reduce_small_kernel<<<(int)((work + 31) / 32), 256, 0, st>>>(ws, bias, C, MN, N, S, act);
        return;
    }
    const int grid = (int)min((long)MMVAE_NUM_SMS * 8, (work + 255) / 256);
    if (vec) splitk_reduce_kernel<true><<<grid, 256, 0, st>>>(ws, bias, C, MN, N, S, act);
    else splitk_reduce_kernel<false><<<grid, 256, 0, st>>>(ws, bias, C, MN, N, S, act);
}

static int pick_splits(int M, int N, int K, int BN) {
    const long tiles = (long)((M + G_BM - 1) / G_BM) * ((N + BN - 1) / BN);
    const int nk = (K + G_BK - 1) / G_BK;
    if (tiles >= MMVAE_NUM_SMS || nk < 8) return 1;
    long s = (2L * MMVAE_NUM_SMS + tiles - 1) / tiles;
    if (s > nk / 4) s = nk / 4;
    if (s > 64) s = 64;
    return s < 1 ? 1 : (int)s;
}

long mmvae_gemm_simt_workspace_bytes(int M, int N, int K) {
    const int BN = (N <= 64) ? 64 : 128;
    const int S = pick_splits(M, N, K, BN);
    return S > 1 ? (long)S * M * N * (long)sizeof(float) : 0;
}

void mmvae_splitk_reduce(const float* ws, const float* bias, float* C, long MN, int N, int S, int act, cudaStream_t st) {
    splitk_reduce_launch(ws, bias, C, MN, N, S, act, st);
}

template <int BN>
static void launch_simt(const float* A, const float* B, const float* bias, float* C, float* ws, int M, int N, int K,
                        int a_k, int b_k, int act, int S, cudaStream_t st) {
    const int nk = (K + G_BK - 1) / G_BK;
    const int tps = (nk + S - 1) / S;
    dim3 grid((N + BN - 1) / BN, (M + G_BM - 1) / G_BM, S);
    if (a_k && b_k) gemm_simt_kernel<BN, true, true><<<grid, G_THREADS, 0, st>>>(A, B, bias, C, ws, M, N, K, act, tps);
    else if (a_k && !b_k) gemm_simt_kernel<BN, true, false><<<grid, G_THREADS, 0, st>>>(A, B, bias, C, ws, M, N, K, act, tps);
    else if (!a_k && b_k) gemm_simt_kernel<BN, false, true><<<grid, G_THREADS, 0, st>>>(A, B, bias, C, ws, M, N, K, act, tps);
    else gemm_simt_kernel<BN, false, false><<<grid, G_THREADS, 0, st>>>(A, B, bias, C, ws, M, N, K, act, tps);
}

int mmvae_gemm_simt(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor,
                    int b_kmajor, int act, float* workspace, long workspace_bytes, cudaStream_t st) {
    const int BN = (N <= 64) ? 64 : 128;
    int S = pick_splits(M, N, K, BN);
    if (S > 1 && (!workspace || workspace_bytes < (long)S * M * N * (long)sizeof(float))) S = 1;
    if (BN == 64) launch_simt<64>(A, B, bias, C, workspace, M, N, K, a_kmajor, b_kmajor, act, S, st);
    else launch_simt<128>(A, B, bias, C, workspace, M, N, K, a_kmajor, b_kmajor, act, S, st);
    MMVAE_LAUNCH_CHECK("gemm_simt");
    if (S > 1) {
        splitk_reduce_launch(workspace, bias, C, (long)M * N, N, S, act, st);
        MMVAE_LAUNCH_CHECK("splitk_reduce");
    }
    return 0;
}

// ---- skinny-K rows x wide-N: y[R,N] = act(x[R,K] * W + b) with K <= 32 (the latent dimension) ----------------------------------
// The first layer of every decoder (pythae Decoder_AE_MLP Linear(L,512); Decoder_VAE_SVHN's ConvTranspose2d(L,128,4,1,0) on a
// 1x1 input = Linear(L, 2048), encoders.py:116-118) at M*K*B rows: 20 FMAs per output, i.e. a STORE-bound kernel (63 MB out at
// C2) that the tiled GEMM above ran at 1 TB/s.  Here a thread keeps its 4 columns of W (K float4) in registers, the block's x
// rows sit in shared memory and are read as broadcasts, and every warp store covers 512 contiguous bytes.
#define SK_ROWS 64
#define SK_THREADS 256
template <int KMAX, bool B_KMAJOR>
__global__ void __launch_bounds__(SK_THREADS)
gemm_skinny_kernel(const float* __restrict__ X, const float* __restrict__ W, const float* __restrict__ bias, float* __restrict__ Y, int R,
                   int N, int K, int act) {
    __shared__ __align__(16) float xs[SK_ROWS][KMAX];
    const int n = (blockIdx.x * SK_THREADS + threadIdx.x) * 4;
    const int r0 = blockIdx.y * SK_ROWS;
    for (int i = threadIdx.x; i < SK_ROWS * KMAX; i += SK_THREADS) {
        const int r = i / KMAX, k = i - r * KMAX;
        xs[r][k] = (r0 + r < R && k < K) ? X[(long)(r0 + r) * K + k] : 0.f;
    }
    float4 w[KMAX];
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n < N) {
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            if (k < K) {
                if (B_KMAJOR) w[k] = make_float4(W[(long)n * K + k], W[(long)(n + 1) * K + k], W[(long)(n + 2) * K + k], W[(long)(n + 3) * K + k]);
                else w[k] = __ldg(reinterpret_cast<const float4*>(W + (long)k * N + n));
            } else {
                w[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        if (bias) b4 = make_float4(bias[n], bias[n + 1], bias[n + 2], bias[n + 3]);
    }
    __syncthreads();
    if (n >= N) return;
    const int rows = min(SK_ROWS, R - r0);
#pragma unroll 2
    for (int r = 0; r < rows; ++r) {
        // packed fp32 FMAs (FFMA2: two columns per instruction, same IEEE result per lane): the loop is issue-bound otherwise
        float2 a01 = make_float2(b4.x, b4.y), a23 = make_float2(b4.z, b4.w);
#pragma unroll
        for (int k = 0; k < KMAX; k += 4) {
            const float4 xv = *reinterpret_cast<const float4*>(&xs[r][k]);          // same address in every lane: broadcast
            const float xk[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 xx = make_float2(xk[q], xk[q]);
                a01 = __ffma2_rn(xx, make_float2(w[k + q].x, w[k + q].y), a01);
                a23 = __ffma2_rn(xx, make_float2(w[k + q].z, w[k + q].w), a23);
            }
        }
        float4 a = make_float4(a01.x, a01.y, a23.x, a23.y);
        a.x = apply_act(a.x, act); a.y = apply_act(a.y, act); a.z = apply_act(a.z, act); a.w = apply_act(a.w, act);
        __stcs(reinterpret_cast<float4*>(Y + (long)(r0 + r) * N + n), a);
    }
}

// returns -1 when the shape is not for this kernel
int mmvae_gemm_skinny(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor, int b_kmajor,
                      int act, cudaStream_t st) {
    if (!a_kmajor || K > 32 || K < 1 || (N & 3) || M < 512 || N < 256) return -1;
    if (((uintptr_t)C & 15) || (!b_kmajor && ((uintptr_t)B & 15))) return -1;
    const dim3 grid((N / 4 + SK_THREADS - 1) / SK_THREADS, (M + SK_ROWS - 1) / SK_ROWS);
    if (grid.y > 65535) return -1;
    if (K <= 20) {
        if (b_kmajor) gemm_skinny_kernel<20, true><<<grid, SK_THREADS, 0, st>>>(A, B, bias, C, M, N, K, act);
        else gemm_skinny_kernel<20, false><<<grid, SK_THREADS, 0, st>>>(A, B, bias, C, M, N, K, act);
    } else {
        if (b_kmajor) gemm_skinny_kernel<32, true><<<grid, SK_THREADS, 0, st>>>(A, B, bias, C, M, N, K, act);
        else gemm_skinny_kernel<32, false><<<grid, SK_THREADS, 0, st>>>(A, B, bias, C, M, N, K, act);
    }
    MMVAE_LAUNCH_CHECK("gemm_skinny");
    return 0;
}

// ---- small GEMMs: one launch, split-K inside a thread-block cluster ---------------------------------------------------------
// The encoders at batch 128, every MADE layer and every weight gradient of a 128-row layer are contractions of 10^6..10^8 MACs:
// latency-bound, 8-14 us each through the kernels above (a K loop on a few CTAs, or a split-K pass plus its reduce launch).
// Here the K range of one 64x64 output tile is split over the S <= 16 CTAs of a cluster (grid z = cluster z = S); each CTA
// leaves its partial tile in its own shared memory, and after one cluster barrier CTA r sums rows [r*64/S, (r+1)*64/S) of the S
// partials through distributed shared memory in rank order (deterministic), applies bias + activation and stores.  No
// workspace, no counters, no second launch.  Exact fp32 FMA arithmetic.
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define GS_BM 64
#define GS_BN 64
#define GS_BKG 128                     // k per shared-memory group: everything a CTA needs for 8 k-tiles is in flight at once
#define GS_LD (GS_BM + G_PAD)          // row stride of an operand group s[k][row]
#define GS_SMEM (2 * GS_BKG * GS_LD * (int)sizeof(float))

// One 64 x GS_BKG operand group in registers (8 float4 per thread), zero-filled outside [0,rows_total) x [k0,k_end).
// KMAJOR (src[row*ld + k]): a warp takes 8 rows x 64 contiguous bytes (whole sectors); the transposing store is 2-way conflicted.
template <bool KMAJOR>
struct GroupLoader {
    static constexpr int NV = GS_BM * GS_BKG / 4 / G_THREADS;
    float4 v[NV];
    __device__ __forceinline__ void load(const float* __restrict__ src, int ld, int row0, int rows_total, int k0, int k_end, bool vec_ok) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int f = threadIdx.x + i * G_THREADS;
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (KMAJOR) {
                const int row = (f / 4) % GS_BM, kq = ((f / (4 * GS_BM)) * 4 + (f % 4)) * 4;
                const int gr = row0 + row, gk = k0 + kq;
                if (gr < rows_total && gk < k_end) {
                    const float* p = src + (long)gr * ld + gk;
                    if (vec_ok && gk + 3 < k_end) t = __ldg(reinterpret_cast<const float4*>(p));
                    else {
                        t.x = p[0];
                        if (gk + 1 < k_end) t.y = p[1];
                        if (gk + 2 < k_end) t.z = p[2];
                        if (gk + 3 < k_end) t.w = p[3];
                    }
                }
            } else {
                const int k = f / (GS_BM / 4), rq = (f % (GS_BM / 4)) * 4;
                const int gk = k0 + k, gr = row0 + rq;
                if (gk < k_end && gr < rows_total) {
                    const float* p = src + (long)gk * ld + gr;
                    if (vec_ok && gr + 3 < rows_total) t = __ldg(reinterpret_cast<const float4*>(p));
                    else {
                        t.x = p[0];
                        if (gr + 1 < rows_total) t.y = p[1];
                        if (gr + 2 < rows_total) t.z = p[2];
                        if (gr + 3 < rows_total) t.w = p[3];
                    }
                }
            }
            v[i] = t;
        }
    }
    __device__ __forceinline__ void store(float* __restrict__ s) const {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int f = threadIdx.x + i * G_THREADS;
            if (KMAJOR) {
                const int row = (f / 4) % GS_BM, kq = ((f / (4 * GS_BM)) * 4 + (f % 4)) * 4;
                s[(kq + 0) * GS_LD + row] = v[i].x;
                s[(kq + 1) * GS_LD + row] = v[i].y;
                s[(kq + 2) * GS_LD + row] = v[i].z;
                s[(kq + 3) * GS_LD + row] = v[i].w;
            } else {
                const int k = f / (GS_BM / 4), rq = (f % (GS_BM / 4)) * 4;
                *reinterpret_cast<float4*>(&s[k * GS_LD + rq]) = v[i];
            }
        }
    }
};

// grid (N tiles, M tiles, S), cluster (1, 1, S).  CTA z contracts k in [z*k_per, (z+1)*k_per) (k_per a multiple of 16).
template <bool A_KMAJOR, bool B_KMAJOR>
__global__ void __launch_bounds__(G_THREADS, 2)
gemm_small_kernel(const float* __restrict__ A, const float* __restrict__ Bm, const float* __restrict__ bias, float* __restrict__ C,
                  int M, int N, int K, int act, int k_per) {
    extern __shared__ __align__(16) float sm[];                           // As[128][68] | Bs[128][68]; later red[64][64]
    float* As = sm;
    float* Bs = sm + GS_BKG * GS_LD;
    const int m0 = blockIdx.y * GS_BM, n0 = blockIdx.x * GS_BN;
    const int k_begin = min(K, (int)blockIdx.z * k_per), k_end = min(K, k_begin + k_per);
    const int lda = A_KMAJOR ? K : M, ldb = B_KMAJOR ? K : N;
    const bool a_vec = (((uintptr_t)A & 15) == 0) && (lda % 4 == 0);
    const bool b_vec = (((uintptr_t)Bm & 15) == 0) && (ldb % 4 == 0);
    const int ty = threadIdx.x / 16, tx = threadIdx.x % 16;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    GroupLoader<A_KMAJOR> la;
    GroupLoader<B_KMAJOR> lb;
    if (k_begin < k_end) {
        la.load(A, lda, m0, M, k_begin, k_end, a_vec);
        lb.load(Bm, ldb, n0, N, k_begin, k_end, b_vec);
        la.store(As);
        lb.store(Bs);
    }
    __syncthreads();
    for (int k0 = k_begin; k0 < k_end; k0 += GS_BKG) {
        const bool more = (k0 + GS_BKG < k_end);
        if (more) {                                                      // the next group's loads fly under this group's FMAs
            la.load(A, lda, m0, M, k0 + GS_BKG, k_end, a_vec);
            lb.load(Bm, ldb, n0, N, k0 + GS_BKG, k_end, b_vec);
        }
        const int kn = min(GS_BKG, (k_end - k0 + 15) & ~15);
        const float* as = As + ty * 4;
        const float* bs = Bs + tx * 4;
#pragma unroll 1
        for (int kb = 0; kb < kn; kb += 16) {
#pragma unroll
            for (int kk = 0; kk < 16; ++kk) {
                const float4 a = *reinterpret_cast<const float4*>(as + (kb + kk) * GS_LD);
                const float4 b = *reinterpret_cast<const float4*>(bs + (kb + kk) * GS_LD);
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
        }
        __syncthreads();
        if (more) {
            la.store(As);
            lb.store(Bs);
            __syncthreads();
        }
    }

    const int S = gridDim.z;
    const bool c_vec = (((uintptr_t)C & 15) == 0) && (N % 4 == 0);
    if (S == 1) {
        const int gn = n0 + tx * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int gm = m0 + ty * 4 + i;
            if (gm >= M) continue;
            float r[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) r[j] = (gn + j < N) ? apply_act(acc[i][j] + (bias ? bias[gn + j] : 0.f), act) : 0.f;
            float* p = C + (long)gm * N + gn;
            if (c_vec && gn + 3 < N) *reinterpret_cast<float4*>(p) = make_float4(r[0], r[1], r[2], r[3]);
            else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (gn + j < N) p[j] = r[j];
            }
        }
        return;
    }
    // the loop's trailing __syncthreads() has retired every read of the operand groups: reuse them for the partial tile
    cg::cluster_group cluster = cg::this_cluster();
    float4* red = reinterpret_cast<float4*>(sm);                         // [64 rows][16 float4]
#pragma unroll
    for (int i = 0; i < 4; ++i) red[(ty * 4 + i) * 16 + tx] = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    cluster.sync();
    const int rank = (int)cluster.block_rank(), rows_per = GS_BM / S;
    for (int e = threadIdx.x; e < rows_per * 16; e += G_THREADS) {
        const int row = rank * rows_per + e / 16, c4 = e % 16;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int z = 0; z < S; ++z) {
            const float4 v = cluster.map_shared_rank(red, z)[row * 16 + c4];
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        const int gm = m0 + row, gn = n0 + c4 * 4;
        if (gm < M && gn < N) {
            float r[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) r[j] = (gn + j < N) ? apply_act(r[j] + (bias ? bias[gn + j] : 0.f), act) : 0.f;
            float* p = C + (long)gm * N + gn;
            if (c_vec && gn + 3 < N) *reinterpret_cast<float4*>(p) = make_float4(r[0], r[1], r[2], r[3]);
            else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (gn + j < N) p[j] = r[j];
            }
        }
    }
    cluster.sync();                                                       // nobody leaves while a peer may still read its partial
}

static int g_small_max_split = 16;
static long g_small_max_macs = 64L << 20;                                 // above this the tensor-core / tiled kernels win (measured)
extern "C" long mmvae_gemm_small_set_limit(long macs) { const long o = g_small_max_macs; g_small_max_macs = macs; return o; }
extern "C" int mmvae_gemm_small_set_max_split(int s) { const int o = g_small_max_split; g_small_max_split = s; return o; }

// S: power of two <= 16, enough CTAs to fill the machine twice, at least two k-tiles per CTA
static int small_splits(long tiles, int nk) {
    const int cap = tiles <= 4 ? g_small_max_split : (g_small_max_split < 8 ? g_small_max_split : 8);   // 16-CTA clusters pay only for 1-4 tiles (measured)
    int S = 1;
    while (S < cap && tiles * S < 2L * MMVAE_NUM_SMS && nk / (2 * S) >= 2) S *= 2;
    return S;
}

template <bool AK, bool BK>
static int small_launch(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int act, dim3 grid, int k_per,
                        cudaStream_t st) {
    static bool attr_done = false;
    auto kern = gemm_small_kernel<AK, BK>;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GS_SMEM);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "gemm_small: kernel attributes: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(G_THREADS); cfg.dynamicSmemBytes = GS_SMEM; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = grid.z;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, A, B, bias, C, M, N, K, act, k_per);
    if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "gemm_small: launch: %s", cudaGetErrorString(e));
    return 0;
}

// returns -1 when the shape is not for this kernel
int mmvae_gemm_small(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor, int b_kmajor,
                     int act, cudaStream_t st) {
    if ((long)M * N * K > g_small_max_macs) return -1;
    const long tm = (M + GS_BM - 1) / GS_BM, tn = (N + GS_BN - 1) / GS_BN;
    if (tm * tn > MMVAE_NUM_SMS && g_small_max_macs < (1L << 40)) return -1;    // enough tiles to fill the machine: the tiled kernels
    if (tm > 65535 || tn > 65535) return -1;
    const int nk = (K + G_BK - 1) / G_BK;
    const int S = small_splits(tm * tn, nk);
    const int tps = ((nk + S - 1) / S) * G_BK;                             // k per CTA
    if (tps > 4 * GS_BKG && g_small_max_macs < (1L << 40)) return -1;      // a long serial K loop on a few CTAs: the split-K kernels
    const dim3 grid((unsigned)tn, (unsigned)tm, (unsigned)S);
    if (a_kmajor && b_kmajor) return small_launch<true, true>(A, B, bias, C, M, N, K, act, grid, tps, st);
    if (a_kmajor && !b_kmajor) return small_launch<true, false>(A, B, bias, C, M, N, K, act, grid, tps, st);
    if (!a_kmajor && b_kmajor) return small_launch<false, true>(A, B, bias, C, M, N, K, act, grid, tps, st);
    return small_launch<false, false>(A, B, bias, C, M, N, K, act, grid, tps, st);
}
