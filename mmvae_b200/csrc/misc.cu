// Library plumbing (error reporting, ABI version), the GEMM dispatcher, and the small elementwise / reduction
// kernels around the contractions: bias gradients, activation backward, MADE weight masking, affine flow
// transforms, fused Adam(amsgrad).
#include <stdarg.h>
#include "common.cuh"

char g_mmvae_err[512] = "";

int mmvae_set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_mmvae_err, sizeof(g_mmvae_err), fmt, ap);
    va_end(ap);
    return code;
}

extern "C" const char* mmvae_last_error(void) { return g_mmvae_err; }
extern "C" int mmvae_abi_version(void) { return MMVAE_ABI_VERSION; }

int mmvae_gemm_simt(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor,
                    int b_kmajor, int act, float* workspace, long workspace_bytes, cudaStream_t st);
int mmvae_gemm_tc(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor,
                  int b_kmajor, int act, float* workspace, long workspace_bytes, cudaStream_t st);
int mmvae_gemm_skinny(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor, int b_kmajor,
                      int act, cudaStream_t st);
int mmvae_gemm_small(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor, int b_kmajor,
                     int act, cudaStream_t st);
long mmvae_gemm_simt_workspace_bytes(int M, int N, int K);
long mmvae_gemm_tc_workspace_bytes(int M, int N, int K, int a_k, int b_k);

extern "C" long mmvae_gemm_workspace_bytes(int M, int N, int K, int a_kmajor, int b_kmajor) {
    const long a = mmvae_gemm_simt_workspace_bytes(M, N, K), b = mmvae_gemm_tc_workspace_bytes(M, N, K, a_kmajor, b_kmajor);
    return a > b ? a : b;
}

static int g_last_gemm_kernel = -1;
extern "C" int mmvae_gemm_last_kernel(void) { return g_last_gemm_kernel; }

extern "C" int mmvae_gemm(const float* A, const float* B, const float* bias, float* C, int M, int N, int K,
                          int a_kmajor, int b_kmajor, int act, float* workspace, long workspace_bytes, int allow_tc,
                          void* stream) {
    MMVAE_REQUIRE(A && B && C, "gemm: null pointer");
    MMVAE_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad sizes M=%d N=%d K=%d", M, N, K);
    MMVAE_REQUIRE(act >= 0 && act <= 2, "gemm: bad activation %d", act);
    cudaStream_t st = (cudaStream_t)stream;
    {
        const int rc = mmvae_gemm_skinny(A, B, bias, C, M, N, K, a_kmajor, b_kmajor, act, st);     // K = latent dimension
        if (rc != -1) { g_last_gemm_kernel = 2; return rc; }
    }
    {
        const int rc = mmvae_gemm_small(A, B, bias, C, M, N, K, a_kmajor, b_kmajor, act, st);      // latency-bound sizes
        if (rc != -1) { g_last_gemm_kernel = 3; return rc; }
    }
    if (allow_tc) {
        const int rc = mmvae_gemm_tc(A, B, bias, C, M, N, K, a_kmajor, b_kmajor, act, workspace, workspace_bytes, st);
        if (rc != -1) { g_last_gemm_kernel = 1; return rc; }
    }
    g_last_gemm_kernel = 0;
    return mmvae_gemm_simt(A, B, bias, C, M, N, K, a_kmajor, b_kmajor, act, workspace, workspace_bytes, st);
}

// ---- column sums (bias gradients): out[n] = sum_m A[m,n].  Two deterministic stages. ---------------------------
// stage 1: grid (column groups of 32, P row-slabs); 256 threads = 8 row-lanes x 32 columns, each block strides over its
// slab.  P is capped so the second stage stays short and the first fills the machine; partials are [P][N].
__global__ void __launch_bounds__(256)
colsum_stage1(const float* __restrict__ A, float* __restrict__ part, int M, int N, int rows_per_slab) {
    __shared__ float sm[8][33];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int n = blockIdx.x * 32 + cx;
    const int m_begin = blockIdx.y * rows_per_slab;
    const int m_end = min(M, m_begin + rows_per_slab);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (n < N) {
        int m = m_begin + ry;
        for (; m + 24 < m_end; m += 32) {                     // 4 independent loads in flight
            s0 += A[(long)m * N + n];
            s1 += A[(long)(m + 8) * N + n];
            s2 += A[(long)(m + 16) * N + n];
            s3 += A[(long)(m + 24) * N + n];
        }
        for (; m < m_end; m += 8) s0 += A[(long)m * N + n];
    }
    sm[ry][cx] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (ry == 0 && n < N) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) t += sm[j][cx];
        part[(long)blockIdx.y * N + n] = t;
    }
}
// stage 2: one block per 32 columns, 8 row-lanes over the P partials
__global__ void __launch_bounds__(256)
colsum_stage2(const float* __restrict__ part, float* __restrict__ out, int P, int N) {
    __shared__ float sm[8][33];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int n = blockIdx.x * 32 + cx;
    float s = 0.f;
    if (n < N)
        for (int p = ry; p < P; p += 8) s += part[(long)p * N + n];
    sm[ry][cx] = s;
    __syncthreads();
    if (ry == 0 && n < N) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) t += sm[j][cx];
        out[n] = t;
    }
}

static int colsum_slabs(int M, int N) {
    const int groups = (N + 31) / 32;
    int P = (4 * MMVAE_NUM_SMS + groups - 1) / groups;         // ~4 blocks per SM in total
    const int maxP = (M + 255) / 256;                          // at least 256 rows per slab
    if (P > maxP) P = maxP;
    return P < 1 ? 1 : P;
}
extern "C" long mmvae_colsum_workspace_bytes(int M, int N) {
    return (long)colsum_slabs(M, N) * N * (long)sizeof(float);
}
extern "C" int mmvae_colsum(const float* A, float* out, float* workspace, int M, int N, void* stream) {
    MMVAE_REQUIRE(A && out && workspace && M > 0 && N > 0, "colsum: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const int P = colsum_slabs(M, N);
    const int rows_per_slab = (((M + P - 1) / P) + 7) & ~7;
    colsum_stage1<<<dim3((N + 31) / 32, P), 256, 0, st>>>(A, workspace, M, N, rows_per_slab);
    MMVAE_LAUNCH_CHECK("colsum_stage1");
    colsum_stage2<<<(N + 31) / 32, 256, 0, st>>>(workspace, out, P, N);
    MMVAE_LAUNCH_CHECK("colsum_stage2");
    return 0;
}

// ---- column sums without any allocation: the stage-1 partials live in a file-scope device array -----------------
// For callers that may not allocate (CUDA-graph capture, serving loops).  Every GPU that loads the library holds the
// array (1 MiB) until exit.  All calls share it: two calls that overlap on the GPU (different streams) race.  Nothing
// in it has to start at zero: stage 1 writes every partial [P][N] that stage 2 reads, so stale contents of an earlier
// call are never observed.
#define MMVAE_COLSUM_STATIC_MAX_FLOATS (1 << 18)
static __device__ __align__(256) float g_colsum_scratch[MMVAE_COLSUM_STATIC_MAX_FLOATS];
#define MMVAE_MAX_DEVICES 64
static float* g_colsum_scratch_ptr[MMVAE_MAX_DEVICES];         // device address of the array, per device (0 = not set up)

// In host code `g_colsum_scratch` is only a stand-in for the symbol: the device address comes from the runtime, once per
// device, with that device current.
static int colsum_static_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "colsum_static: cudaGetDevice: %s", cudaGetErrorString(e));
    MMVAE_REQUIRE(dev >= 0 && dev < MMVAE_MAX_DEVICES, "colsum_static: device ordinal %d out of range", dev);
    float* p = __atomic_load_n(&g_colsum_scratch_ptr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        e = cudaGetSymbolAddress((void**)&p, g_colsum_scratch);
        if (e != cudaSuccess || !p)
            return mmvae_set_error(1000 + (int)e, "colsum_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        __atomic_store_n(&g_colsum_scratch_ptr[dev], p, __ATOMIC_RELEASE);   // same value from every thread
    }
    *out = p;
    return 0;
}
extern "C" long mmvae_colsum_static_max_bytes(void) { return (long)MMVAE_COLSUM_STATIC_MAX_FLOATS * (long)sizeof(float); }
extern "C" int mmvae_colsum_static(const float* A, float* out, int M, int N, void* stream) {
    MMVAE_REQUIRE(A && out && M > 0 && N > 0, "colsum_static: bad arguments");
    const int P = colsum_slabs(M, N);
    MMVAE_REQUIRE((long)P * N <= (long)MMVAE_COLSUM_STATIC_MAX_FLOATS,
                  "colsum_static: M=%d N=%d needs %ld scratch bytes, the static scratch holds %ld; use mmvae_colsum with a "
                  "caller-provided workspace", M, N, (long)P * N * (long)sizeof(float), mmvae_colsum_static_max_bytes());
    float* scratch = nullptr;
    const int rc = colsum_static_scratch(&scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int rows_per_slab = (((M + P - 1) / P) + 7) & ~7;
    colsum_stage1<<<dim3((N + 31) / 32, P), 256, 0, st>>>(A, scratch, M, N, rows_per_slab);
    MMVAE_LAUNCH_CHECK("colsum_static stage1");
    colsum_stage2<<<(N + 31) / 32, 256, 0, st>>>(scratch, out, P, N);
    MMVAE_LAUNCH_CHECK("colsum_static stage2");
    return 0;
}

// ---- activation backward, in place: g *= act'(y) ---------------------------------------------------------------
__global__ void act_bwd_kernel(const float* g, const float* __restrict__ y, float* out, long n, int act) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float v = y[i], gi = g[i];
        out[i] = (act == MMVAE_ACT_RELU) ? (v > 0.f ? gi : 0.f) : gi * v * (1.f - v);
    }
}
extern "C" int mmvae_act_bwd(const float* g, const float* y, float* out, long n, int act, void* stream) {
    MMVAE_REQUIRE(g && y && out && n > 0, "act_bwd: bad arguments");
    MMVAE_REQUIRE(act == MMVAE_ACT_RELU || act == MMVAE_ACT_SIGMOID, "act_bwd: bad activation %d", act);
    const int grid = (int)min((long)MMVAE_NUM_SMS * 16, (n + 255) / 256);
    act_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g, y, out, n, act);
    MMVAE_LAUNCH_CHECK("act_bwd");
    return 0;
}

// ---- MADE: out = w * mask (layers.py:24-25), used for the weight and for its gradient ---------------------------
__global__ void mask_weight_kernel(const float* __restrict__ w, const float* __restrict__ mask, float* __restrict__ out,
                                   long n) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        out[i] = w[i] * mask[i];
}
extern "C" int mmvae_mask_weight(const float* w, const float* mask, float* out, long n, void* stream) {
    MMVAE_REQUIRE(w && mask && out && n > 0, "mask_weight: bad arguments");
    const int grid = (int)min((long)MMVAE_NUM_SMS * 4, (n + 255) / 256);
    mask_weight_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w, mask, out, n);
    MMVAE_LAUNCH_CHECK("mask_weight");
    return 0;
}

// ---- fused Adam / AMSGrad over one flat parameter buffer (torch.optim.Adam semantics; main.py:114-115) ----------
__global__ void adam_step_kernel(float* step) { step[0] += 1.f; }
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, float* __restrict__ vmax, long n, float lr, float b1, float b2,
                            float eps, const float* __restrict__ step, int amsgrad) {
    const float t = step[0];                                   // device-resident step count: CUDA-graph replayable
    const float bc1 = 1.f - powf(b1, t);
    const float bc2_sqrt = sqrtf(1.f - powf(b2, t));
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float gi = g[i];
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        float vv = vi;
        if (amsgrad) {
            vv = fmaxf(vmax[i], vi);
            vmax[i] = vv;
        }
        const float denom = sqrtf(vv) / bc2_sqrt + eps;
        p[i] -= (lr / bc1) * (mi / denom);
    }
}
extern "C" int mmvae_adam_amsgrad(float* p, const float* g, float* m, float* v, float* vmax, long n, float lr,
                                  float beta1, float beta2, float eps, float* step_dev, int amsgrad, void* stream) {
    MMVAE_REQUIRE(p && g && m && v && (vmax || !amsgrad) && n > 0 && step_dev, "adam: bad arguments");
    const int grid = (int)min((long)MMVAE_NUM_SMS * 8, (n + 255) / 256);
    adam_step_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev);
    adam_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, vmax, n, lr, beta1, beta2, eps, step_dev, amsgrad);
    MMVAE_LAUNCH_CHECK("adam");
    return 0;
}
