// The SEQUENTIAL direction of a MADE flow block (MAF.inverse, maf_model.py:91-104; IAF.forward, iaf_model.py:60-78) as ONE
// kernel per block.  The reference (and the GEMM formulation of flows.py) runs L full MADE passes, one per latent
// column, because column i of the output needs the MADE output of the columns already produced.  With the MADE's
// sequential degrees a hidden unit of degree d depends only on columns <= d, so it is final as soon as column d is:
// every hidden unit is evaluated exactly ONCE over the whole sweep (1/L of the reference's arithmetic) and the hidden
// activations never leave the SM.
//
// Layout: one thread = one row; the packed, degree-sorted weights of the block (~100 KB for L=20, H=128) sit in shared
// memory and are read as warp-wide broadcasts; the row's activations h1,h2,h3 [H] and outputs x [L] live in a
// lane-private column of shared memory ([k][lane]: conflict-free), so the kernel needs no barrier after the weight load.
// Units of degree d form group d (<= 8 units, padded to 8 accumulators): group d of layer 1 reads x[0..d], of layers
// 2 / 3 the first cnt[d] units of the layer below, and output column i the first cnt[i-1] units of h3.
//
// Pack (floats, built by mmvae_b200/ops.py:made_seq_pack from the masked weights):
//   b1g,b2g,b3g [G][8] | b4 [pad4(2L)] | W1g: group d -> [d+1][8] | W2g, W3g: group d -> [cnt[d]][8] | W4g: column i ->
//   [cnt[i-1]][2] (mu, log_var)          with G = L-1 groups and cnt[d] = number of hidden units of degree <= d.
#include "common.cuh"

#define MS_THREADS 64
#define MS_MAXL 32

__global__ void __launch_bounds__(MS_THREADS, 1)   // launched with 64 threads, or 32 when two warps' activations do not fit
made_seq_kernel(const float* __restrict__ y, const float* __restrict__ ladj_in, float* __restrict__ out,
                float* __restrict__ ladj_out, const float* __restrict__ pack, int pack_floats, const int* __restrict__ cnt_g,
                int R, int L, int H, int mode, int flip_in, int flip_out) {
    extern __shared__ __align__(16) float sm[];
    __shared__ int cnt[MS_MAXL + 1];          // cnt[0] = 0, cnt[d + 1] = units of degree <= d
    __shared__ int o2[MS_MAXL], o4[MS_MAXL + 1];
    const int G = L - 1;
    if (threadIdx.x == 0) {
        cnt[0] = 0;
        for (int d = 0; d < G; ++d) cnt[d + 1] = cnt_g[d];
        int a = 0, b = 0;
        for (int d = 0; d < G; ++d) { o2[d] = a; a += 8 * cnt[d + 1]; }
        for (int i = 0; i < L; ++i) { o4[i] = b; b += 2 * cnt[i]; }
        o2[G] = a;                             // total of W2g / W3g (MS_MAXL >= L > G)
    }
    for (int i = threadIdx.x * 4; i < pack_floats; i += blockDim.x * 4)
        *reinterpret_cast<float4*>(sm + i) = *reinterpret_cast<const float4*>(pack + i);
    __syncthreads();
    const float* b1g = sm;
    const float* b2g = b1g + 8 * G;
    const float* b3g = b2g + 8 * G;
    const float* b4 = b3g + 8 * G;
    const float* W1g = b4 + ((2 * L + 3) & ~3);
    const float* W2g = W1g + 8 * (G * (G + 1) / 2);
    const float* W3g = W2g + o2[G];
    const float* W4g = W3g + o2[G];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* h1 = sm + pack_floats + warp * (3 * H + L) * 32 + lane;
    float* h2 = h1 + H * 32;
    float* h3 = h2 + H * 32;
    float* xs = h3 + H * 32;

    for (long tile = blockIdx.x; tile * blockDim.x < R; tile += gridDim.x) {
        const long r = tile * blockDim.x + threadIdx.x;
        const bool live = r < R;
        const long rr = live ? r : R - 1;
        float ladj = ladj_in[rr];
        for (int i = 0; i < L; ++i) {
            // column i: (mu, log_var) from the h3 units of degree < i
            const int n = cnt[i];
            float mu = b4[i], lv = b4[L + i];
            const float2* w4 = reinterpret_cast<const float2*>(W4g + o4[i]);
#pragma unroll 4
            for (int k = 0; k < n; ++k) {
                const float h = h3[k * 32];
                const float2 w = w4[k];
                mu = fmaf(w.x, h, mu);
                lv = fmaf(w.y, h, lv);
            }
            const float yi = y[rr * L + (flip_in ? L - 1 - i : i)];
            float u;
            if (mode == 0) { u = (yi - mu) * expf(-lv); ladj -= lv; }
            else           { u = yi * expf(lv) + mu;    ladj += lv; }
            xs[i * 32] = u;
            if (live) out[r * L + (flip_out ? L - 1 - i : i)] = u;
            if (i >= G) continue;
            // the hidden units of degree i, layer by layer
            const int u0 = cnt[i], n1 = cnt[i + 1], nu = n1 - u0;
            float acc[8];
#pragma unroll
            for (int layer = 0; layer < 3; ++layer) {
                const float* bg = (layer == 0 ? b1g : layer == 1 ? b2g : b3g) + 8 * i;
                const float* W = layer == 0 ? W1g + 8 * (i * (i + 1) / 2) : (layer == 1 ? W2g : W3g) + o2[i];
                const float* src = layer == 0 ? xs : layer == 1 ? h1 : h2;
                float* dst = layer == 0 ? h1 : layer == 1 ? h2 : h3;
                const int kn = layer == 0 ? i + 1 : n1;
#pragma unroll
                for (int q = 0; q < 8; ++q) acc[q] = bg[q];
#pragma unroll 4
                for (int k = 0; k < kn; ++k) {
                    const float v = src[k * 32];
                    const float4 wa = *reinterpret_cast<const float4*>(W + 8 * k);
                    const float4 wb = *reinterpret_cast<const float4*>(W + 8 * k + 4);
                    acc[0] = fmaf(wa.x, v, acc[0]); acc[1] = fmaf(wa.y, v, acc[1]);
                    acc[2] = fmaf(wa.z, v, acc[2]); acc[3] = fmaf(wa.w, v, acc[3]);
                    acc[4] = fmaf(wb.x, v, acc[4]); acc[5] = fmaf(wb.y, v, acc[5]);
                    acc[6] = fmaf(wb.z, v, acc[6]); acc[7] = fmaf(wb.w, v, acc[7]);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q < nu) dst[(u0 + q) * 32] = fmaxf(acc[q], 0.f);
            }
        }
        if (live) ladj_out[r] = ladj;
    }
}

static long made_seq_pack_floats_host(const int* cnt, int L) {
    const int G = L - 1;
    long s = 0, s4 = 0;
    for (int d = 0; d < G; ++d) s += cnt[d];
    for (int i = 1; i < L; ++i) s4 += cnt[i - 1];
    long n = 3L * 8 * G + ((2 * L + 3) & ~3) + 8L * (G * (G + 1) / 2) + 2L * 8 * s + 2L * s4;
    return (n + 3) & ~3L;
}

extern "C" long mmvae_made_seq_pack_floats(const int* cnt_host, int L) {
    if (!cnt_host || L < 2 || L > MS_MAXL) return -1;
    return made_seq_pack_floats_host(cnt_host, L);
}

static long made_seq_smem(const int* cnt_host, int L, int H, int warps) {
    return (made_seq_pack_floats_host(cnt_host, L) + (long)warps * (3 * H + L) * 32) * 4;
}

// shared memory of the launch configuration the library picks: two warps per CTA, one when that does not fit
extern "C" long mmvae_made_seq_smem_bytes(const int* cnt_host, int L, int H) {
    if (!cnt_host || L < 2 || L > MS_MAXL) return -1;
    const long two = made_seq_smem(cnt_host, L, H, 2);
    return two <= 227 * 1024 ? two : made_seq_smem(cnt_host, L, H, 1);
}

extern "C" int mmvae_made_seq_fwd(const float* y, const float* ladj_in, float* out, float* ladj_out, const float* pack,
                                  const int* cnt_dev, const int* cnt_host, int R, int L, int H, int mode, int flip_in,
                                  int flip_out, void* stream) {
    MMVAE_REQUIRE(y && ladj_in && out && ladj_out && pack && cnt_dev && cnt_host, "made_seq_fwd: null pointer");
    MMVAE_REQUIRE(R > 0 && L >= 2 && L <= MS_MAXL && H > 0 && (mode == 0 || mode == 1), "made_seq_fwd: bad sizes R=%d L=%d H=%d", R, L, H);
    MMVAE_REQUIRE(y != out, "made_seq_fwd: in-place call");
    int prev = 0;
    for (int d = 0; d < L - 1; ++d) {
        MMVAE_REQUIRE(cnt_host[d] >= prev && cnt_host[d] - prev <= 8 && cnt_host[d] <= H,
                      "made_seq_fwd: degree group %d has %d units (at most 8 per degree)", d, cnt_host[d] - prev);
        prev = cnt_host[d];
    }
    const long pf = made_seq_pack_floats_host(cnt_host, L);
    const long smem = mmvae_made_seq_smem_bytes(cnt_host, L, H);
    MMVAE_REQUIRE(smem <= 227 * 1024, "made_seq_fwd: %ld bytes of shared memory needed (L=%d H=%d)", smem, L, H);
    static long attr_set = 0;
    if (smem > attr_set) {
        cudaError_t e = cudaFuncSetAttribute(made_seq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "made_seq_fwd: %s", cudaGetErrorString(e));
        attr_set = smem;
    }
    const int threads = smem == made_seq_smem(cnt_host, L, H, 2) ? 64 : 32;
    const int tiles = ceil_div_i(R, threads);
    const int grid = tiles < MMVAE_NUM_SMS ? tiles : MMVAE_NUM_SMS;
    made_seq_kernel<<<grid, threads, smem, (cudaStream_t)stream>>>(y, ladj_in, out, ladj_out, pack, (int)pf, cnt_dev, R, L, H,
                                                                      mode, flip_in, flip_out);
    MMVAE_LAUNCH_CHECK("made_seq_fwd");
    return 0;
}
