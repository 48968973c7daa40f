// The SEQUENTIAL direction of a MADE flow block (MAF.inverse, maf_model.py:91-104; IAF.forward, iaf_model.py:60-78) as ONE
// kernel per block.  The reference (and the GEMM formulation of flows.py) runs L full MADE passes, one per latent
// column, because column i of the output needs the MADE output of the columns already produced.  With the MADE's
// sequential degrees a hidden unit of degree d depends only on columns <= d, so it is final as soon as column d is:
// every hidden unit is evaluated exactly ONCE over the whole sweep (1/L of the reference's arithmetic) and the hidden
// activations never leave the SM.
//
// Layout: a warp owns 16 rows, lane = (k-half, row): the two half-warps split every dot product over even / odd k and
// add their halves with one shuffle.  The packed, degree-sorted weights of the block (~100 KB for L=20, H=128) sit in
// shared memory and are read as half-warp broadcasts; a row's activations h1,h2,h3 [H] and outputs x [L] live in a
// warp-private column of shared memory ([k][row]: conflict-free), so after the weight load only __syncwarp is needed
// and four warps fit beside the weights (one per scheduler).
// Units of degree d form group d (<= 8 units, padded to 8 accumulators): group d of layer 1 reads x[0..d], of layers
// 2 / 3 the first cnt[d] units of the layer below, and output column i the first cnt[i-1] units of h3.
//
// Pack (floats, built by mmvae_b200/ops.py:made_seq_pack from the masked weights):
//   b1g,b2g,b3g [G][8] | b4 [pad4(2L)] | W1g: group d -> [d+1][8] | W2g, W3g: group d -> [cnt[d]][8] | W4g: column i ->
//   [cnt[i-1]][2] (mu, log_var)          with G = L-1 groups and cnt[d] = number of hidden units of degree <= d.
#include "common.cuh"

#define MS_MAX_WARPS 4
#define MS_ROWS 16                            // rows per warp: lane = (k-half, row)
#define MS_MAXL 32

__global__ void __launch_bounds__(MS_MAX_WARPS * 32, 1)
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
    // A warp owns 16 rows; its two half-warps split every dot product over k (even / odd k) and add the halves with one
    // shuffle, so four warps fit beside the weights and each scheduler of the SM has a warp to issue from.
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, half = lane >> 4, rl = lane & 15;
    const int rows_cta = (blockDim.x >> 5) * MS_ROWS;
    float* h1 = sm + pack_floats + warp * (3 * H + L) * MS_ROWS + rl;
    float* h2 = h1 + H * MS_ROWS;
    float* h3 = h2 + H * MS_ROWS;
    float* xs = h3 + H * MS_ROWS;

    for (long tile = blockIdx.x; tile * rows_cta < R; tile += gridDim.x) {
        const long r = tile * rows_cta + warp * MS_ROWS + rl;
        const bool live = r < R;
        const long rr = live ? r : R - 1;
        float ladj = ladj_in[rr];
        for (int i = 0; i < L; ++i) {
            // column i: (mu, log_var) from the h3 units of degree < i
            const int n = cnt[i];
            float mu = half ? 0.f : b4[i], lv = half ? 0.f : b4[L + i];
            const float2* w4 = reinterpret_cast<const float2*>(W4g + o4[i]);
#pragma unroll 4
            for (int k = half; k < n; k += 2) {
                const float h = h3[k * MS_ROWS];
                const float2 w = w4[k];
                mu = fmaf(w.x, h, mu);
                lv = fmaf(w.y, h, lv);
            }
            mu += __shfl_xor_sync(0xffffffffu, mu, 16);
            lv += __shfl_xor_sync(0xffffffffu, lv, 16);
            const float yi = y[rr * L + (flip_in ? L - 1 - i : i)];
            float u;
            if (mode == 0) { u = (yi - mu) * expf(-lv); ladj -= lv; }
            else           { u = yi * expf(lv) + mu;    ladj += lv; }
            if (!half) {
                xs[i * MS_ROWS] = u;
                if (live) out[r * L + (flip_out ? L - 1 - i : i)] = u;
            }
            __syncwarp();
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
                for (int q = 0; q < 8; ++q) acc[q] = half ? 0.f : bg[q];
#pragma unroll 4
                for (int k = half; k < kn; k += 2) {
                    const float v = src[k * MS_ROWS];
                    const float4 wa = *reinterpret_cast<const float4*>(W + 8 * k);
                    const float4 wb = *reinterpret_cast<const float4*>(W + 8 * k + 4);
                    acc[0] = fmaf(wa.x, v, acc[0]); acc[1] = fmaf(wa.y, v, acc[1]);
                    acc[2] = fmaf(wa.z, v, acc[2]); acc[3] = fmaf(wa.w, v, acc[3]);
                    acc[4] = fmaf(wb.x, v, acc[4]); acc[5] = fmaf(wb.y, v, acc[5]);
                    acc[6] = fmaf(wb.z, v, acc[6]); acc[7] = fmaf(wb.w, v, acc[7]);
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) acc[q] += __shfl_xor_sync(0xffffffffu, acc[q], 16);
#pragma unroll
                for (int qq = 0; qq < 4; ++qq) {           // half h stores the units q = 2*qq + h
                    const float v = half ? acc[2 * qq + 1] : acc[2 * qq];
                    const int q = 2 * qq + half;
                    if (q < nu) dst[(u0 + q) * MS_ROWS] = fmaxf(v, 0.f);
                }
                __syncwarp();
            }
        }
        if (live && !half) ladj_out[r] = ladj;
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
    return (made_seq_pack_floats_host(cnt_host, L) + (long)warps * (3 * H + L) * MS_ROWS) * 4;
}

static int made_seq_warps(const int* cnt_host, int L, int H) {
    for (int w = MS_MAX_WARPS; w > 1; w >>= 1)
        if (made_seq_smem(cnt_host, L, H, w) <= 227 * 1024) return w;
    return 1;
}

// shared memory of the launch configuration the library picks: four warps (16 rows each) per CTA, fewer when that does not fit
extern "C" long mmvae_made_seq_smem_bytes(const int* cnt_host, int L, int H) {
    if (!cnt_host || L < 2 || L > MS_MAXL) return -1;
    return made_seq_smem(cnt_host, L, H, made_seq_warps(cnt_host, L, H));
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
    const int warps = made_seq_warps(cnt_host, L, H);
    const int threads = 32 * warps;
    const int tiles = ceil_div_i(R, warps * MS_ROWS);
    const int grid = tiles < MMVAE_NUM_SMS ? tiles : MMVAE_NUM_SMS;
    made_seq_kernel<<<grid, threads, smem, (cudaStream_t)stream>>>(y, ladj_in, out, ladj_out, pack, (int)pf, cnt_dev, R, L, H,
                                                                      mode, flip_in, flip_out);
    MMVAE_LAUNCH_CHECK("made_seq_fwd");
    return 0;
}
