// Last decoder layer on the tensor cores: ConvTranspose2d(32 -> COUT <= 4, k4, s2, p1) + bias + activation on 16x16
// inputs (encoders.py:124-136, `ConvTranspose2d(fBase, 3, 4, 2, 1) + Sigmoid`), NHWC input, NCHW output.
//
// The layer is too thin for an implicit GEMM with Cout as the N dimension (N = 3), and as a direct SIMT kernel it is
// bound by the fp32 FMA pipe (convt_small.cu: 198 us for 7680 images, 48 % FMA-pipe utilisation, HBM floor 53 us).  Here
// the GEMM is the LOWERED form kept entirely on the SM:
//     col[pixel, (ky,kx,co)] = sum_ci x[pixel, ci] * w[ci, co, ky, kx]          M = 128 pixels, N = 16*COUT, K = 32
// one 128-pixel tile = half an input image, K = Cin = one 128-byte swizzle row, so a tile is ONE TMA box and 4 k-steps
// x 3 passes (3xTF32) = 12 UMMAs of 32 + N/4 clocks.  The col block never leaves the SM: each epilogue thread owns one
// input pixel (one TMEM lane) and adds its 16*COUT values into the output image held in shared memory
// (y[co, 2iy-1+ky, 2ix-1+kx]) in four barrier-separated tap phases (fixed summation order: deterministic); after the
// two tiles of an image four finalize warps add the bias, apply the activation, stream the finished NCHW image out with
// coalesced 128-bit stores and zero the buffer again, while the epilogue warps scatter the next image into a second one.
//
// Warp roles (448 threads, one persistent CTA per SM): warp 0 TMA producer, warp 1 UMMA issuer, warps 2-5 hi/lo
// converter (the landed fp32 tile is the tf32 "hi" operand by truncation; lo = rn_tf32(x - trunc_tf32(x)) is written
// beside it), warps 6-9 epilogue (scatter), warps 10-13 finalize.  The weights (hi and lo, 16*COUT rows x 32) are split once per CTA and written in the
// 128-byte-swizzled K-major layout by ordinary stores.  TMEM: 4 accumulator pairs (big = hi*hi, small = lo*hi + hi*lo)
// of 2 x 64 columns.
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

#define L4_THREADS 448                  // warps 0 producer, 1 issuer, 2-5 converter, 6-9 epilogue (scatter), 10-13 finalize
#define L4_STAGES 4
#define L4_NACC 4
#define L4_TILE (128 * 32 * 4)          // one A tile: 128 pixels x 32 channels fp32 = 16 KB
#define L4_BROWS 64                     // rows reserved per weight panel (16*COUT <= 64)
#define L4_PITCH 36                     // floats per row of the smem output image: position = column + 1, 8-byte aligned pairs

int mmvae_tc_tmap_rows(CUtensorMap* m, const float* base, long rows, long cols, int box_rows);   // gemm_tc.cu
int mmvae_tc_dbg_bits();                                                                        // gemm_tc.cu

__device__ __forceinline__ float l4_tf32_lo(float a) {       // as tf32_lo in gemm_tc.cu
    const float r = a - __uint_as_float(__float_as_uint(a) & 0xFFFFE000u);
    return __uint_as_float(__float_as_uint(r) + 0x1000u);
}
__device__ __forceinline__ void l4_bar_epi() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
// sigmoid through ex2.approx / rcp.approx (2 MUFU ops; |error| < 2e-7 absolute, inside the 2e-6 parity bound): the
// IEEE expf + division of the SIMT kernel would cost ~40 issue slots per output pixel in the four epilogue warps
__device__ __forceinline__ float l4_act_fast(float v, int act) {
    if (act == MMVAE_ACT_RELU) return fmaxf(v, 0.f);
    if (act == MMVAE_ACT_SIGMOID) return __fdividef(1.f, 1.f + __expf(-v));
    return v;
}
// TMEM -> registers without the wait (several loads in flight, one tcgen05.wait::ld for all of them)
__device__ __forceinline__ void l4_tmem_ld16_nowait(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void l4_tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Evaluator mode (no gradient): the finished image never reaches HBM.  The finalize warps, which hold every output pixel of an
// image after bias + sigmoid, reduce log p(x | image) on the spot (mmvae.py:154-156, multi_vaes.py:306-311: Normal / Laplace
// with unit scale) and write one (hi, lo) pair per image: 12 KB of stores and the 12 KB re-read of loglik_fwd_kernel per latent
// row are gone (at K = 1000 that is a quarter of the decoder's HBM traffic).  lp == NULL: plain forward.
struct L4LogLik {
    const float* x;          // [B, COUT*32*32] data, row (img / x_div) % B
    float* lp;               // [n_img] or NULL
    float* lp_lo;            // [n_img] residual of the fp64 row sum, or NULL
    int B, x_div, lik;
    float scale;
    double row_const;        // -D * log sqrt(2 pi)  /  -D * log 2
};

template <int COUT>
__global__ void __launch_bounds__(L4_THREADS, 1)
convt_l4_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const float* __restrict__ w,
                       const float* __restrict__ bias, float* __restrict__ y, int n_img, int act, int dbg, const L4LogLik ll) {
    constexpr int NB = 16 * COUT;                 // UMMA N
    constexpr int OUT_FLOATS = COUT * 32 * 32;
    constexpr int OUT_SMEM = COUT * 32 * L4_PITCH;     // floats of the padded smem image
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* a_ring = smem;                                          // L4_STAGES x (hi 16 KB + lo 16 KB)
    uint8_t* b_hi = a_ring + L4_STAGES * 2 * L4_TILE;                // 64 rows x 128 B
    uint8_t* b_lo = b_hi + L4_BROWS * 128;
    float* out0 = (float*)(b_lo + L4_BROWS * 128);                   // 2 x [COUT][32][L4_PITCH]: image n in buffer n & 1
    uint64_t* bars = (uint64_t*)(out0 + 2 * OUT_SMEM);
    uint64_t* full = bars;
    uint64_t* conv = bars + L4_STAGES;
    uint64_t* empty = bars + 2 * L4_STAGES;
    uint64_t* tfull = bars + 3 * L4_STAGES;
    uint64_t* tempty = tfull + L4_NACC;
    uint64_t* img_full = tempty + L4_NACC;        // [2] scattered image complete (epilogue -> converter warps)
    uint64_t* img_empty = img_full + 2;           // [2] image stored and zeroed again (converter -> epilogue warps)
    uint32_t* tmem_slot = (uint32_t*)(img_empty + 2);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler: the single-warp roles run on the uniform datapath

    if (threadIdx.x == 0) {
        for (int s = 0; s < L4_STAGES; ++s) { mbar_init(full + s, 1); mbar_init(conv + s, 4); mbar_init(empty + s, 1); }
        for (int a = 0; a < L4_NACC; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 4); }
        for (int b = 0; b < 2; ++b) { mbar_init(img_full + b, 1); mbar_init(img_empty + b, 4); }
        mbar_fence_init();
        tma_prefetch_desc(&tmap_x);
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    // weights: B[n = (ky*4+kx)*COUT + co][k = ci] = w[ci][co][ky][kx], K-major rows of 128 bytes, 16-byte chunk c of row
    // r stored at chunk c ^ (r & 7) (SWIZZLE_128B); hi = the fp32 value (the tensor core truncates it), lo = residual
    for (int i = threadIdx.x; i < L4_BROWS * 32; i += L4_THREADS) {
        const int r = i >> 5, k = i & 31;
        float v = 0.f;
        if (r < NB) {
            const int tap = r / COUT, co = r - tap * COUT;
            v = __ldg(w + ((long)k * COUT + co) * 16 + tap);
        }
        const int off = r * 128 + (((k >> 2) ^ (r & 7)) << 4) + (k & 3) * 4;
        *(float*)(b_hi + off) = v;
        *(float*)(b_lo + off) = l4_tf32_lo(v);
    }
    for (int i = threadIdx.x; i < 2 * OUT_SMEM; i += L4_THREADS) out0[i] = 0.f;
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const int my_imgs = (n_img - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // images blockIdx.x, +grid, ...
    const int my_tiles = 2 * my_imgs;

    if (warp == 0) {
        // ===== producer: one 128-pixel box per tile =====
        for (int it = 0; it < my_tiles; ++it) {
            const int s = it % L4_STAGES, img = blockIdx.x + (it >> 1) * gridDim.x;
            mbar_wait(empty + s, ((it / L4_STAGES) & 1) ^ 1);
            mbar_arrive_expect_tx_elect(full + s, L4_TILE);
            tma_load_2d_elect(smem_u32(a_ring + s * 2 * L4_TILE), &tmap_x, full + s, 0, img * 256 + (it & 1) * 128);
        }
    } else if (warp == 1) {
        // ===== UMMA issuer =====
        constexpr uint32_t idesc = umma_idesc_tf32(128, NB, 0, 0);
        const uint64_t da0 = umma_smem_desc(smem_u32(a_ring), 16, 1024, 2);
        const uint64_t dbh = umma_smem_desc(smem_u32(b_hi), 16, 1024, 2);
        const uint64_t dbl = umma_smem_desc(smem_u32(b_lo), 16, 1024, 2);
        for (int it = 0; it < my_tiles; ++it) {
            const int s = it % L4_STAGES, acc = it % L4_NACC;
            mbar_wait(tempty + acc, ((it / L4_NACC) & 1) ^ 1);
            mbar_wait(conv + s, (it / L4_STAGES) & 1);
            tc_fence_after_sync();
            const uint64_t dah = da0 + (uint64_t)(s * (2 * L4_TILE >> 4)), dal = dah + (L4_TILE >> 4);
            const uint32_t d_big = tmem_base + acc * 128, d_small = d_big + 64;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                umma_tf32_elect(d_small, dal + kk * 2, dbh + kk * 2, idesc, kk != 0);
                umma_tf32_elect(d_small, dah + kk * 2, dbl + kk * 2, idesc, 1);
                umma_tf32_elect(d_big, dah + kk * 2, dbh + kk * 2, idesc, kk != 0);
            }
            umma_commit_elect(empty + s);
            umma_commit_elect(tfull + acc);
        }
    } else if (warp < 6) {
        // ===== converter: lo tile beside the landed tile =====
        const int tq = threadIdx.x - 64;
        for (int it = 0; it < my_tiles; ++it) {
            const int s = it % L4_STAGES;
            mbar_wait(full + s, (it / L4_STAGES) & 1);
            const uint32_t hi = smem_u32(a_ring + s * 2 * L4_TILE), lo = hi + L4_TILE;
#pragma unroll
            for (int j = 0; j < ((dbg & 4) ? 0 : L4_TILE / 16 / 128); ++j) {
                float4 v;
                const uint32_t off = (uint32_t)(tq + j * 128) * 16u;
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(hi + off));
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(lo + off), "f"(l4_tf32_lo(v.x)),
                             "f"(l4_tf32_lo(v.y)), "f"(l4_tf32_lo(v.z)), "f"(l4_tf32_lo(v.w)) : "memory");
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(conv + s);
        }
    } else if (warp < 10) {
        // ===== epilogue: TMEM lane = input pixel; its 16*COUT values are added into the smem image of its output =====
        const int lane_base = (warp & 3) * 32, e = (warp - 6) * 32 + lane;
        const int p = lane_base + lane;                     // pixel of the tile
        const int ix = p & 15, iyl = p >> 4;                // column, row within the half image
        for (int it = 0; it < my_tiles; ++it) {
            const int acc = it % L4_NACC, half = it & 1, n = it >> 1;
            float* out = out0 + (n & 1) * OUT_SMEM;
            if (!half) mbar_wait(img_empty + (n & 1), ((n >> 1) & 1) ^ 1);   // image n-2 has left this buffer
            mbar_wait(tfull + acc, (it / L4_NACC) & 1);
            tc_fence_after_sync();
            const uint32_t tb = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(acc * 128);
            const int iy = half * 8 + iyl;
            const int oy0 = 2 * iy - 1, ox0 = 2 * ix - 1;
            float val[NB];
            {
                uint32_t rb[NB], rs[NB];
#pragma unroll
                for (int c0 = 0; c0 < NB; c0 += 16) {
                    l4_tmem_ld16_nowait(tb + (uint32_t)c0, rb + c0);
                    l4_tmem_ld16_nowait(tb + 64u + (uint32_t)c0, rs + c0);
                }
                l4_tmem_wait_ld();
#pragma unroll
                for (int j = 0; j < NB; ++j) val[j] = __uint_as_float(rb[j]) + __uint_as_float(rs[j]);
            }
            tc_fence_before_sync();                            // all TMEM reads of this accumulator pair are done
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + acc);
            // Column direction in registers: output columns (2ix-1, 2ix) receive taps kx = 0,1 of this pixel and kx = 2,3
            // of its left neighbour (the previous lane; lanes 0-15 / 16-31 are two image rows), fetched by shuffles.
            // Row direction through the smem image in two phases a = 0,1 (taps ky = 2a, 2a+1 -> output rows
            // 2(iy+a)-1, 2(iy+a)): within a phase every thread owns its own 2x2 block, so plain 64-bit read-modify-writes
            // are safe; the row block below is touched by the thread of the next image row in the other phase, which the
            // barrier orders -> fixed summation order, bit-reproducible results.  Columns are stored shifted by one
            // (position = ox + 1) so that the pair (2ix-1, 2ix) is 8-byte aligned; position 0 collects column -1 (unused).
            const bool has_left = ix != 0;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                if (dbg & 1) break;                            // ablation (tools/l4_time.py): no scatter
                float2 add[2][COUT];
                float last[2][COUT];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int ky = 2 * a + i;
#pragma unroll
                    for (int co = 0; co < COUT; ++co) {
                        const float l2 = __shfl_up_sync(0xffffffffu, val[(ky * 4 + 2) * COUT + co], 1);
                        const float l3 = __shfl_up_sync(0xffffffffu, val[(ky * 4 + 3) * COUT + co], 1);
                        add[i][co].x = val[(ky * 4 + 0) * COUT + co] + (has_left ? l2 : 0.f);
                        add[i][co].y = val[(ky * 4 + 1) * COUT + co] + (has_left ? l3 : 0.f);
                        last[i][co] = val[(ky * 4 + 2) * COUT + co];          // column 31 (ix = 15 only)
                    }
                }
                float2 old[2][COUT];
                float oldl[2][COUT];
                bool ok[2];
                int idx[2];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int oy = oy0 + 2 * a + i;
                    ok[i] = (unsigned)oy < 32u;
                    idx[i] = (ok[i] ? oy : 0) * L4_PITCH + 2 * ix;
#pragma unroll
                    for (int co = 0; co < COUT; ++co) {
                        old[i][co] = *reinterpret_cast<const float2*>(out + co * 32 * L4_PITCH + idx[i]);
                        oldl[i][co] = (ix == 15) ? out[co * 32 * L4_PITCH + idx[i] + 2] : 0.f;
                    }
                }
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    if (ok[i]) {
#pragma unroll
                        for (int co = 0; co < COUT; ++co) {
                            *reinterpret_cast<float2*>(out + co * 32 * L4_PITCH + idx[i]) =
                                make_float2(old[i][co].x + add[i][co].x, old[i][co].y + add[i][co].y);
                            if (ix == 15) out[co * 32 * L4_PITCH + idx[i] + 2] = oldl[i][co] + last[i][co];
                        }
                    }
                }
                l4_bar_epi();
            }
            if (half && e == 0) mbar_arrive(img_full + (n & 1));   // every pixel of image n has been scattered
        }
    } else {
        // ===== finalize: the images the epilogue warps have scattered (bias + activation + NCHW store + re-zero), on their
        // own four warps so that neither the conversion of the next tiles nor the scatter of image n+1 waits for the
        // store of image n (two smem images; img_full / img_empty hand them over) =====
        const int tq = threadIdx.x - 320;
        __shared__ double ll_red[2][4];
        float bv[COUT];
#pragma unroll
        for (int c = 0; c < COUT; ++c) bv[c] = bias ? __ldg(bias + c) : 0.f;
        auto finalize = [&](int n) {
            float* out = out0 + (n & 1) * OUT_SMEM;
            mbar_wait(img_full + (n & 1), (n >> 1) & 1);
            if (!(dbg & 2)) {
                const int img = blockIdx.x + n * gridDim.x;
                // the smem rows are shifted by one column (position = ox + 1): thread (row, j) reads the aligned float4
                // at positions 4j..4j+3 = columns 4j-1..4j+2 and takes column 4j+3 from its right neighbour's first
                // element (the lane above; position 32 for the last float4 of a row) -> aligned 128-bit global stores
                float4* dst = reinterpret_cast<float4*>(y + (long)img * OUT_FLOATS);
                const float4* xr = ll.lp ? reinterpret_cast<const float4*>(ll.x + (long)((img / ll.x_div) % ll.B) * OUT_FLOATS) : nullptr;
                float part = 0.f;                              // this thread's 24 pixel terms (fp32, as the reference's)
                float4 xv[OUT_FLOATS / 4 / 128];
                if (ll.lp) {                                   // all data loads in flight before the first use
#pragma unroll
                    for (int i = 0; i < OUT_FLOATS / 4 / 128; ++i) xv[i] = __ldg(xr + tq + i * 128);
                }
#pragma unroll
                for (int i = 0; i < OUT_FLOATS / 4 / 128; ++i) {
                    const int q = tq + i * 128;                // float4 of the NCHW image: 256 per channel -> channel = i >> 1
                    const int row = q >> 3, j = q & 7;
                    float4* src = reinterpret_cast<float4*>(out + row * L4_PITCH + 4 * j);
                    const float4 m = *src;
                    float nx = __shfl_down_sync(0xffffffffu, m.x, 1);
                    if (j == 7) nx = out[row * L4_PITCH + 32];
                    const float b = bv[i >> 1];
                    float4 o;
                    o.x = l4_act_fast(m.y + b, act); o.y = l4_act_fast(m.z + b, act);
                    o.z = l4_act_fast(m.w + b, act); o.w = l4_act_fast(nx + b, act);
                    if (ll.lp) {
                        const float d0 = xv[i].x - o.x, d1 = xv[i].y - o.y, d2 = xv[i].z - o.z, d3 = xv[i].w - o.w;
                        if (ll.lik == MMVAE_LIK_NORMAL) part += -0.5f * ((d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3));
                        else part += -((fabsf(d0) + fabsf(d1)) + (fabsf(d2) + fabsf(d3)));
                    } else {
                        __stcs(dst + q, o);
                    }
                    *src = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (j == 7) out[row * L4_PITCH + 32] = 0.f;
                }
                if (ll.lp) {                                   // 128 partial sums of 24 terms -> fp64 total (ll_red alternates per image)
                    const float ws = warp_sum(part);
                    if (lane == 0) ll_red[n & 1][warp - 10] = (double)ws;
                    asm volatile("bar.sync 3, 128;" ::: "memory");
                    if (tq == 0) {
                        const double* r = ll_red[n & 1];
                        const double v = (double)ll.scale * (((r[0] + r[1]) + (r[2] + r[3])) + ll.row_const);
                        const float hi = (float)v;
                        ll.lp[img] = hi;
                        if (ll.lp_lo) ll.lp_lo[img] = (float)(v - (double)hi);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(img_empty + (n & 1));
        };
        for (int n = 0; n < my_imgs; ++n) finalize(n);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

template <int COUT>
static int l4_launch(const CUtensorMap& tm, const float* w, const float* bias, float* y, int N, int act, cudaStream_t st,
                     const L4LogLik& ll = L4LogLik{nullptr, nullptr, nullptr, 1, 1, 0, 1.f, 0.0}) {
    static bool attr_done = false;
    constexpr int SMEM = L4_STAGES * 2 * L4_TILE + 2 * L4_BROWS * 128 + 2 * COUT * 32 * L4_PITCH * 4 + 256 + 1024;
    auto kern = convt_l4_fwd_tc_kernel<COUT>;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_l4_fwd_tc: smem attribute: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    const int grid = N < MMVAE_NUM_SMS ? N : MMVAE_NUM_SMS;
    kern<<<grid, L4_THREADS, SMEM, st>>>(tm, w, bias, y, N, act, (mmvae_tc_dbg_bits() >> 10) & 7, ll);   // ablation bits, 0 in production
    MMVAE_LAUNCH_CHECK("convt_l4_fwd_tc");
    return 0;
}

// -> 0 launched, -1 not applicable (the caller uses the SIMT kernel), > 0 error
int mmvae_convt_l4_fwd_tc(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int Cin,
                          int Cout, int act, cudaStream_t st) {
    if (H != 16 || W != 16 || Cin != 32 || (Cout != 3 && Cout != 1)) return -1;
    if (((uintptr_t)x & 15) || ((uintptr_t)y & 15) || (mmvae_tc_dbg_bits() & 512)) return -1;
    CUtensorMap tm;
    const int rc = mmvae_tc_tmap_rows(&tm, x, (long)N * 256, 32, 128);
    if (rc < 0) return -1;                                   // no tcgen05 / TMA on this device
    if (rc) return mmvae_set_error(1, "convt_l4_fwd_tc: cuTensorMapEncodeTiled failed with %d", rc);
    return Cout == 3 ? l4_launch<3>(tm, w, bias, y, N, act, st) : l4_launch<1>(tm, w, bias, y, N, act, st);
}

// forward fused with the pixel log-likelihood (evaluators): lp[img] (+ lp_lo) = scale * sum_pix log p(xdata[(img / x_div) % B] | y_img)
extern "C" int mmvae_convt_small_fwd_loglik_supported(int H, int W, int Cin, int Cout, int lik) {
    return H == 16 && W == 16 && Cin == 32 && (Cout == 3 || Cout == 1) && (lik == MMVAE_LIK_NORMAL || lik == MMVAE_LIK_LAPLACE) &&
           mmvae_has_tcgen05() && !(mmvae_tc_dbg_bits() & 512);
}
extern "C" int mmvae_convt_small_fwd_loglik(const float* x, const float* w, const float* bias, const float* xdata, float* lp,
                                            float* lp_lo, int N, int H, int W, int Cin, int Cout, int act, int B, int x_div, int lik,
                                            float lik_scaling, void* stream) {
    MMVAE_REQUIRE(x && w && xdata && lp && N > 0 && B > 0 && x_div > 0, "convt_small_fwd_loglik: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_small_fwd_loglik_supported(H, W, Cin, Cout, lik), "convt_small_fwd_loglik: unsupported shape / likelihood");
    MMVAE_REQUIRE((((uintptr_t)x | (uintptr_t)xdata) & 15) == 0, "convt_small_fwd_loglik: unaligned pointer");
    CUtensorMap tm;
    const int rc = mmvae_tc_tmap_rows(&tm, x, (long)N * 256, 32, 128);
    if (rc) return mmvae_set_error(1, "convt_small_fwd_loglik: cuTensorMapEncodeTiled failed with %d", rc);
    const int D = Cout * 32 * 32;
    const L4LogLik ll{xdata, lp, lp_lo, B, x_div, lik, lik_scaling,
                      lik == MMVAE_LIK_NORMAL ? -(double)D * (double)MMVAE_LOG_SQRT_2PI : -(double)D * (double)MMVAE_LOG2};
    cudaStream_t st = (cudaStream_t)stream;
    return Cout == 3 ? l4_launch<3>(tm, w, bias, nullptr, N, act, st, ll) : l4_launch<1>(tm, w, bias, nullptr, N, act, st, ll);
}

// =====================================================================================================================
// dgrad of the same layer on tcgen05: gx[pixel, ci] = mask(x) * sum_{(ky,kx,co)} G[pixel,(ky,kx,co)] * w[ci,co,ky,kx]
// with G = the gather (im2col) of the pre-activation gradient, G[pixel,(ky,kx,co)] = g_pre[co, 2iy-1+ky, 2ix-1+kx].
// M = 128 pixels (half an image), N = Cin = 32, K = 16*COUT = 48.  G never exists in memory: per image the builder
// warps turn g (and y for the activation derivative) into a zero-bordered g_pre image in shared memory and gather the
// two 128 x 48 tiles from it straight into the swizzled K-major operand layout, hi and lo (3xTF32) in one pass.
// UMMAs per tile: 6 k-steps x 2 (hi*[hi|lo] as one N = 64 instruction + lo*hi, N = 32).  Epilogue: one pixel per
// thread, ReLU mask of the layer below (the pixel's 128-byte row of x, fetched before the accumulator wait) and a
// full-line store of the pixel's 32 gradients.
// Warp roles (448 threads): 0 bulk-copy producer (g / y images, L2 prefetch of the mask rows), 1 UMMA issuer,
// 2-5 and 10-13 builder (one group per image half), 6-9 epilogue.
// =====================================================================================================================
#define L4D_PITCH 36                      // floats per row of the padded g_pre image (34 used)
#define L4D_ASTAGE (2 * 2 * L4_TILE)      // one A stage: 2 swizzle atoms (k 0-31, 32-63) x (hi + lo) x 16 KB
#define L4D_NACC 4
#define L4D_EPI_PITCH 36                  // floats per staged row: 16-byte aligned, conflict-free for 128-bit accesses

#define L4D_THREADS 448                   // warps 0 producer, 1 issuer, 2-5 + 10-13 builder, 6-9 epilogue
__device__ __forceinline__ void l4_bar_build() { asm volatile("bar.sync 2, 256;" ::: "memory"); }

template <int COUT>
__global__ void __launch_bounds__(L4D_THREADS, 1)
convt_l4_dgrad_tc_kernel(const float* __restrict__ g, const float* __restrict__ y_act, const float* __restrict__ w,
                         const float* __restrict__ x_mask, float* __restrict__ gx, int n_img, int act, int dbg) {
    constexpr int KT = 16 * COUT;                       // 48
    constexpr int NKK = (KT + 7) / 8;                   // k-steps of 8: 6
    constexpr int IMG = COUT * 1024;                    // floats per g / y image
    constexpr int PIMG = COUT * 34 * L4D_PITCH;         // padded g_pre image
    static_assert(KT <= 64 && KT % 4 == 0, "two swizzle atoms");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* a_ring = smem;                                          // 2 stages x 64 KB
    uint8_t* b_pan = a_ring + 2 * L4D_ASTAGE;                        // per atom: [hi 32 rows | lo 32 rows] x 128 B = 8 KB
    float* raw = (float*)(b_pan + 2 * 8192);                         // 2 stages x (g image + y image)
    float* pimg = raw + 2 * 2 * IMG;
    float* epi_pad = pimg + ((PIMG + 3) & ~3);                       // 4 warps x 32 rows x L4D_EPI_PITCH
    uint64_t* bars = (uint64_t*)(epi_pad + 4 * 32 * L4D_EPI_PITCH);
    uint64_t* raw_full = bars;            // [2]
    uint64_t* raw_empty = bars + 2;       // [2] count 4
    uint64_t* a_full = bars + 4;          // [2] count 4
    uint64_t* a_empty = bars + 6;         // [2] count 1 (commit)
    uint64_t* tfull = bars + 8;           // [NACC]
    uint64_t* tempty = bars + 8 + L4D_NACC;
    uint32_t* tmem_slot = (uint32_t*)(tempty + L4D_NACC);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler: the single-warp roles run on the uniform datapath
    const bool has_y = y_act != nullptr && act != MMVAE_ACT_NONE;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) { mbar_init(raw_full + s, 1); mbar_init(raw_empty + s, 8); mbar_init(a_full + s, 4); mbar_init(a_empty + s, 1); }
        for (int a = 0; a < L4D_NACC; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 4); }
        mbar_fence_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
    // weights: B[n = ci][k = (ky*4+kx)*COUT + co] = w[ci][co][ky][kx]; atom a = k / 32; rows 0-31 hi, rows 32-63 lo
    for (int i = threadIdx.x; i < 32 * 64; i += L4D_THREADS) {
        const int n = i >> 6, k = i & 63;
        float v = 0.f;
        if (k < KT) {
            const int tap = k / COUT, co = k - tap * COUT;
            v = __ldg(w + ((long)n * COUT + co) * 16 + tap);
        }
        const int atom = k >> 5, kk = k & 31;
        const int off = atom * 8192 + n * 128 + (((kk >> 2) ^ (n & 7)) << 4) + (kk & 3) * 4;
        *(float*)(b_pan + off) = v;
        *(float*)(b_pan + off + 4096) = l4_tf32_lo(v);
    }
    for (int i = threadIdx.x; i < PIMG; i += L4D_THREADS) pimg[i] = 0.f;     // the border stays zero
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const int my_imgs = (n_img - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    if (warp == 0) {
        // ===== producer: the g and y images of every image of this CTA, two in flight =====
        for (int j = 0; j < my_imgs; ++j) {
            const int s = j & 1;
            const long img = blockIdx.x + (long)j * gridDim.x;
            // the ReLU-mask rows the epilogue will read two images from now: pull them into L2 (32 KB per image), so
            // that its register-staged loads see L2 latency instead of HBM latency
            if (x_mask && lane == 0 && j + 2 < my_imgs && !(dbg & 1)) {
                const float* pf = x_mask + (img + 2L * gridDim.x) * 256 * 32;
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pf), "r"(256 * 32 * 4) : "memory");
            }
            mbar_wait(raw_empty + s, ((j >> 1) & 1) ^ 1);
            if (lane == 0) {
                mbar_arrive_expect_tx(raw_full + s, (has_y ? 2 : 1) * IMG * 4);
                bulk_load_1d(raw + s * 2 * IMG, g + img * IMG, IMG * 4, raw_full + s);
                if (has_y) bulk_load_1d(raw + s * 2 * IMG + IMG, y_act + img * IMG, IMG * 4, raw_full + s);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ===== UMMA issuer =====
        constexpr uint32_t idesc64 = umma_idesc_tf32(128, 64, 0, 0), idesc32 = umma_idesc_tf32(128, 32, 0, 0);
        const uint64_t da0 = umma_smem_desc(smem_u32(a_ring), 16, 1024, 2);
        const uint64_t db0 = umma_smem_desc(smem_u32(b_pan), 16, 1024, 2);
        for (int it = 0; it < 2 * my_imgs; ++it) {
            const int s = it & 1, acc = it % L4D_NACC;
            mbar_wait(tempty + acc, ((it / L4D_NACC) & 1) ^ 1);
            mbar_wait(a_full + s, (it >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t d = tmem_base + acc * 64;                 // [big 32 | small 32]
#pragma unroll
            for (int kk = 0; kk < NKK; ++kk) {
                const int atom = kk >> 2, kin = kk & 3;
                const uint64_t dah = da0 + (uint64_t)((s * L4D_ASTAGE + atom * 2 * L4_TILE) >> 4) + kin * 2;
                const uint64_t dal = dah + (L4_TILE >> 4);
                const uint64_t db = db0 + (uint64_t)((atom * 8192) >> 4) + kin * 2;
                umma_tf32_elect(d, dah, db, idesc64, kk != 0);        // hi*hi -> big, hi*lo -> small
                umma_tf32_elect(d + 32, dal, db, idesc32, 1);         // lo*hi -> small
            }
            umma_commit_elect(a_empty + s);
            umma_commit_elect(tfull + acc);
        }
    } else if (warp < 6 || warp >= 10) {
        // ===== builder (two groups of four warps): all eight build the g_pre image, then group bg gathers the operand
        // tile of image half bg (which always lives in A stage bg) =====
        const int bg = warp >= 10 ? 1 : 0;
        const int t = threadIdx.x - (bg ? 320 : 64);                 // 0..127 = pixel row of the group's tile
        const int ix = t & 15, iyl = t >> 4;
        for (int j = 0; j < my_imgs; ++j) {
            const int rs = j & 1;
            mbar_wait(raw_full + rs, (j >> 1) & 1);
            const float* gi = raw + rs * 2 * IMG;
            const float* yi = gi + IMG;
#pragma unroll
            for (int q = 0; q < ((dbg & 4) ? 0 : IMG / 256); ++q) {      // dbg: ablation bits of tools/l4_time.py
                const int e = t + 128 * bg + q * 256, co = e >> 10, oy = (e >> 5) & 31, ox = e & 31;
                float v = gi[e];
                if (has_y) {
                    const float yy = yi[e];
                    v *= (act == MMVAE_ACT_SIGMOID) ? yy * (1.f - yy) : (yy > 0.f ? 1.f : 0.f);
                }
                pimg[(co * 34 + oy + 1) * L4D_PITCH + ox + 1] = v;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(raw_empty + rs);
            l4_bar_build();                                          // the whole g_pre image is in place
            {
                const int half = bg, it = 2 * j + half, s = bg;
                mbar_wait(a_empty + s, ((it >> 1) & 1) ^ 1);
                const int iy = half * 8 + iyl;
                const float* p0 = pimg + (2 * iy) * L4D_PITCH + 2 * ix;          // P[co][2iy+ky][2ix+kx]
                uint8_t* row = a_ring + s * L4D_ASTAGE + t * 128;
#pragma unroll
                for (int c = 0; c < ((dbg & 2) ? 0 : KT / 4); ++c) {   // 16-byte chunk c = k 4c..4c+3
                    float v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int k = 4 * c + u, tap = k / COUT, co = k - tap * COUT, ky = tap >> 2, kx = tap & 3;
                        v[u] = p0[(co * 34 + ky) * L4D_PITCH + kx];
                    }
                    const int atom = c >> 3, cc = c & 7;
                    // plain stores (no asm barrier): the compiler batches the gather loads of all 12 chunks
                    uint8_t* dst = row + atom * 2 * L4_TILE + ((cc ^ (t & 7)) << 4);
                    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
                    *reinterpret_cast<float4*>(dst + L4_TILE) =
                        make_float4(l4_tf32_lo(v[0]), l4_tf32_lo(v[1]), l4_tf32_lo(v[2]), l4_tf32_lo(v[3]));
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full + s);
            }
            l4_bar_build();                                          // everyone has gathered: the image may be overwritten
        }
    } else {
        // ===== epilogue: TMEM lane = pixel; the warp's 32 rows x 128 B are one contiguous 4 KB block of gx (and of x):
        // rows go through a padded smem tile so that the mask loads and the stores are fully coalesced 128-bit accesses
        // (one 128-byte row per thread cost > 100 us: 32 cache lines per instruction) =====
        const int lane_base = (warp & 3) * 32;
        float* pad = epi_pad + (warp & 3) * 32 * L4D_EPI_PITCH;
        // the mask rows of tile it+1 are requested before tile it is processed (a full tile of work hides the latency)
        auto blk_of = [&](int it) {
            const long img = blockIdx.x + (long)(it >> 1) * gridDim.x;
            return (img * 256 + (it & 1) * 128 + lane_base) * 8;               // float4 index of the warp's 4 KB block
        };
        float4 mk_next[8];
        if (x_mask && !(dbg & 1) && my_imgs > 0) {
#pragma unroll
            for (int q = 0; q < 8; ++q) mk_next[q] = __ldcs(reinterpret_cast<const float4*>(x_mask) + blk_of(0) + lane + 32 * q);
        }
        for (int it = 0; it < 2 * my_imgs; ++it) {
            const int acc = it % L4D_NACC;
            const long blk = blk_of(it);
            float4 mk[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) mk[q] = mk_next[q];
            if (x_mask && !(dbg & 1) && it + 1 < 2 * my_imgs) {
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    mk_next[q] = __ldcs(reinterpret_cast<const float4*>(x_mask) + blk_of(it + 1) + lane + 32 * q);
            }
            mbar_wait(tfull + acc, (it / L4D_NACC) & 1);
            tc_fence_after_sync();
            const uint32_t tb = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(acc * 64);
            uint32_t rb[32], rsm[32];
            l4_tmem_ld16_nowait(tb, rb);
            l4_tmem_ld16_nowait(tb + 16, rb + 16);
            l4_tmem_ld16_nowait(tb + 32, rsm);
            l4_tmem_ld16_nowait(tb + 48, rsm + 16);
            l4_tmem_wait_ld();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + acc);
#pragma unroll
            for (int q = 0; q < 8; ++q)
                *reinterpret_cast<float4*>(pad + lane * L4D_EPI_PITCH + 4 * q) =
                    make_float4(__uint_as_float(rb[4 * q]) + __uint_as_float(rsm[4 * q]),
                                __uint_as_float(rb[4 * q + 1]) + __uint_as_float(rsm[4 * q + 1]),
                                __uint_as_float(rb[4 * q + 2]) + __uint_as_float(rsm[4 * q + 2]),
                                __uint_as_float(rb[4 * q + 3]) + __uint_as_float(rsm[4 * q + 3]));
            __syncwarp();
            float4* dst = reinterpret_cast<float4*>(gx) + blk;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (dbg & 1) continue;
                const int idx = lane + 32 * q;                       // float4 of the block: row idx >> 3, chunk idx & 7
                float4 o = *reinterpret_cast<const float4*>(pad + (idx >> 3) * L4D_EPI_PITCH + 4 * (idx & 7));
                if (x_mask) {
                    o.x = mk[q].x > 0.f ? o.x : 0.f; o.y = mk[q].y > 0.f ? o.y : 0.f;
                    o.z = mk[q].z > 0.f ? o.z : 0.f; o.w = mk[q].w > 0.f ? o.w : 0.f;
                }
                __stcs(dst + idx, o);
            }
            __syncwarp();
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 256);
}

// -> 0 launched, -1 not applicable (the caller uses the SIMT kernel), > 0 error
int mmvae_convt_l4_dgrad_tc(const float* g, const float* y_act, const float* w, const float* x_mask, float* gx, int N,
                            int H, int W, int Cin, int Cout, int act, cudaStream_t st) {
    if (H != 16 || W != 16 || Cin != 32 || Cout != 3) return -1;
    if (((uintptr_t)g & 15) || ((uintptr_t)gx & 15) || ((uintptr_t)y_act & 15) || ((uintptr_t)x_mask & 15)) return -1;
    if (mmvae_tc_dbg_bits() & 512) return -1;
    CUtensorMap tm;
    if (mmvae_tc_tmap_rows(&tm, gx, (long)N * 256, 32, 128) < 0) return -1;      // only asks: tcgen05 device?
    constexpr int COUT = 3;
    constexpr int SMEM = 2 * L4D_ASTAGE + 2 * 8192 + 2 * 2 * COUT * 1024 * 4 + ((COUT * 34 * L4D_PITCH + 3) & ~3) * 4 + 4 * 32 * L4D_EPI_PITCH * 4 + 256 + 1024;
    static bool attr_done = false;
    auto kern = convt_l4_dgrad_tc_kernel<COUT>;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_l4_dgrad_tc: smem attribute: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    const int grid = N < MMVAE_NUM_SMS ? N : MMVAE_NUM_SMS;
    kern<<<grid, L4D_THREADS, SMEM, st>>>(g, y_act, w, x_mask, gx, N, act, (mmvae_tc_dbg_bits() >> 10) & 7);
    MMVAE_LAUNCH_CHECK("convt_l4_dgrad_tc");
    return 0;
}

// =====================================================================================================================
// wgrad of the same layer on tcgen05: gw[ci,co,ky,kx] = sum_pixels x[pixel,ci] * G[pixel,(ky,kx,co)] (+ gb[co] = sum g_pre)
// as D[m = (ky,kx,co)][n = ci] with K = pixels.  Per 32-pixel k-block ("atom"): A = G^T, K-major rows m with the 32
// pixels contiguous, gathered from the zero-bordered g_pre image in shared memory (hi and lo) by ONE builder warp per
// atom (8 warps = the 8 atoms of an image); B = the x rows as an MN-major TMA box (32 pixels x 32 channels), lo written
// beside it by the same warp.  8 UMMAs per atom (hi*[hi|lo] N = 64 + lo*hi N = 32, M = 128 with rows >= 48 unused).
// The TMEM accumulation chain is drained into registers every 4 images (128 UMMAs per accumulator: the tensor core
// truncates its fp32 accumulate, see gemm_tc.cu), double-buffered so the UMMAs never wait for the drain.  Each CTA writes
// one partial in the layout of convt_small_wgrad_reduce (fixed-order, deterministic).
// Warp roles (448 threads): 0 producer (g / y bulk copies, x TMA boxes), 1 UMMA issuer, 2-5 + 10-13 builders, 6-9 drain.
// =====================================================================================================================
#define L4W_THREADS 448
#define L4W_STAGES 8                      // one stage per atom of an image = per builder warp
#define L4W_ASTAGE 12288                  // A_hi rows 0..47 at +0 (6 KB), A_lo at +6144 (UMMA reads 128 rows: the tail is unused)
#define L4W_ALO 6144
#define L4W_BSTAGE 8192                   // B_hi 4 KB | B_lo 4 KB
#define L4W_CHAIN 4                       // images per TMEM accumulation chain

int mmvae_tc_tmap_mn(CUtensorMap* m, const float* base, long krows, long ncols, int box_rows);   // gemm_tc.cu

template <int COUT>
__global__ void __launch_bounds__(L4W_THREADS, 1)
convt_l4_wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmap_x, const float* __restrict__ x_base,
                         const float* __restrict__ g, const float* __restrict__ y_act, float* __restrict__ part, int n_img, int act, int dbg) {
    constexpr int KT = 16 * COUT;                       // 48 rows of D
    constexpr int IMG = COUT * 1024;
    constexpr int PIMG = COUT * 34 * L4D_PITCH;
    constexpr int TOT = 32 * COUT * 16;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* a_ring = smem;
    uint8_t* b_ring = a_ring + L4W_STAGES * L4W_ASTAGE;
    float* raw = (float*)(b_ring + L4W_STAGES * L4W_BSTAGE);
    float* pimg = raw + 2 * 2 * IMG;
    float* bias_part = pimg + ((PIMG + 3) & ~3);                     // [8 warps][COUT]
    uint64_t* bars = (uint64_t*)(bias_part + 32);
    uint64_t* raw_full = bars;                 // [2]
    uint64_t* raw_empty = bars + 2;            // [2] count 8
    uint64_t* b_full = bars + 4;               // [STAGES] TMA
    uint64_t* st_full = bars + 4 + L4W_STAGES; // [STAGES] count 1 (the atom's builder warp)
    uint64_t* st_empty = bars + 4 + 2 * L4W_STAGES;
    uint64_t* tfull = bars + 4 + 3 * L4W_STAGES;        // [2]
    uint64_t* tempty = tfull + 2;                       // [2] count 4
    uint32_t* tmem_slot = (uint32_t*)(tempty + 2);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler: the single-warp roles run on the uniform datapath
    const bool has_y = y_act != nullptr && act != MMVAE_ACT_NONE;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) { mbar_init(raw_full + s, 1); mbar_init(raw_empty + s, 8); mbar_init(tfull + s, 1); mbar_init(tempty + s, 4); }
        for (int s = 0; s < L4W_STAGES; ++s) { mbar_init(b_full + s, 1); mbar_init(st_full + s, 1); mbar_init(st_empty + s, 1); }
        mbar_fence_init();
        tma_prefetch_desc(&tmap_x);
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 128); tmem_relinquish(); }
    for (int i = threadIdx.x; i < PIMG; i += L4W_THREADS) pimg[i] = 0.f;
    // rows 48..127 of the A tiles are read by the M = 128 UMMAs but never used: keep them finite
    for (int i = threadIdx.x; i < (L4W_STAGES * L4W_ASTAGE + L4W_STAGES * L4W_BSTAGE) / 4; i += L4W_THREADS) ((float*)a_ring)[i] = 0.f;
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const int my_imgs = (n_img - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int chains = (my_imgs + L4W_CHAIN - 1) / L4W_CHAIN;

    if (warp == 0) {
        // ===== producer =====
        auto load_raw = [&](int j) {
            const int s = j & 1;
            const long img = blockIdx.x + (long)j * gridDim.x;
            mbar_wait(raw_empty + s, ((j >> 1) & 1) ^ 1);
            if (lane == 0) {
                mbar_arrive_expect_tx(raw_full + s, (has_y ? 2 : 1) * IMG * 4);
                bulk_load_1d(raw + s * 2 * IMG, g + img * IMG, IMG * 4, raw_full + s);
                if (has_y) bulk_load_1d(raw + s * 2 * IMG + IMG, y_act + img * IMG, IMG * 4, raw_full + s);
            }
            __syncwarp();
        };
        // x boxes can only be requested when their stage is free (one use per image), so their latency is exposed unless
        // the rows are already in L2: pull the x block of the image after next (32 KB) into L2 now
        auto prefetch_x = [&](int j) {
            if (lane == 0 && j < my_imgs) {
                const float* pf = x_base + (blockIdx.x + (long)j * gridDim.x) * 256 * 32;
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pf), "r"(256 * 32 * 4) : "memory");
            }
        };
        prefetch_x(0);
        prefetch_x(1);
        if (my_imgs > 0) load_raw(0);
        for (int j = 0; j < my_imgs; ++j) {
            prefetch_x(j + 2);
            if (j + 1 < my_imgs) load_raw(j + 1);
            const long img = blockIdx.x + (long)j * gridDim.x;
#pragma unroll 1
            for (int a = 0; a < 8; ++a) {
                const int s = a, u = j;
                mbar_wait(st_empty + s, (u & 1) ^ 1);
                mbar_arrive_expect_tx_elect(b_full + s, 4096);
                tma_load_2d_elect(smem_u32(b_ring + s * L4W_BSTAGE), &tmap_x, b_full + s, 0, (int)(img * 256 + a * 32));
            }
        }
    } else if (warp == 1) {
        // ===== UMMA issuer =====
        constexpr uint32_t idesc64 = umma_idesc_tf32(128, 64, 0, 1), idesc32 = umma_idesc_tf32(128, 32, 0, 1);
        const uint64_t da0 = umma_smem_desc(smem_u32(a_ring), 16, 1024, 2);
        const uint64_t db0 = umma_smem_desc(smem_u32(b_ring), 4096, 512, 1);      // MN-major, SWIZZLE_128B_BASE32B
        for (int j = 0; j < my_imgs; ++j) {
            const int c = j / L4W_CHAIN, acc = c & 1;
            if (j % L4W_CHAIN == 0) {
                mbar_wait(tempty + acc, ((c >> 1) & 1) ^ 1);
                tc_fence_after_sync();
            }
            const uint32_t d = tmem_base + acc * 64;
#pragma unroll 1
            for (int a = 0; a < 8; ++a) {
                const int s = a, u = j;
                mbar_wait(st_full + s, u & 1);
                tc_fence_after_sync();
                const uint64_t dah = da0 + (uint64_t)((s * L4W_ASTAGE) >> 4), dal = dah + (L4W_ALO >> 4);
                const uint64_t db = db0 + (uint64_t)((s * L4W_BSTAGE) >> 4);
#pragma unroll
                for (int kk = 0; kk < ((dbg & 4) ? 1 : 4); ++kk) {
                    const uint32_t first = !(j % L4W_CHAIN == 0 && a == 0 && kk == 0);
                    umma_tf32_elect(d, dah + kk * 2, db + kk * (1024 >> 4), idesc64, first);       // hi*hi | hi*lo
                    umma_tf32_elect(d + 32, dal + kk * 2, db + kk * (1024 >> 4), idesc32, 1);      // lo*hi
                }
                umma_commit_elect(st_empty + s);
            }
            if (j % L4W_CHAIN == L4W_CHAIN - 1 || j == my_imgs - 1) umma_commit_elect(tfull + acc);
        }
    } else if (warp < 6 || warp >= 10) {
        // ===== builders: warp w owns atom w (pixels 32w..32w+31) of every image =====
        const int w = warp < 6 ? warp - 2 : warp - 6;                // 0..7
        const int tb = w * 32 + lane;                                // 0..255
        const int c = lane & 7, ml = lane >> 3;                      // 16-byte pixel chunk of the atom, row within a group of 4
        const int iy = 2 * w + (c >> 2), ix0 = (c & 3) * 4;
        float bs[COUT];
#pragma unroll
        for (int co = 0; co < COUT; ++co) bs[co] = 0.f;
        for (int j = 0; j < my_imgs; ++j) {
            const int rs = j & 1;
            mbar_wait(raw_full + rs, (j >> 1) & 1);
            const float* gi = raw + rs * 2 * IMG;
            const float* yi = gi + IMG;
#pragma unroll
            for (int q = 0; q < ((dbg & 2) ? 0 : IMG / 256); ++q) {          // dbg: ablation bits of tools/l4_time.py
                const int e = tb + q * 256, co = e >> 10, oy = (e >> 5) & 31, ox = e & 31;
                float v = gi[e];
                if (has_y) {
                    const float yy = yi[e];
                    v *= (act == MMVAE_ACT_SIGMOID) ? yy * (1.f - yy) : (yy > 0.f ? 1.f : 0.f);
                }
                pimg[(co * 34 + oy + 1) * L4D_PITCH + ox + 1] = v;
                bs[q >> 2] += v;                                     // 1024 elements per channel = 4 values of q
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(raw_empty + rs);
            l4_bar_build();
            const int s = w, u = j;
            mbar_wait(st_empty + s, (u & 1) ^ 1);
            uint8_t* ah = a_ring + s * L4W_ASTAGE;
            const float* p0 = pimg + (2 * iy) * L4D_PITCH + 2 * ix0;
#pragma unroll
            for (int jj = 0; jj < ((dbg & 1) ? 0 : KT / 4); ++jj) {
                const int m = 4 * jj + ml, tap = m / COUT, co = m - tap * COUT, ky = tap >> 2, kx = tap & 3;
                const float* pp = p0 + (co * 34 + ky) * L4D_PITCH + kx;
                const float4 v = make_float4(pp[0], pp[2], pp[4], pp[6]);
                uint8_t* dst = ah + m * 128 + ((c ^ (m & 7)) << 4);
                *reinterpret_cast<float4*>(dst) = v;
                *reinterpret_cast<float4*>(dst + L4W_ALO) = make_float4(l4_tf32_lo(v.x), l4_tf32_lo(v.y), l4_tf32_lo(v.z), l4_tf32_lo(v.w));
            }
            mbar_wait(b_full + s, u & 1);
            float4* bh = reinterpret_cast<float4*>(b_ring + s * L4W_BSTAGE);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float4 v = bh[lane + 32 * q];
                bh[256 + lane + 32 * q] = make_float4(l4_tf32_lo(v.x), l4_tf32_lo(v.y), l4_tf32_lo(v.z), l4_tf32_lo(v.w));
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(st_full + s);
            l4_bar_build();
        }
#pragma unroll
        for (int co = 0; co < COUT; ++co) {
            const float t = warp_sum(bs[co]);
            if (lane == 0) bias_part[w * COUT + co] = t;
        }
        l4_bar_build();
        if (w == 0 && lane < COUT) {
            float t = 0.f;
            for (int k = 0; k < 8; ++k) t += bias_part[k * COUT + lane];
            part[(long)blockIdx.x * (TOT + COUT) + TOT + lane] = t;
        }
    } else {
        // ===== drain: TMEM lane = row m of D; registers accumulate the chains =====
        const int lane_base = (warp & 3) * 32, m = lane_base + lane;
        float accv[32];
#pragma unroll
        for (int n = 0; n < 32; ++n) accv[n] = 0.f;
        for (int c = 0; c < chains; ++c) {
            const int acc = c & 1;
            mbar_wait(tfull + acc, (c >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t tbm = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(acc * 64);
            uint32_t rb[32], rsm[32];
            l4_tmem_ld16_nowait(tbm, rb);
            l4_tmem_ld16_nowait(tbm + 16, rb + 16);
            l4_tmem_ld16_nowait(tbm + 32, rsm);
            l4_tmem_ld16_nowait(tbm + 48, rsm + 16);
            l4_tmem_wait_ld();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty + acc);
#pragma unroll
            for (int n = 0; n < 32; ++n) accv[n] += __uint_as_float(rb[n]) + __uint_as_float(rsm[n]);
        }
        if (m < KT) {
            const int tap = m / COUT, co = m - tap * COUT;
            float* dst = part + (long)blockIdx.x * (TOT + COUT);
#pragma unroll
            for (int n = 0; n < 32; ++n) dst[(n * COUT + co) * 16 + tap] = accv[n];
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 128);
}

// partials for convt_small_wgrad_reduce.  -> 0 launched, -1 not applicable, > 0 error
int mmvae_convt_l4_wgrad_tc(const float* x, const float* g, const float* y_act, float* part, int P, int N, int H, int W,
                            int Cin, int Cout, int act, cudaStream_t st) {
    if (H != 16 || W != 16 || Cin != 32 || Cout != 3) return -1;
    if (((uintptr_t)x & 15) || ((uintptr_t)g & 15) || ((uintptr_t)y_act & 15) || (mmvae_tc_dbg_bits() & 512)) return -1;
    if (P != (N < MMVAE_NUM_SMS ? N : MMVAE_NUM_SMS)) return -1;
    CUtensorMap tm;
    const int rc = mmvae_tc_tmap_mn(&tm, x, (long)N * 256, 32, 32);
    if (rc < 0) return -1;
    if (rc) return mmvae_set_error(1, "convt_l4_wgrad_tc: cuTensorMapEncodeTiled failed with %d", rc);
    constexpr int COUT = 3;
    constexpr int SMEM = L4W_STAGES * (L4W_ASTAGE + L4W_BSTAGE) + 2 * 2 * COUT * 1024 * 4 +
                         ((COUT * 34 * L4D_PITCH + 3) & ~3) * 4 + 128 + 512 + 1024;   // the A over-read of the last stage lands in the B ring
    static bool attr_done = false;
    auto kern = convt_l4_wgrad_tc_kernel<COUT>;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_l4_wgrad_tc: smem attribute: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    kern<<<P, L4W_THREADS, SMEM, st>>>(tm, x, g, y_act, part, N, act, (mmvae_tc_dbg_bits() >> 10) & 7);
    MMVAE_LAUNCH_CHECK("convt_l4_wgrad_tc");
    return 0;
}
