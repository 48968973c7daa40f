// Kernel (4), tensor-core path: C[M,N] = act(A*B + bias) in fp32 accuracy on the 5th-gen tensor cores.
//   * persistent: one CTA per SM walks a static tile list; operands stream HBM -> shared memory by TMA
//     (cp.async.bulk.tensor.2d, 128-byte swizzle, OOB zero-fill for tails) through a ring that runs across tiles;
//   * tcgen05.mma kind::tf32, UMMA 128 x BN x 8, fp32 accumulators in TMEM, double-buffered so the epilogue of tile i
//     overlaps the MMAs of tile i+1;
//   * fp32 accuracy by the error-compensated 3xTF32 split done IN the kernel: a converter warpgroup writes
//     lo = rn_tf32(a - trunc_tf32(a)) next to each landed tile (the tensor core reads the top 19 bits of the raw fp32
//     word as the hi part — measured: it truncates), and the issuer runs lo*hi + hi*lo into a "small" accumulator and
//     hi*hi into a "big" one (the TMEM accumulate truncates too — measured ~2^-25 relative bias per MMA — so the big
//     chain is kept short: <= TC_MAX_CHAIN k-blocks per split, the two are added in fp32 in the epilogue);
//   * K-major operands use SWIZZLE_128B, MN-major operands SWIZZLE_128B_BASE32B (the only MN-major tf32 layout the
//     hardware has) with the instruction-descriptor transpose bits, so dgrad/wgrad need no transposed copies;
//   * deterministic split-K (partials + splitk_reduce) for the huge-K weight gradients; optional transposed store.
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2-5 = hi/lo converter, warps 6-9 = epilogue.
// Reference ops replaced: the same torch.nn.Linear / Conv2d / ConvTranspose2d contractions as gemm_simt.cu.
#include <cuda.h>
#include <cudaTypedefs.h>
#include "common.cuh"
#include "sm100_ptx.cuh"

#define TC_BM 128
#define TC_BK 32                       // fp32 elements per k-block = one 128-byte swizzle row
#define TC_THREADS 320
#define TC_A_TILE (TC_BM * TC_BK * 4)  // 16 KB
#define TC_EPI_PITCH 36               // floats per staged row: 16-byte aligned, conflict-free for 128-bit accesses
#define TC_EPI_BYTES (4 * 32 * TC_EPI_PITCH * 4)     // one 32 x 32 transposition pad per epilogue warp
#define TC_MAX_CHAIN 32                // k-blocks accumulated in one TMEM chain when split-K is available

static int g_tc_dbg = 0;
static long g_tc_launches = 0;         // launches of gemm_tc_kernel (tests assert that a shape really ran on tcgen05)
static int g_tc_mode = 0;              // 0 = 3xTF32 (fp32 accuracy), 1 = single TF32 pass (debug / speed probe)

template <int BN> struct TcCfg {
    static constexpr int B_TILE = BN * TC_BK * 4;
    static constexpr int STAGE = 2 * TC_A_TILE + 2 * B_TILE;
    static constexpr int STAGES = (BN == 128) ? 3 : 4;
    static constexpr int SMEM = STAGES * STAGE + 1024 + 256 + TC_EPI_BYTES;
    static constexpr int TMEM_COLS = 4 * BN;          // {big, small} x 2 accumulator buffers
};

// lo part of the 3xTF32 split: residual of the hardware's truncation of `a`, rounded to nearest at tf32 precision by
// adding half a tf32 ulp to the bit pattern (the tensor core truncates the low 13 bits itself), so that the split error is
// unbiased.  Three full-rate instructions per element; cvt.rna.tf32.f32 compiles to six (inf/nan guards) and the
// converter warps share issue slots with the MMA issuer.
__device__ __forceinline__ float tf32_lo(float a) {
    const float r = a - __uint_as_float(__float_as_uint(a) & 0xFFFFE000u);
    return __uint_as_float(__float_as_uint(r) + 0x1000u);
}
// lo tile <- split(hi tile), n4 float4 per thread of a 128-thread group, through shared-space (not generic) accesses
__device__ __forceinline__ void tf32_split_tile(uint32_t hi_addr, uint32_t lo_addr, int tq, int n4) {
#pragma unroll 4
    for (int j = 0; j < n4; ++j) {
        float4 v;
        const uint32_t off = (uint32_t)(tq + j * 128) * 16u;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(hi_addr + off));
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(lo_addr + off), "f"(tf32_lo(v.x)), "f"(tf32_lo(v.y)),
                     "f"(tf32_lo(v.z)), "f"(tf32_lo(v.w)) : "memory");
    }
}

// bf16x3 mode (passes == 2; K-major x K-major operands): a landed row of 32 fp32 (128 bytes, 16-byte chunks XOR-swizzled by the
// row) is split IN PLACE into 16 words of packed bf16 hi pairs | 16 words of packed bf16 mid pairs -- the same 128 bytes are then
// a K-major bf16 operand row whose hi half starts at byte 0 and whose mid half at byte 64 (convt_ts.cu uses the same split).
__device__ __forceinline__ void bf16_split_row(uint32_t row_addr, int sw) {
    float f[32];
#pragma unroll
    for (int c = 0; c < 8; ++c)
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[4 * c]), "=f"(f[4 * c + 1]), "=f"(f[4 * c + 2]), "=f"(f[4 * c + 3])
                     : "r"(row_addr + ((uint32_t)(c ^ sw) << 4)));
    uint32_t hp[16], mp[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hp[j]) : "f"(f[2 * j + 1]), "f"(f[2 * j]));
        const float r0 = f[2 * j] - __uint_as_float(hp[j] << 16), r1 = f[2 * j + 1] - __uint_as_float(hp[j] & 0xffff0000u);
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(mp[j]) : "f"(r1), "f"(r0));
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + ((uint32_t)(c ^ sw) << 4)), "r"(hp[4 * c]), "r"(hp[4 * c + 1]),
                     "r"(hp[4 * c + 2]), "r"(hp[4 * c + 3]) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + ((uint32_t)((4 + c) ^ sw) << 4)), "r"(mp[4 * c]), "r"(mp[4 * c + 1]),
                     "r"(mp[4 * c + 2]), "r"(mp[4 * c + 3]) : "memory");
    }
}

__device__ __forceinline__ float apply_act_tc(float v, int act) {
    if (act == MMVAE_ACT_RELU) return fmaxf(v, 0.f);
    if (act == MMVAE_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
    return v;
}

// Implicit-GEMM transposed convolution (k4 s2 p1): the same pipeline with the A operand fetched by 4-D TMA boxes over the
// NHWC activation / gradient tensor instead of from an im2col buffer (out-of-bound rows and the padding ring arrive as
// zeros from the TMA unit; stride-2 gathers use the tensor map's elementStrides):
//   mode 1 forward : per output-parity class (py,px) a GEMM rows x Cout x 4*Cin; the class is the N-tile index, the A tile
//                    of k-block (tap,chunk) is the input window shifted by (dy,dx) in {-1,0,1}^2; the epilogue scatters
//                    rows to y[n, 2i+py, 2j+px, :] with bias + activation;
//   mode 2 dgrad   : rows x Cin x 16*Cout; the A tile of k-block (tap,chunk) is g[n, 2i-1+ky, 2j-1+kx, chunk] (stride-2 box);
//                    the epilogue applies the ReLU mask of the layer below;
//   mode 3 wgrad   : (16*Cout) x Cin x rows, both operands MN-major, A = stride-2 boxes of g, B = x; split-K over rows,
//                    transposed store into gWp[Cin, (tap,co)].
struct TcConv {
    int mode;          // 0 = plain GEMM
    int H, W;          // input (coarse) grid
    int chunks;        // 32-channel chunks per tap (mode 1: Cin/32; modes 2: Cout/32)
    int Cout;          // mode 3: channels per tap of the M index
};

struct TcTile { int m0, n0, kb_begin, nk, z, nt; };

__device__ __forceinline__ TcTile tc_tile(int t, int tiles_m, int tiles_n, int BN, int nkb, int kb_per) {
    TcTile r;
    const int nt = t % tiles_n;
    const int mt = (t / tiles_n) % tiles_m;
    r.z = t / (tiles_n * tiles_m);
    r.nt = nt;
    r.m0 = mt * TC_BM;
    r.n0 = nt * BN;
    r.kb_begin = r.z * kb_per;
    r.nk = min(nkb, r.kb_begin + kb_per) - r.kb_begin;
    return r;
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const float* __restrict__ bias, float* __restrict__ C, float* __restrict__ ws, int M, int N, int K,
               int act, int kb_per_split, int splits, int c_transposed, int passes_dbg, const TcConv cv,
               const float* __restrict__ mask) {
    using Cfg = TcCfg<BN>;
    const int passes = passes_dbg & 0xff, dbg = passes_dbg >> 8;      // dbg: ablation bits of tools/f9_ablate.py, 0 in production
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* bars = (uint64_t*)(smem + STAGES * Cfg::STAGE);
    uint64_t* full = bars;                      // [STAGES] TMA landed
    uint64_t* conv = bars + STAGES;             // [STAGES] lo tiles written
    uint64_t* empty = bars + 2 * STAGES;        // [STAGES] MMAs that read the stage completed
    uint64_t* tfull = bars + 3 * STAGES;        // [2] accumulator buffer complete
    uint64_t* tempty = bars + 3 * STAGES + 2;   // [2] accumulator buffer drained by the epilogue
    uint32_t* tmem_slot = (uint32_t*)(bars + 3 * STAGES + 4);
    float* epi = (float*)(smem + STAGES * Cfg::STAGE + 256);          // epilogue transposition pads (16-byte aligned)

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler: the single-warp roles run on the uniform datapath
    const int nkb = (K + TC_BK - 1) / TC_BK;
    const int tiles_m = (M + TC_BM - 1) / TC_BM, tiles_n = (N + BN - 1) / BN;
    const int total = tiles_m * tiles_n * splits;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full + s, 1); mbar_init(conv + s, 4); mbar_init(empty + s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 4); }
        mbar_fence_init();
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
    }
    if (warp == 1) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (whole warp converged, one elected lane issues; see sm100_ptx.cuh): the stage ring runs
        // straight through tile boundaries =====
        const uint32_t smem_base = smem_u32(smem);
        int s = 0, ph = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
            const TcTile tl = tc_tile(t, tiles_m, tiles_n, BN, nkb, kb_per_split);
            for (int i = 0; i < tl.nk; ++i) {
                mbar_wait(empty + s, ph ^ 1);
                const uint32_t a_hi = smem_base + s * Cfg::STAGE, b_hi = a_hi + 2 * TC_A_TILE;
                if (dbg & 6) {                                  // ablation: no operand traffic
                    mbar_arrive_elect(full + s);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                    continue;
                }
                mbar_arrive_expect_tx_elect(full + s, TC_A_TILE + Cfg::B_TILE);
                const int kbi = tl.kb_begin + i;
                const int k0 = kbi * TC_BK;
                if (cv.mode == 0) {
                    if (A_MN) {
#pragma unroll
                        for (int j = 0; j < TC_BM / 32; ++j)
                            tma_load_2d_elect(a_hi + j * 4096, &tmap_a, full + s, tl.m0 + 32 * j, k0);
                    } else {
                        tma_load_2d_elect(a_hi, &tmap_a, full + s, k0, tl.m0);
                    }
                } else if (cv.mode == 3) {
                    // k-block = 32 consecutive pixels (n, y, x) of the coarse grid; M group j = 32 channels of one tap
                    const int HW = cv.H * cv.W, p0 = kbi * 32;
                    const int n0 = p0 / HW, y0 = (p0 - n0 * HW) / cv.W;
#pragma unroll
                    for (int j = 0; j < TC_BM / 32; ++j) {
                        const int m = tl.m0 + 32 * j, tap = m / cv.Cout, c0 = m - tap * cv.Cout;
                        tma_load_4d_elect(a_hi + j * 4096, &tmap_a, full + s, c0, (tap & 3) - 1, 2 * y0 + (tap >> 2) - 1, n0);
                    }
                } else {
                    const int HW = cv.H * cv.W;
                    const int n0 = tl.m0 / HW, y0 = (tl.m0 - n0 * HW) / cv.W;
                    const int tap = kbi / cv.chunks, c0 = (kbi - tap * cv.chunks) * 32;
                    if (cv.mode == 1) {                      // class (py,px) = N tile; tap (ty,tx) -> window shift
                        const int py = tl.nt >> 1, px = tl.nt & 1;
                        const int dy = (tap >> 1) ? 2 * py - 1 : 0, dx = (tap & 1) ? 2 * px - 1 : 0;
                        tma_load_4d_elect(a_hi, &tmap_a, full + s, c0, dx, y0 + dy, n0);
                    } else {                                 // mode 2: tap (ky,kx) of the stride-2 gather
                        tma_load_4d_elect(a_hi, &tmap_a, full + s, c0, (tap & 3) - 1, 2 * y0 + (tap >> 2) - 1, n0);
                    }
                }
                if (B_MN) {
#pragma unroll
                    for (int j = 0; j < BN / 32; ++j)
                        tma_load_2d_elect(b_hi + j * 4096, &tmap_b, full + s, tl.n0 + 32 * j, k0);
                } else {
                    tma_load_2d_elect(b_hi, &tmap_b, full + s, k0, tl.n0);
                }
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (whole warp converged, one elected lane issues) =====
        constexpr uint32_t idesc = umma_idesc_tf32(TC_BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
        constexpr uint32_t idesc2 = umma_idesc_tf32(TC_BM, 2 * BN <= 256 ? 2 * BN : 256, A_MN ? 1 : 0, B_MN ? 1 : 0);
        // K-major (SWIZZLE_128B): 8 tf32 = 32 bytes inside the 128-byte row, 8-row groups 1024 B apart.
        // MN-major (SWIZZLE_128B_BASE32B): 8 k-rows of 128 B = 1024 B per MMA, 4-row swizzle atoms
        // 512 B apart (SBO), 32-element MN groups one TMA box = 4096 B apart (LBO).
        constexpr uint32_t a_lbo = A_MN ? 4096 : 16, b_lbo = B_MN ? 4096 : 16;
        constexpr uint32_t a_sbo = A_MN ? 512 : 1024, b_sbo = B_MN ? 512 : 1024;
        constexpr uint32_t a_lt = A_MN ? 1 : 2, b_lt = B_MN ? 1 : 2;
        constexpr uint32_t a_step = (A_MN ? 1024 : 32) >> 4, b_step = (B_MN ? 1024 : 32) >> 4;   // descriptor units
        const uint32_t smem_base = smem_u32(smem);
        // descriptors of stage 0; a stage / operand / k-slice offset is an add on the 14-bit address field
        const uint64_t da0 = umma_smem_desc(smem_base, a_lbo, a_sbo, a_lt);
        const uint64_t db0 = umma_smem_desc(smem_base + 2 * TC_A_TILE, b_lbo, b_sbo, b_lt);
        int s = 0, ph = 0, ti = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++ti) {
            const TcTile tl = tc_tile(t, tiles_m, tiles_n, BN, nkb, kb_per_split);
            const int acc = ti & 1;
            mbar_wait(tempty + acc, ((ti >> 1) & 1) ^ 1);          // epilogue has drained this buffer
            tc_fence_after_sync();
            const uint32_t d_big = tmem_base + acc * (2 * BN), d_small = d_big + BN;
            for (int i = 0; i < tl.nk; ++i) {
                mbar_wait(passes >= 2 ? conv + s : full + s, ph);
                tc_fence_after_sync();
                const uint64_t dah = da0 + (uint64_t)(s * (Cfg::STAGE >> 4)), dal = dah + (TC_A_TILE >> 4);
                const uint64_t dbh = db0 + (uint64_t)(s * (Cfg::STAGE >> 4)), dbl = dbh + (Cfg::B_TILE >> 4);
                if (!A_MN && !B_MN && passes == 2) {
                    // bf16x3: hi*hi + hi*mid + mid*hi, K = 16 per MMA (two k-steps per 32-float k-block); hi / mid halves of the
                    // split rows are 64 bytes apart, a k-step 32 bytes
                    constexpr uint32_t idesc_bf = umma_idesc_bf16(TC_BM, BN, 0, 0);
#pragma unroll
                    for (int kk = 0; kk < 2; ++kk) {
                        const uint32_t first = (i | kk) != 0;
                        umma_bf16_ss_elect(d_big, dah + kk * 2, dbh + kk * 2, idesc_bf, first);
                        umma_bf16_ss_elect(d_big, dah + kk * 2, dbh + 4 + kk * 2, idesc_bf, 1);
                        umma_bf16_ss_elect(d_big, dah + 4 + kk * 2, dbh + kk * 2, idesc_bf, 1);
                    }
                    umma_commit_elect(empty + s);
                    if (++s == STAGES) { s = 0; ph ^= 1; }
                    continue;
                }
#pragma unroll
                for (int kk = 0; kk < TC_BK / 8; ++kk) {
                    if ((dbg & 16) && (i | kk) != 0) break;
                    const uint32_t first = (i | kk) != 0;
                    if (passes == 3 && BN <= 64 && !(dbg & 8192)) {            // bit 8192: A/B switch of tools/gemm_ablate.py
                        // The lo tile of B follows its hi tile at the row-group (K-major) / column-group (MN-major)
                        // pitch, and `small` follows `big` in TMEM: ONE instruction of N = 2*BN computes
                        // hi*hi -> big and hi*lo -> small.  A UMMA costs 32 + N/4 clocks (the 32 = reading the A tile),
                        // so two instructions instead of three save 22 % (BN = 64) / 27 % (BN = 32) of the issue stream.
                        umma_tf32_elect(d_big, dah + kk * a_step, dbh + kk * b_step, idesc2, first);
                        umma_tf32_elect(d_small, dal + kk * a_step, dbh + kk * b_step, idesc, 1);
                        continue;
                    }
                    if (passes == 3) {
                        umma_tf32_elect(d_small, dal + kk * a_step, dbh + kk * b_step, idesc, first);
                        umma_tf32_elect(d_small, dah + kk * a_step, dbl + kk * b_step, idesc, 1);
                    }
                    umma_tf32_elect(d_big, dah + kk * a_step, dbh + kk * b_step, idesc, first);
                }
                umma_commit_elect(empty + s);                  // frees the stage once these MMAs have read it
                if (++s == STAGES) { s = 0; ph ^= 1; }
            }
            umma_commit_elect(tfull + acc);
        }
    } else if (warp < 6) {
        // ===== converter (warps 2..5): lo = rn_tf32(a - trunc_tf32(a)), same swizzled position in the lo tile =====
        const int tq = threadIdx.x - 64;               // 0..127
        if (!A_MN && !B_MN && passes == 2) {
            int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const TcTile tl = tc_tile(t, tiles_m, tiles_n, BN, nkb, kb_per_split);
                for (int i = 0; i < tl.nk; ++i, ++it) {
                    const int s = it % STAGES, ph = (it / STAGES) & 1;
                    mbar_wait(full + s, ph);
                    const uint32_t st = smem_u32(smem) + s * Cfg::STAGE;
                    bf16_split_row(st + tq * 128, tq & 7);                                  // A: one row per thread
                    if (tq < BN) bf16_split_row(st + 2 * TC_A_TILE + tq * 128, tq & 7);      // B: BN rows
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(conv + s);
                }
            }
        } else if (passes == 3) {
            int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const TcTile tl = tc_tile(t, tiles_m, tiles_n, BN, nkb, kb_per_split);
                for (int i = 0; i < tl.nk; ++i, ++it) {
                    const int s = it % STAGES, ph = (it / STAGES) & 1;
                    mbar_wait(full + s, ph);
                    if (dbg & 8) { __syncwarp(); if (lane == 0) mbar_arrive(conv + s); continue; }
                    const uint32_t st = smem_u32(smem) + s * Cfg::STAGE;
                    tf32_split_tile(st, st + TC_A_TILE, tq, TC_A_TILE / 16 / 128);
                    tf32_split_tile(st + 2 * TC_A_TILE, st + 2 * TC_A_TILE + Cfg::B_TILE, tq, Cfg::B_TILE / 16 / 128);
                    fence_proxy_async_smem();          // generic-proxy writes -> visible to the tensor core's reads
                    __syncwarp();
                    if (lane == 0) mbar_arrive(conv + s);
                }
            }
        }
    } else {
        // ===== epilogue (warps 6..9): TMEM -> registers -> global, overlapped with the next tile's MMAs =====
        const int lane_base = (warp & 3) * 32;         // a warp may only touch TMEM lanes 32*(warp%4) .. +31
        constexpr int NCH = BN / 32;
        const int col = (lane & 7) * 4, rsub = lane >> 3;       // this lane's place in the transposed (row-coalesced) block
        // fused ReLU backward of the layer below (mode 2): the mask values (the layer's input, streamed from HBM) are
        // fetched one 32-column chunk ahead of their use -- loaded at the point of use their latency made dgrad
        // epilogue-bound (L3 dgrad 424 us with, 263 us without the mask)
        const bool use_mask = mask && splits == 1 && (N & 3) == 0 && !(dbg & 32);
        float4 mk[2][8];
        auto load_mask = [&](const TcTile& tl, int c, float4* dst) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int grow = tl.m0 + lane_base + 4 * k + rsub, n = tl.n0 + c + col;
                dst[k] = (grow < M && n + 4 <= N) ? __ldg(reinterpret_cast<const float4*>(mask + (long)grow * N + n))
                                                  : make_float4(1.f, 1.f, 1.f, 1.f);
            }
        };
        int ti = 0;
        if (use_mask && blockIdx.x < total) load_mask(tc_tile(blockIdx.x, tiles_m, tiles_n, BN, nkb, kb_per_split), 0, mk[0]);
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++ti) {
            const TcTile tl = tc_tile(t, tiles_m, tiles_n, BN, nkb, kb_per_split);
            const int acc = ti & 1;
            mbar_wait(tfull + acc, (ti >> 1) & 1);
            tc_fence_after_sync();
            const int row = tl.m0 + lane_base + lane;
            float* out = (splits > 1) ? ws + (long)tl.z * M * N : C;
            const uint32_t tb = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(acc * 2 * BN);
            // The chunk loop is unrolled by two only (the mask buffers alternate with the chunk parity): fully unrolled the
            // epilogue of a 128-column tile is ~8K instructions that each run once per tile, and the epilogue warps stall
            // on instruction fetch (ncu: no_instruction 8.3 per issue, 12 us per tile with one k-block per tile).
            constexpr int CSTEP = (NCH >= 2) ? 2 : 1;
#pragma unroll 1
            for (int cp = 0; cp < NCH; cp += CSTEP)
#pragma unroll
            for (int hh = 0; hh < CSTEP; ++hh) {
                const int ci = cp + hh;
                const int c = ci * 32;
                // mask buffer parity: chunks alternate; with one chunk per tile (BN = 32) the buffer is reloaded in place
                constexpr int kOne = (NCH == 1);
                const int cur = kOne ? 0 : hh, nxt = kOne ? 0 : (hh ^ 1);
                if (use_mask && !kOne) {
                    if (ci + 1 < NCH) load_mask(tl, c + 32, mk[nxt]);
                    else if (t + (int)gridDim.x < total)
                        load_mask(tc_tile(t + gridDim.x, tiles_m, tiles_n, BN, nkb, kb_per_split), 0, mk[nxt]);
                }
                float v[32];
                tmem_ld_32x32(tb + (uint32_t)c, v);
                if (passes == 3) {
                    float u[32];
                    tmem_ld_32x32(tb + (uint32_t)(BN + c), u);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += u[j];
                }
                if (ci + 1 == NCH) {                    // last TMEM read of this buffer: hand it back to the issuer
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(tempty + acc);
                }
                if (dbg & 1) continue;
                if (c_transposed) {
                    if (row < M) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int n = tl.n0 + c + j;
                            if (n < N) out[(long)n * M + row] = v[j];       // lanes = consecutive rows: coalesced
                        }
                    }
                    continue;
                }
                // bias + activation by column (the thread holds one row), then transpose the warp's 32 x 32 block through
                // shared memory so that global traffic is row-coalesced: each 128-bit access instruction then covers
                // 4 rows x 128 contiguous bytes (whole 32-byte sectors) instead of 32 rows x 16 bytes
                if (splits == 1 && (bias || act)) {
                    // one coalesced bias load per 32-column chunk, broadcast by shuffles: 32 predicated scalar loads
                    // per thread serialised at L2 latency (17 us per 128-column tile, tools/gemm_fixed_cost.py)
                    const int nb = (cv.mode == 1) ? c : tl.n0 + c;
                    const float bl = (bias && nb + lane < N) ? __ldg(bias + nb + lane) : 0.f;
                    if (act == MMVAE_ACT_RELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + __shfl_sync(0xffffffffu, bl, j), 0.f);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = apply_act_tc(v[j] + __shfl_sync(0xffffffffu, bl, j), act);
                    }
                }
                float* pad = epi + (warp & 3) * 32 * TC_EPI_PITCH;
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<float4*>(pad + lane * TC_EPI_PITCH + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                __syncwarp();
                if (use_mask && kOne) load_mask(tl, c, mk[0]);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int r = 4 * k + rsub, grow = tl.m0 + lane_base + r;
                    if (grow >= M) continue;
                    float4 o = *reinterpret_cast<const float4*>(pad + r * TC_EPI_PITCH + col);
                    if (cv.mode == 1) {
                        // scatter: row (n,i,j) of class (py,px) -> y[n, 2i+py, 2j+px, :] (NHWC, Cout == BN columns)
                        const int HW = cv.H * cv.W, n_img = grow / HW, rem = grow - n_img * HW;
                        const int ii = rem / cv.W, jj = rem - ii * cv.W;
                        const long orow = ((long)(n_img * 2 * cv.H + 2 * ii + (tl.nt >> 1)) * (2 * cv.W) + 2 * jj + (tl.nt & 1));
                        *reinterpret_cast<float4*>(C + orow * BN + c + col) = o;
                        continue;
                    }
                    const int n = tl.n0 + c + col;
                    float* p = out + (long)grow * N + n;
                    if ((N & 3) == 0 && n + 4 <= N) {
                        if (use_mask) {
                            const float4 m = mk[cur][k];
                            o.x = m.x > 0.f ? o.x : 0.f; o.y = m.y > 0.f ? o.y : 0.f;
                            o.z = m.z > 0.f ? o.z : 0.f; o.w = m.w > 0.f ? o.w : 0.f;
                        }
                        if (!(dbg & 64)) *reinterpret_cast<float4*>(p) = o;
                    } else {
                        const float ov[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            if (n + q < N) {
                                float val = ov[q];
                                if (mask && splits == 1 && __ldg(mask + (long)grow * N + n + q) <= 0.f) val = 0.f;
                                p[q] = val;
                            }
                    }
                }
                __syncwarp();                                  // the pad is reused by the next 32-column chunk
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}
// ------------------------------------------------------------------------------------------------------ host side
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;
static int g_tc_state = -1;            // -1 unknown, 0 unavailable, 1 available

static int tc_init() {
    if (g_tc_state >= 0) return g_tc_state;
    g_tc_state = 0;
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess || major != 10) return 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn ||
        q != cudaDriverEntryPointSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    g_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    g_tc_state = 1;
    return 1;
}

extern "C" int mmvae_has_tcgen05(void) { return tc_init(); }
extern "C" long mmvae_gemm_tc_launch_count(void) { return g_tc_launches; }
extern "C" int mmvae_gemm_tc_set_mode(int mode) {
    const int old = g_tc_mode;
    g_tc_mode = (mode & 0xff) == 1 ? 1 : 0;
    g_tc_dbg = mode >> 8;                 // ablation bits for tools/f9_ablate.py (0 in production)
    return old;
}

// row-major [rows, cols] fp32 matrix, box = 32 columns (128 bytes, one swizzle row) x box_rows
static int make_tmap(CUtensorMap* m, const float* base, long rows, long cols, int box_rows, bool mn_major) {
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * sizeof(float)};
    cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE,
                          mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

// used by convt_l4_tc.cu: K-major [rows, cols] map with a 32 x box_rows box; -1 = no tcgen05/TMA on this device
int mmvae_tc_tmap_rows(CUtensorMap* m, const float* base, long rows, long cols, int box_rows) {
    if (!tc_init()) return -1;
    return make_tmap(m, base, rows, cols, box_rows, false);
}
int mmvae_tc_dbg_bits() { return g_tc_dbg; }
// used by convt_ts.cu: any tiled map with 128-byte swizzle.  dims/box/estr: `rank` entries, innermost first; strides:
// rank-1 byte pitches.  is_bf16: element type (else fp32).  -1 = no tcgen05/TMA on this device, else the CUresult.
static int tc_encode_impl(CUtensorMap* m, int is_bf16, int rank, const void* base, const cuuint64_t* dims, const cuuint64_t* strides,
                          const cuuint32_t* box, const cuuint32_t* estr, CUtensorMapSwizzle swz) {
    if (!tc_init()) return -1;
    CUresult r = CUDA_SUCCESS;
    for (int attempt = 0; attempt < 2; ++attempt) {
        r = g_encode(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank, (void*)base,
                     dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        // a thread whose first CUDA call is this driver entry point (an autograd worker) has no context bound yet
        if (r != CUDA_ERROR_INVALID_CONTEXT) break;
        (void)cudaFree(nullptr);
    }
    return (int)r;
}
int mmvae_tc_encode(CUtensorMap* m, int is_bf16, int rank, const void* base, const cuuint64_t* dims, const cuuint64_t* strides,
                    const cuuint32_t* box, const cuuint32_t* estr) {
    return tc_encode_impl(m, is_bf16, rank, base, dims, strides, box, estr, CU_TENSOR_MAP_SWIZZLE_128B);
}
// same, rows of the box laid out linearly in shared memory (no swizzle): tiles that only threads read, channel-contiguous
int mmvae_tc_encode_linear(CUtensorMap* m, int is_bf16, int rank, const void* base, const cuuint64_t* dims,
                           const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* estr) {
    return tc_encode_impl(m, is_bf16, rank, base, dims, strides, box, estr, CU_TENSOR_MAP_SWIZZLE_NONE);
}
// MN-major operand [krows, ncols] (ncols contiguous), box = 32 columns x box_rows k-rows, SWIZZLE_128B_ATOM_32B
int mmvae_tc_tmap_mn(CUtensorMap* m, const float* base, long krows, long ncols, int box_rows) {
    if (!tc_init()) return -1;
    return make_tmap(m, base, krows, ncols, box_rows, true);
}

struct TcPlan { int BN, S, kb_per; bool swap; };

static TcPlan tc_plan(int M, int N, int K, bool can_swap) {
    TcPlan p;
    p.swap = can_swap && M <= 64 && N > M && N >= 32 && N % 4 == 0;
    const int Mm = p.swap ? N : M, Nn = p.swap ? M : N;
    p.BN = Nn <= 32 ? 32 : (Nn <= 64 ? 64 : 128);
    const long tiles = (long)((Mm + TC_BM - 1) / TC_BM) * ((Nn + p.BN - 1) / p.BN);
    const int nkb = (K + TC_BK - 1) / TC_BK;
    int S = 1;
    if (tiles < MMVAE_NUM_SMS && nkb >= 8) {                         // fill the machine
        long s = (2L * MMVAE_NUM_SMS + tiles - 1) / tiles;
        if (s > nkb / 2) s = nkb / 2;
        S = (int)(s < 1 ? 1 : s);
    }
    if ((nkb + S - 1) / S > TC_MAX_CHAIN && nkb > 2 * TC_MAX_CHAIN)   // keep the truncating TMEM chain short
        S = (nkb + TC_MAX_CHAIN - 1) / TC_MAX_CHAIN;
    p.kb_per = (nkb + S - 1) / S;
    p.S = (nkb + p.kb_per - 1) / p.kb_per;
    return p;
}

static bool tc_shape_ok(int M, int N, int K, int a_k, int b_k) {
    if ((g_tc_dbg & 128) && ((!a_k && M < 32) || (!b_k && N < 32))) return false;   // tools/gemm_small.py A/B switch
    if (K < ((g_tc_dbg & 256) ? 8 : 32)) return false;              // K = latent size: store-bound, the SIMT kernel wins (measured)
    if ((long)M * N * K < (1L << 20)) return false;                 // launch-latency-sized: SIMT is as good
    // Below ~0.4 GMAC the persistent tcgen05 kernel is latency-bound (TMEM allocation, barrier set-up, a serial k-block
    // chain on a handful of CTAs: 20-37 us where the SIMT kernel takes 9-21 us, tools/gemm_tiny.py), unless K is long
    // (split-K chains of the weight gradients) or an operand is thin (measured faster here, tools/gemm_small.py).
    // Debug bit 32768 restores the old gate.
    if (!(g_tc_dbg & 32768) && (long)M * N * K < (3L << 27) && K < 4096 && M >= 32 && N >= 32) return false;
    // TMA: 16-byte row pitch.  An inner extent below the 32-float box is fine (zero-filled past the edge): thin
    // MN-major operands of the huge-K weight gradients (N = 20) run here 3x faster than on the SIMT kernel.
    if (a_k ? (K % 4) : (M % 4 || M < 8)) return false;
    if (b_k ? (K % 4) : (N % 4 || N < 8)) return false;
    return true;
}

long mmvae_gemm_tc_workspace_bytes(int M, int N, int K, int a_k, int b_k) {
    if (!tc_shape_ok(M, N, K, a_k, b_k)) return 0;
    // the launch plans with can_swap = (no bias, no activation), which this query does not know: cover both plans
    const TcPlan p = tc_plan(M, N, K, true), q = tc_plan(M, N, K, false);
    const int S = p.S > q.S ? p.S : q.S;
    return S > 1 ? (long)S * M * N * (long)sizeof(float) : 0;
}

void mmvae_splitk_reduce(const float* ws, const float* bias, float* C, long MN, int N, int S, int act, cudaStream_t st);

template <int BN, bool A_MN, bool B_MN>
static int tc_launch(const CUtensorMap& ta, const CUtensorMap& tb, const float* bias, float* C, float* ws, int M, int N,
                     int K, int act, const TcPlan& p, int transposed, cudaStream_t st, const TcConv& cv = TcConv{0, 0, 0, 0, 0},
                     const float* mask = nullptr) {
    static bool attr_done = false;
    auto kern = gemm_tc_kernel<BN, A_MN, B_MN>;
    if (tc_init() == 0) return mmvae_set_error(1, "gemm_tc: tcgen05/TMA path unavailable on this device");
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<BN>::SMEM);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "gemm_tc: smem attribute: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    const long total = (long)((N + BN - 1) / BN) * ((M + TC_BM - 1) / TC_BM) * p.S;
    const int grid = (int)(total < MMVAE_NUM_SMS ? total : MMVAE_NUM_SMS);      // persistent: one CTA per SM
    // Debug bit 65536 runs K-major x K-major plain GEMMs with the bf16x3 split of convt_ts.cu (hi*hi + hi*mid + mid*hi, half the
    // MMA instructions of 3xTF32).  Measured (tools/gemm_bf16_ab.py): 7680x784x512 54.1 -> 48.9 us, 122880x128x512 107 -> 93 us
    // at 3.3e-6 instead of 9e-7 relative error -- this kernel is not bound by MMA issue, so 3xTF32 stays the default.
    const int passes = g_tc_mode == 1 ? 1 : ((!A_MN && !B_MN && cv.mode == 0 && (g_tc_dbg & 65536)) ? 2 : 3);
    kern<<<grid, TC_THREADS, TcCfg<BN>::SMEM, st>>>(ta, tb, bias, C, ws, M, N, K, act, p.kb_per, p.S, transposed,
                                                    passes | (g_tc_dbg << 8), cv, mask);
    MMVAE_LAUNCH_CHECK("gemm_tc");
    ++g_tc_launches;
    return 0;
}

// returns -1 when the shape / alignment is not for this path (the dispatcher then uses the SIMT kernel)
int mmvae_gemm_tc(const float* A, const float* B, const float* bias, float* C, int M, int N, int K, int a_kmajor,
                  int b_kmajor, int act, float* workspace, long workspace_bytes, cudaStream_t st) {
    if (!tc_init() || !tc_shape_ok(M, N, K, a_kmajor, b_kmajor)) return -1;
    if (((uintptr_t)A & 15) || ((uintptr_t)B & 15) || ((uintptr_t)C & 15)) return -1;
    TcPlan p = tc_plan(M, N, K, bias == nullptr && act == MMVAE_ACT_NONE);
    if (p.S > 1 && (!workspace || workspace_bytes < (long)p.S * M * N * (long)sizeof(float))) {
        p.S = 1;
        p.kb_per = (K + TC_BK - 1) / TC_BK;
    }
    // after an operand swap the kernel computes C^T = B^T A^T and stores it transposed
    const float* Ak = p.swap ? B : A;
    const float* Bk = p.swap ? A : B;
    const int Mk = p.swap ? N : M, Nk = p.swap ? M : N;
    const bool a_mn = p.swap ? !b_kmajor : !a_kmajor, b_mn = p.swap ? !a_kmajor : !b_kmajor;
    CUtensorMap ta, tb;
    int rc;
    rc = a_mn ? make_tmap(&ta, Ak, K, Mk, TC_BK, true) : make_tmap(&ta, Ak, Mk, K, TC_BM, false);
    if (rc) return mmvae_set_error(1, "gemm_tc: cuTensorMapEncodeTiled(A) failed with %d (M=%d N=%d K=%d)", rc, M, N, K);
    rc = b_mn ? make_tmap(&tb, Bk, K, Nk, TC_BK, true) : make_tmap(&tb, Bk, Nk, K, p.BN, false);
    if (rc) return mmvae_set_error(1, "gemm_tc: cuTensorMapEncodeTiled(B) failed with %d (M=%d N=%d K=%d)", rc, M, N, K);
    const int tr = p.swap ? 1 : 0;
#define TC_GO(BNv, AM, BM_)                                                                                    \
    rc = tc_launch<BNv, AM, BM_>(ta, tb, bias, C, workspace, Mk, Nk, K, act, p, tr, st)
#define TC_PICK(BNv)                                                         \
    do {                                                                     \
        if (a_mn && b_mn) TC_GO(BNv, true, true);                            \
        else if (a_mn) TC_GO(BNv, true, false);                              \
        else if (b_mn) TC_GO(BNv, false, true);                              \
        else TC_GO(BNv, false, false);                                       \
    } while (0)
    if (p.BN == 32) TC_PICK(32);
    else if (p.BN == 64) TC_PICK(64);
    else TC_PICK(128);
#undef TC_PICK
#undef TC_GO
    if (rc) return rc;
    if (p.S > 1) mmvae_splitk_reduce(workspace, bias, C, (long)M * N, N, p.S, act, st);
    return 0;
}

// ------------------------------------------------------------------------------ implicit-GEMM transposed convolution
// NHWC tensor [Nimg, Hh, Ww, C] as a 4-D map (C, Ww, Hh, Nimg); box = 32 channels x (bw x bh x bn pixels), every
// `stride`-th pixel in x and y.  K-major operand tiles use SWIZZLE_128B, MN-major ones the 32-byte-atom variant.
static int make_tmap4(CUtensorMap* m, const float* base, long Nimg, int Hh, int Ww, int C, int bw, int bh, int bn,
                      int stride, bool mn_major) {
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)Ww, (cuuint64_t)Hh, (cuuint64_t)Nimg};
    cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)Ww * C * 4, (cuuint64_t)Hh * Ww * C * 4};
    cuuint32_t box[4] = {32u, (cuuint32_t)(bw * stride), (cuuint32_t)(bh * stride), (cuuint32_t)bn};
    cuuint32_t estr[4] = {1u, (cuuint32_t)stride, (cuuint32_t)stride, 1u};
    CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (void*)base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE,
                          mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

// pixel box covering `rows` consecutive (n,y,x) positions of an H x W grid: whole images when the grid is small,
// else full-width bands of one image
static bool conv_box(int H, int W, int rows, int* bw, int* bh, int* bn) {
    const int HW = H * W;
    if (HW <= rows) {
        if (rows % HW) return false;
        *bw = W; *bh = H; *bn = rows / HW;
    } else {
        if (rows % W || HW % rows) return false;
        *bw = W; *bh = rows / W; *bn = 1;
    }
    return *bw * 2 <= 256 && *bh * 2 <= 256 && *bn <= 256;
}

extern "C" int mmvae_convt_tc_supported(int H, int W, int Cin, int Cout) {
    int bw, bh, bn;
    if (!tc_init()) return 0;
    if (Cin % 32 || Cout % 32 || Cin > 128 || Cout > 128 || (Cin & (Cin - 1)) || (Cout & (Cout - 1))) return 0;
    if (!conv_box(H, W, TC_BM, &bw, &bh, &bn) || !conv_box(H, W, 32, &bw, &bh, &bn)) return 0;
    return 1;
}

static TcPlan plan_fixed(int BN, int nkb) {
    TcPlan p;
    p.BN = BN; p.S = 1; p.kb_per = nkb; p.swap = false;
    return p;
}

// ---------------------------------------------------------------------------------------- forward, 9-shift form
// All four output-parity classes of a 128-row tile in ONE pass.  The 16 (class, tap) products of a transposed 4x4/s2
// convolution touch only 9 distinct input windows (shifts (dy,dx) in {-1,0,1}^2): the centre window feeds all four
// classes, an edge window two, a corner one.  Per 32-channel chunk the producer therefore loads 9 A tiles (not 16), the
// converter splits 9, and the issuer runs the centre as ONE UMMA of N = 4*Cout, edges as N = 2*Cout (classes are laid
// out in TMEM in Gray order (0,0),(0,1),(1,1),(1,0) so three of the four edge pairs are adjacent) and corners as N = Cout.
// Weights arrive shift-major: w9[(rowoff[s] + q)*Cout + co][ci], q = rank of the class among those the shift feeds.
__device__ __constant__ int8_t F9_DY[9] = {0, -1, 1, 0, 0, -1, -1, 1, 1};
__device__ __constant__ int8_t F9_DX[9] = {0, 0, 0, 1, -1, -1, 1, 1, -1};
__device__ __constant__ int8_t F9_NCLS[9] = {4, 2, 2, 2, 2, 1, 1, 1, 1};
__device__ __constant__ int8_t F9_ROW[9] = {0, 4, 6, 8, 10, 12, 13, 14, 15};
// MMA groups per shift: (first TMEM class slot, B row offset in classes, classes per MMA); shift 4 (dx=-1) has two
__device__ __constant__ int8_t F9_G0[9][3] = {{0, 0, 4}, {0, 0, 2}, {2, 0, 2}, {1, 0, 2}, {0, 0, 1}, {0, 0, 1}, {1, 0, 1}, {2, 0, 1}, {3, 0, 1}};
// Gray slot -> (py, px)
__device__ __constant__ int8_t F9_PY[4] = {0, 0, 1, 1};
__device__ __constant__ int8_t F9_PX[4] = {0, 1, 1, 0};

// Shared-memory plan.  The stage-release chain (TMA landing -> hi/lo split -> MMAs -> commit -> producer) is several
// thousand clocks long, so throughput = bytes in flight / chain latency: the ring must hold as many k-blocks as possible.
// A windows (hi + lo, 32 KB) rotate through SA slots.  Weights rotate through EIGHT class slots (one class = COUT rows
// x 32 channels): in the order of F9_NCLS = 4,2,2,2,2,1,1,1,1 the k-blocks of one channel chunk tile the eight slots
// exactly twice (F9_ROW % 8), so every shift has a fixed, contiguous home and the slots a k-block overwrites were last
// used F9_BACK[shift] k-blocks earlier.  Barriers are per k-block ticket (NT of each kind); MMAs retire in order, so
// "ticket j released" implies everything older is released too.
__device__ __constant__ int8_t F9_BACK[9] = {5, 4, 3, 3, 4, 4, 5, 5, 6};
template <int COUT> struct F9Cfg {
    static constexpr int CLS = COUT * TC_BK * 4;                   // one class of weights for one k-block
    static constexpr int B_HALF = 8 * CLS;                         // hi ring; the lo ring follows
    static constexpr int A_SLOT = 2 * TC_A_TILE;
    static constexpr int SA = (COUT == 32) ? 5 : 3;
    static constexpr int NT = 8;
    static constexpr int SMEM = SA * A_SLOT + 2 * B_HALF + 1024 + 256;
    static constexpr int NACC = (COUT == 32) ? 2 : 1;              // 8*COUT TMEM columns per accumulator set
    static constexpr int TMEM_COLS = 512;
};

template <int COUT>
__global__ void __launch_bounds__(TC_THREADS, 1)
convt_fwd9_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const float* __restrict__ bias, float* __restrict__ Y, int M, int H, int W, int chunks, int act,
                  int passes_dbg) {
    using Cfg = F9Cfg<COUT>;
    constexpr int SA = Cfg::SA, NT = Cfg::NT, NACC = Cfg::NACC;
    const int passes = passes_dbg & 0xff, dbg = passes_dbg >> 8;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* bring = smem + SA * Cfg::A_SLOT;
    uint64_t* bars = (uint64_t*)(bring + 2 * Cfg::B_HALF);
    uint64_t* full = bars;
    uint64_t* conv = bars + NT;
    uint64_t* empty = bars + 2 * NT;
    uint64_t* tfull = bars + 3 * NT;
    uint64_t* tempty = bars + 3 * NT + 2;
    uint32_t* tmem_slot = (uint32_t*)(bars + 3 * NT + 4);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler: the single-warp roles run on the uniform datapath
    const int tiles = (M + TC_BM - 1) / TC_BM, nk = 9 * chunks, HW = H * W;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NT; ++s) { mbar_init(full + s, 1); mbar_init(conv + s, 4); mbar_init(empty + s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull + a, 1); mbar_init(tempty + a, 4); }
        mbar_fence_init();
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
    }
    if (warp == 1) { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // producer: whole warp converged, one elected lane issues (sm100_ptx.cuh); the nine shifts are unrolled so that
        // every table entry is an immediate
        const uint32_t a_ring = smem_u32(smem), b_ring = smem_u32(bring);
        int it = 0, tk = 0, as = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
            const int m0 = t * TC_BM, n0 = m0 / HW, y0 = (m0 - n0 * HW) / W;
            for (int ck = 0; ck < chunks; ++ck) {
#pragma unroll
                for (int sh = 0; sh < 9; ++sh) {
                    constexpr int8_t kDy[9] = {0, -1, 1, 0, 0, -1, -1, 1, 1}, kDx[9] = {0, 0, 0, 1, -1, -1, 1, 1, -1};
                    constexpr int8_t kNcls[9] = {4, 2, 2, 2, 2, 1, 1, 1, 1}, kRow[9] = {0, 4, 6, 8, 10, 12, 13, 14, 15};
                    constexpr int8_t kBack[9] = {5, 4, 3, 3, 4, 4, 5, 5, 6};
                    // the A slot was last used SA k-blocks ago, the weight slots kBack[sh] k-blocks ago
                    const int back = kBack[sh] < SA ? kBack[sh] : SA, j = it - back;
                    if (j >= 0) mbar_wait(empty + (j & (NT - 1)), (j / NT) & 1);
                    const uint32_t a_hi = a_ring + as * Cfg::A_SLOT, b_hi = b_ring + (kRow[sh] & 7) * Cfg::CLS;
                    const bool la = !(dbg & 2) || it < SA, lb = !(dbg & 4) || it < 9;
                    mbar_arrive_expect_tx_elect(full + tk, (la ? TC_A_TILE : 0) + (lb ? kNcls[sh] * Cfg::CLS : 0));
                    if (la) tma_load_4d_elect(a_hi, &tmap_a, full + tk, ck * 32, kDx[sh], y0 + kDy[sh], n0);
#pragma unroll
                    for (int q = 0; q < kNcls[sh]; ++q)
                        if (lb) tma_load_2d_elect(b_hi + q * Cfg::CLS, &tmap_b, full + tk, ck * 32, (kRow[sh] + q) * COUT);
                    ++it;
                    tk = (tk + 1) & (NT - 1);
                    if (++as == SA) as = 0;
                }
            }
        }
    } else if (warp == 1) {
        // MMA issuer: whole warp converged, one elected lane issues
        constexpr uint32_t idesc1 = umma_idesc_tf32(TC_BM, COUT, 0, 0), idesc2 = umma_idesc_tf32(TC_BM, 2 * COUT, 0, 0),
                           idesc4 = umma_idesc_tf32(TC_BM, 4 * COUT, 0, 0);
        const uint64_t da0 = umma_smem_desc(smem_u32(smem), 16, 1024, 2);
        const uint64_t db0 = umma_smem_desc(smem_u32(bring), 16, 1024, 2);
        int tk = 0, ph = 0, as = 0, ti = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++ti) {
            const int acc = ti % NACC, use = ti / NACC;
            mbar_wait(tempty + acc, (use & 1) ^ 1);
            tc_fence_after_sync();
            const uint32_t d_big = tmem_base + acc * (8 * COUT), d_small = d_big + 4 * COUT;
            for (int ck = 0; ck < chunks; ++ck) {
#pragma unroll
                for (int sh = 0; sh < 9; ++sh) {
                    constexpr int8_t kRow[9] = {0, 4, 6, 8, 10, 12, 13, 14, 15};
                    // MMA groups per shift: first TMEM class slot and classes per MMA; shift 4 (dx=-1) = two single-class MMAs
                    constexpr int8_t kSlot[9] = {0, 0, 2, 1, 0, 0, 1, 2, 3}, kNc[9] = {4, 2, 2, 2, 1, 1, 1, 1, 1};
                    mbar_wait(passes == 3 ? conv + tk : full + tk, ph);
                    tc_fence_after_sync();
                    const uint64_t dah = da0 + (uint64_t)(as * (Cfg::A_SLOT >> 4)), dal = dah + (TC_A_TILE >> 4);
                    const uint64_t dbh = db0 + (uint64_t)((kRow[sh] & 7) * (Cfg::CLS >> 4)), dbl = dbh + (Cfg::B_HALF >> 4);
                    const uint32_t idesc = kNc[sh] == 4 ? idesc4 : (kNc[sh] == 2 ? idesc2 : idesc1);
#pragma unroll
                    for (int kk = 0; kk < TC_BK / 8; ++kk) {
                        if ((dbg & 16) && (ck | sh | kk) != 0) break;
                        const uint32_t first = (ck | sh | kk) != 0;      // the centre window (first) spans all four classes
#pragma unroll
                        for (int gi = 0; gi < (sh == 4 ? 2 : 1); ++gi) {
                            const uint32_t col = (gi ? 3 : kSlot[sh]) * COUT, boff = (gi * Cfg::CLS >> 4) + kk * 2;
                            if (passes == 3) {
                                umma_tf32_elect(d_small + col, dal + kk * 2, dbh + boff, idesc, first);
                                umma_tf32_elect(d_small + col, dah + kk * 2, dbl + boff, idesc, 1);
                            }
                            umma_tf32_elect(d_big + col, dah + kk * 2, dbh + boff, idesc, first);
                        }
                    }
                    umma_commit_elect(empty + tk);
                    tk = (tk + 1) & (NT - 1);
                    if (tk == 0) ph ^= 1;
                    if (++as == SA) as = 0;
                }
            }
            umma_commit_elect(tfull + acc);
        }
    } else if (warp < 6) {
        const int tq = threadIdx.x - 64;
        if (passes == 3) {
            int it = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
                for (int kb = 0; kb < nk; ++kb, ++it) {
                    const int tk = it % NT, ph = (it / NT) & 1, sh = kb % 9;
                    const int nb4 = F9_NCLS[sh] * Cfg::CLS / 16 / 128;                 // float4 per thread in the weights
                    mbar_wait(full + tk, ph);
                    if (dbg & 8) { __syncwarp(); if (lane == 0) mbar_arrive(conv + tk); continue; }
                    const uint32_t a_hi = smem_u32(smem) + (it % SA) * Cfg::A_SLOT;
                    const uint32_t b_hi = smem_u32(bring) + (F9_ROW[sh] & 7) * Cfg::CLS;
                    tf32_split_tile(a_hi, a_hi + TC_A_TILE, tq, TC_A_TILE / 16 / 128);
                    tf32_split_tile(b_hi, b_hi + Cfg::B_HALF, tq, nb4);
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(conv + tk);
                }
            }
        }
    } else {
        const int lane_base = (warp & 3) * 32;
        // the bias stays in registers for the whole kernel: reloading 32 values per 32-column chunk cost 30-90 us of
        // issue slots that the MMA issuer and the converter warps share with these warps
        float bv[COUT];
#pragma unroll
        for (int j = 0; j < COUT; ++j) bv[j] = bias ? __ldg(bias + j) : 0.f;
        int ti = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++ti) {
            const int acc = ti % NACC, use = ti / NACC;
            mbar_wait(tfull + acc, use & 1);
            tc_fence_after_sync();
            const int row = t * TC_BM + lane_base + lane;
            const int n_img = row / HW, rem = row - n_img * HW, ii = rem / W, jj = rem - ii * W;
            const uint32_t tb = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(acc * 8 * COUT);
            float* const y00 = Y + ((long)(n_img * 2 * H + 2 * ii) * (2 * W) + 2 * jj) * COUT;     // class (0,0) position
#pragma unroll
            for (int c = 0; c < 4 * COUT; c += 32) {
                float v[32];
                tmem_ld_32x32(tb + (uint32_t)c, v);
                if (passes == 3) {
                    float u[32];
                    tmem_ld_32x32(tb + (uint32_t)(4 * COUT + c), u);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += u[j];
                }
                if (c + 32 >= 4 * COUT) {
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(tempty + acc);
                }
                if (row < M && !(dbg & 1)) {
                    // each thread owns one output row of 32 contiguous floats (one full 128-byte line): direct stores
                    constexpr int kPy[4] = {0, 0, 1, 1}, kPx[4] = {0, 1, 1, 0};       // Gray slot -> (py, px)
                    const int slot = c / COUT, cc = c - slot * COUT;
                    float* p = y00 + ((long)kPy[slot] * (2 * W) + kPx[slot]) * COUT + cc;
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = apply_act_tc(v[j] + bv[cc + j], act);
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        __stcs(reinterpret_cast<float4*>(p + j), make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
}

template <int COUT>
static int fwd9_launch(const CUtensorMap& ta, const CUtensorMap& tb, const float* bias, float* y, int M, int H, int W,
                       int chunks, int act, cudaStream_t st) {
    static bool attr_done = false;
    auto kern = convt_fwd9_kernel<COUT>;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F9Cfg<COUT>::SMEM);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_fwd9: smem attribute: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    const int tiles = (M + TC_BM - 1) / TC_BM;
    const int grid = tiles < MMVAE_NUM_SMS ? tiles : MMVAE_NUM_SMS;
    kern<<<grid, TC_THREADS, F9Cfg<COUT>::SMEM, st>>>(ta, tb, bias, y, M, H, W, chunks, act,
                                                     (g_tc_mode == 1 ? 1 : 3) | (g_tc_dbg << 8));
    MMVAE_LAUNCH_CHECK("convt_fwd9");
    return 0;
}

// forward.  x: [Nimg,H,W,Cin] NHWC; wf: forward weights from mmvae_convt_weight_classes (16*Cin*Cout floats; shift-major
// for Cout in {32,64}, class-major for Cout = 128); y: [Nimg,2H,2W,Cout] NHWC = act(convT(x) + bias)
extern "C" int mmvae_convt_tc_fwd(const float* x, const float* wf, const float* bias, float* y, int Nimg, int H, int W,
                                  int Cin, int Cout, int act, void* stream) {
    MMVAE_REQUIRE(x && wf && y && Nimg > 0, "convt_tc_fwd: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_tc_supported(H, W, Cin, Cout), "convt_tc_fwd: unsupported shape H=%d W=%d Cin=%d Cout=%d", H, W, Cin, Cout);
    int bw, bh, bn;
    conv_box(H, W, TC_BM, &bw, &bh, &bn);
    CUtensorMap ta, tb;
    int rc = make_tmap4(&ta, x, Nimg, H, W, Cin, bw, bh, bn, 1, false);
    if (rc) return mmvae_set_error(1, "convt_tc_fwd: tensor map (x) failed with %d", rc);
    const int M = Nimg * H * W;
    cudaStream_t st = (cudaStream_t)stream;
    if (Cout <= 64) {
        rc = make_tmap(&tb, wf, 16L * Cout, Cin, Cout, false);
        if (rc) return mmvae_set_error(1, "convt_tc_fwd: tensor map (w9) failed with %d", rc);
        return Cout == 32 ? fwd9_launch<32>(ta, tb, bias, y, M, H, W, Cin / 32, act, st)
                          : fwd9_launch<64>(ta, tb, bias, y, M, H, W, Cin / 32, act, st);
    }
    rc = make_tmap(&tb, wf, 4L * Cout, 4L * Cin, Cout, false);
    if (rc) return mmvae_set_error(1, "convt_tc_fwd: tensor map (w) failed with %d", rc);
    const int N = 4 * Cout, K = 4 * Cin;
    const TcPlan p = plan_fixed(Cout, K / TC_BK);
    const TcConv cv{1, H, W, Cin / 32, Cout};
    return tc_launch<128, false, false>(ta, tb, bias, y, nullptr, M, N, K, act, p, 0, st, cv);
}

// dgrad.  g: [Nimg,2H,2W,Cout] NHWC gradient w.r.t. the layer's pre-activation; wp: [Cin, 16*Cout] (tap-major);
// gx: [Nimg,H,W,Cin]; x_mask: NULL or the layer input (gx zeroed where x <= 0)
extern "C" int mmvae_convt_tc_dgrad(const float* g, const float* wp, const float* x_mask, float* gx, int Nimg, int H,
                                    int W, int Cin, int Cout, void* stream) {
    MMVAE_REQUIRE(g && wp && gx && Nimg > 0, "convt_tc_dgrad: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_tc_supported(H, W, Cin, Cout), "convt_tc_dgrad: unsupported shape");
    int bw, bh, bn;
    conv_box(H, W, TC_BM, &bw, &bh, &bn);
    CUtensorMap ta, tb;
    int rc = make_tmap4(&ta, g, Nimg, 2 * H, 2 * W, Cout, bw, bh, bn, 2, false);
    if (rc) return mmvae_set_error(1, "convt_tc_dgrad: tensor map (g) failed with %d", rc);
    rc = make_tmap(&tb, wp, Cin, 16L * Cout, Cin, false);
    if (rc) return mmvae_set_error(1, "convt_tc_dgrad: tensor map (w) failed with %d", rc);
    const int M = Nimg * H * W, N = Cin, K = 16 * Cout;
    const TcPlan p = plan_fixed(Cin, K / TC_BK);
    const TcConv cv{2, H, W, Cout / 32, Cout};
    cudaStream_t st = (cudaStream_t)stream;
    if (Cin == 32) rc = tc_launch<32, false, false>(ta, tb, nullptr, gx, nullptr, M, N, K, 0, p, 0, st, cv, x_mask);
    else if (Cin == 64) rc = tc_launch<64, false, false>(ta, tb, nullptr, gx, nullptr, M, N, K, 0, p, 0, st, cv, x_mask);
    else rc = tc_launch<128, false, false>(ta, tb, nullptr, gx, nullptr, M, N, K, 0, p, 0, st, cv, x_mask);
    return rc;
}

static TcPlan plan_wgrad(int Cin, long rows) {
    TcPlan p;
    p.BN = Cin; p.swap = false;
    const int nkb = (int)((rows + 31) / 32);
    int S = (nkb + TC_MAX_CHAIN - 1) / TC_MAX_CHAIN;
    p.kb_per = (nkb + S - 1) / S;
    p.S = (nkb + p.kb_per - 1) / p.kb_per;
    return p;
}
extern "C" long mmvae_convt_tc_wgrad_workspace_bytes(int Nimg, int H, int W, int Cin, int Cout) {
    const TcPlan p = plan_wgrad(Cin, (long)Nimg * H * W);
    return p.S > 1 ? (long)p.S * 16 * Cout * Cin * (long)sizeof(float) : 0;
}
// wgrad.  x: [Nimg,H,W,Cin]; g: [Nimg,2H,2W,Cout]; gwp: [Cin, 16*Cout] (tap-major, as mmvae_permute_weight's layout)
extern "C" int mmvae_convt_tc_wgrad(const float* x, const float* g, float* gwp, float* workspace, long workspace_bytes,
                                    int Nimg, int H, int W, int Cin, int Cout, void* stream) {
    MMVAE_REQUIRE(x && g && gwp && Nimg > 0, "convt_tc_wgrad: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_tc_supported(H, W, Cin, Cout), "convt_tc_wgrad: unsupported shape");
    int bw, bh, bn;
    conv_box(H, W, 32, &bw, &bh, &bn);
    const long rows = (long)Nimg * H * W;
    TcPlan p = plan_wgrad(Cin, rows);
    MMVAE_REQUIRE(p.S == 1 || (workspace && workspace_bytes >= mmvae_convt_tc_wgrad_workspace_bytes(Nimg, H, W, Cin, Cout)),
                  "convt_tc_wgrad: workspace too small");
    CUtensorMap ta, tb;
    int rc = make_tmap4(&ta, g, Nimg, 2 * H, 2 * W, Cout, bw, bh, bn, 2, true);
    if (rc) return mmvae_set_error(1, "convt_tc_wgrad: tensor map (g) failed with %d", rc);
    rc = make_tmap(&tb, x, rows, Cin, TC_BK, true);
    if (rc) return mmvae_set_error(1, "convt_tc_wgrad: tensor map (x) failed with %d", rc);
    const int M = 16 * Cout, N = Cin;
    const int K = (int)(((rows + 31) / 32) * 32);
    const TcConv cv{3, H, W, Cout / 32, Cout};
    cudaStream_t st = (cudaStream_t)stream;
    if (Cin == 32) rc = tc_launch<32, true, true>(ta, tb, nullptr, gwp, workspace, M, N, K, 0, p, 1, st, cv);
    else if (Cin == 64) rc = tc_launch<64, true, true>(ta, tb, nullptr, gwp, workspace, M, N, K, 0, p, 1, st, cv);
    else rc = tc_launch<128, true, true>(ta, tb, nullptr, gwp, workspace, M, N, K, 0, p, 1, st, cv);
    if (rc) return rc;
    if (p.S > 1) mmvae_splitk_reduce(workspace, nullptr, gwp, (long)M * N, N, p.S, 0, st);
    return 0;
}

// forward weights.  Cout = 128: class-major wf[(py*2+px)*Cout + co][(ty*2+tx)*Cin + ci] = w[ci][co][ky][kx] with
// ky = ty ? 3 - 3*py : 1 + py  (the tap that reaches output row 2i+py from input row i + (ty ? 2*py-1 : 0)); same in x.
// Cout <= 64: shift-major w9[(F9_ROW[s] + q)*Cout + co][ci] (see convt_fwd9_kernel), class slot in Gray order.
__global__ void convt_weight_classes_kernel(const float* __restrict__ w, float* __restrict__ wf, int Cin, int Cout) {
    const long n = 16L * Cin * Cout;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cin), t = (int)((i / Cin) % 4), co = (int)((i / (4L * Cin)) % Cout), cls = (int)(i / (4L * Cin * Cout));
        const int py = cls >> 1, px = cls & 1, ty = t >> 1, tx = t & 1;
        const int ky = ty ? 3 - 3 * py : 1 + py, kx = tx ? 3 - 3 * px : 1 + px;
        wf[i] = w[(((long)ci * Cout + co) * 4 + ky) * 4 + kx];
    }
}
__global__ void convt_weight_shifts_kernel(const float* __restrict__ w, float* __restrict__ w9, int Cin, int Cout) {
    const long n = 16L * Cin * Cout;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cin), co = (int)((i / Cin) % Cout), r = (int)(i / ((long)Cin * Cout));   // r: class row 0..15
        int sh = 8;
        while (F9_ROW[sh] > r) --sh;
        const int q = r - F9_ROW[sh], dy = F9_DY[sh], dx = F9_DX[sh];
        // q-th Gray slot (ascending) among the classes this shift feeds
        int slot = -1, seen = 0;
        for (int g = 0; g < 4; ++g) {
            const bool oky = dy == 0 || (dy < 0 ? F9_PY[g] == 0 : F9_PY[g] == 1);
            const bool okx = dx == 0 || (dx < 0 ? F9_PX[g] == 0 : F9_PX[g] == 1);
            if (oky && okx) { if (seen == q) { slot = g; break; } ++seen; }
        }
        const int py = F9_PY[slot], px = F9_PX[slot];
        const int ky = dy == 0 ? 1 + py : (py == 0 ? 3 : 0), kx = dx == 0 ? 1 + px : (px == 0 ? 3 : 0);
        w9[i] = w[(((long)ci * Cout + co) * 4 + ky) * 4 + kx];
    }
}
extern "C" int mmvae_convt_weight_classes(const float* w, float* wf, int Cin, int Cout, void* stream) {
    MMVAE_REQUIRE(w && wf && Cin > 0 && Cout > 0, "convt_weight_classes: bad arguments");
    const long n = 16L * Cin * Cout;
    const int grid = (int)min((long)MMVAE_NUM_SMS * 4, (n + 255) / 256);
    if (Cout <= 64) convt_weight_shifts_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w, wf, Cin, Cout);
    else convt_weight_classes_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w, wf, Cin, Cout);
    MMVAE_LAUNCH_CHECK("convt_weight_classes");
    return 0;
}
