// ConvTranspose2d(k4, s2, p1) forward and data gradient, second generation (replaces the shared-memory-operand kernels of
// gemm_tc.cu for the decoder's middle layers, encoders.py:113-124 and their autograd backward).
//
// What bounded the first generation (profiles/r1_decoder_kernels_ncu_full_4d1e464.txt): a tcgen05.mma whose A operand is in
// shared memory costs max(32 + N/4, N/2) clocks — the 32 is the read of the 128-row A tile — so with N = Cout <= 128 and three
// tf32 passes the issue stream was operand-read-bound (tensor pipe 26-47 % active); every A window was fetched 9 / 16 times from
// L2; the hi/lo split went through shared memory.  Here (measured in tools/ts_probe.cu, profiles/r2_umma_ts_probe.txt):
//   * the A operand is read from TENSOR MEMORY ("TS" form): N/2 clocks per MMA, no shared-memory read of A at all;
//   * operands are split into bf16 (hi, mid) and the product is hi*hi + hi*mid + mid*hi in kind::f16 — K = 16 per MMA at the
//     clocks tf32 needs for K = 8; the dropped terms (mid*mid, hi*lo, lo*hi) are <= 3 * 2^-18 of |a||b| per product;
//   * ONE halo tile per 128-pixel tile and 32-channel chunk arrives by TMA (out-of-image rows / columns = the unit's zero fill);
//     the 9 (forward) or 4x4 (dgrad, per output-parity class) shifted windows are sub-views of it;
//   * eight converter warps copy, per shifted window, each pixel's row into the TMEM lane that owns the pixel (conflict-free
//     128-bit shared loads + tcgen05.st).  Forward: the landed fp32 tile is first split IN PLACE into packed bf16 pairs (row =
//     32 fp32 -> 16 hi words | 16 mid words, same 128 bytes, same swizzle), so the 9 window copies carry no arithmetic; dgrad
//     (4 windows per class tile): the split happens inside the copy -- no write-back, barrier or proxy fence;
//   * weights are prepared once per step as bf16 (hi | mid) K-major rows (mmvae_convt_ts_pack_weights) and stay RESIDENT in
//     shared memory when they fit (128 KB: the Cin*Cout = 2048 layer), else stream through a block ring;
//   * the issuing warp runs on the UNIFORM datapath (warp index through a lane-0 shuffle, TMEM addresses off a literal base):
//     one UTCHMMA per MMA, no election / R2UR chain per instruction;
//   * forward epilogue: TMEM -> registers -> bias + activation -> the warp's swizzled 4 KB tile -> one TMA store per 32-channel
//     chunk of a parity class (y as a rank-5 tensor); dgrad epilogue: warp-private transposition pads, ReLU mask fused.
// Warp roles (480 threads): 0 = activation (A) TMA producer, 1 = weight (B) TMA producer, 2 = TMEM owner + MMA issuer,
// 3-10 = converters (two warps per TMEM lane quarter, each half of the 32 channels), 11-14 = epilogue.
// TMEM: columns 0..255 accumulators (double-buffered when they need <= 128 columns; otherwise released per class slot),
// 256..511 a ring of eight A slots (32 columns: 16 hi-pair + 16 mid-pair columns = 32 channels).
// Where the clocks of a tile go: tools/ts_trace.py (TRACE instantiation, clock64 around every wait of every role).
#include <cuda.h>
#include <cudaTypedefs.h>
#include "common.cuh"
#include "sm100_ptx.cuh"

#define TS_THREADS 480
#define TS_NCONV 256                  // converter threads (warps 3..10)
#define TS_NAS 8                      // A slots in TMEM
#define TS_ND 16                      // per-window completion tickets
#define TS_A_COL 256                  // first TMEM column of the A ring
#define TS_FWD 0
#define TS_DGRAD 1

int mmvae_tc_dbg_bits();
int mmvae_tc_encode(CUtensorMap* m, int is_bf16, int rank, const void* base, const cuuint64_t* dims, const cuuint64_t* strides,
                    const cuuint32_t* box, const cuuint32_t* estr);
int mmvae_tc_encode_linear(CUtensorMap* m, int is_bf16, int rank, const void* base, const cuuint64_t* dims,
                           const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* estr);

// ---- window schedule ----------------------------------------------------------------------------------------------------
// forward: one halo tile per chunk, 9 windows (dy,dx); window s feeds kNcls[s] output-parity classes whose accumulators are
// adjacent in TMEM (Gray order of (py,px): (0,0),(0,1),(1,1),(1,0)) except dx = -1 (slots 0 and 3: two MMA groups).
// Weight blocks (one class x 32 channels: Cout rows of 128 bytes) come in consumption order: block kRow[s] + q.
// dgrad: per chunk one halo tile per output-parity class of g (stride-2 TMA box), 4 windows (ey,ex) each; block = cls*4+ey*2+ex.
template <int MODE> struct TsSched;
template <> struct TsSched<TS_FWD> {
    static constexpr int STAGES = 1, WINDOWS = 9, BLOCKS = 16;
};
template <> struct TsSched<TS_DGRAD> {
    static constexpr int STAGES = 4, WINDOWS = 4, BLOCKS = 16;
};

struct TsArgs {
    int M;            // pixels of the coarse (input) grid = Nimg * H * W
    int H, W;         // coarse grid
    int chunks;       // 32-channel chunks of the contraction (fwd: Cin/32, dgrad: Cout/32)
    int act;
    int stage_bytes;  // one landed halo tile, rounded up to 1024
    int stage_tx;     // bytes the TMA writes per tile
    int nstg;         // halo tiles in flight
    int nbs;          // weight-block slots (streaming) / 16*chunks (resident)
    int resident;
    unsigned long long* trace;   // tools/ts_trace.py: per-role clock accounting of CTA 0 (TRACE instantiation only), else NULL
    int dbg;          // ablation bits (tools/ts_time.py), 0 in production: 1 no global stores, 2 no window copies, 4 no split,
                      // 8 no MMAs, 16 no activation loads
};

// in-place split of one landed row: 32 fp32 -> 16 words of packed bf16 hi pairs | 16 words of packed bf16 mid pairs.
// cvt.rn.bf16x2.f32 d, a, b: upper half = a, lower half = b; the lower half holds the even k.
__device__ __forceinline__ void ts_split_row(uint32_t row_addr, int sw) {
    float f[32];
#pragma unroll
    for (int c = 0; c < 8; ++c)
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[4 * c]), "=f"(f[4 * c + 1]), "=f"(f[4 * c + 2]), "=f"(f[4 * c + 3])
                     : "r"(row_addr + ((uint32_t)(c ^ sw) << 4)));
    uint32_t hp[16], mp[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        bf16_split_pair(f[2 * j], f[2 * j + 1], hp[j], mp[j]);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + ((uint32_t)(c ^ sw) << 4)), "r"(hp[4 * c]), "r"(hp[4 * c + 1]),
                     "r"(hp[4 * c + 2]), "r"(hp[4 * c + 3]) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_addr + ((uint32_t)((4 + c) ^ sw) << 4)), "r"(mp[4 * c]), "r"(mp[4 * c + 1]),
                     "r"(mp[4 * c + 2]), "r"(mp[4 * c + 3]) : "memory");
    }
}

// Windows are handed from the converters to the issuer, and released back, in BATCHES (one mbarrier round trip per batch;
// per-window hand-offs cost more than the 170-clock MMA work of a window: tools/ts_time.py ablation): forward 3 batches of
// 3 windows per chunk, dgrad 2 batches of 2 windows per class tile.  The eight A slots in TMEM remain a ring of windows.
template <int MODE> struct TsBatch;
template <> struct TsBatch<TS_FWD> {
    static constexpr int BW = 3, PER_STAGE = 3, PER_CHUNK = 3;
    __host__ __device__ static constexpr int of_window(int w) { return (w / 9) * 3 + (w % 9) / 3; }
};
template <> struct TsBatch<TS_DGRAD> {
    static constexpr int BW = 2, PER_STAGE = 2, PER_CHUNK = 8;
    __host__ __device__ static constexpr int of_window(int w) { return w >> 1; }
};

__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}

__device__ __forceinline__ void tmem_zero_32x32(uint32_t taddr) {          // 32 lanes x 32 columns of zeros
    const uint32_t z = 0;
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
                 "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(z) : "memory");
}

// NB: rows of one weight block = MMA N of one class (forward: Cout) / of the tile (dgrad: Cin)
#define TS_T0() const long long _c0 = TRACE ? clock64() : 0
#define TS_T1(slot) do { if (TRACE) tacc[slot] += clock64() - _c0; } while (0)
#define TS_TRACE_OUT(role) do { if (TRACE && blockIdx.x == 0 && lane == 0) { tacc[7] = clock64() - t_role; \
        for (int k = 0; k < 8; ++k) ar.trace[(role) * 8 + k] = (unsigned long long)tacc[k]; } } while (0)

template <int MODE, int NB, bool TRACE = false>
__global__ void __launch_bounds__(TS_THREADS, 1)
convt_ts_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                const __grid_constant__ CUtensorMap tmap_y, const float* __restrict__ bias, const float* __restrict__ mask,
                float* __restrict__ out, const TsArgs ar) {
    using S = TsSched<MODE>;
    using Bt = TsBatch<MODE>;
    constexpr int BLOCK_BYTES = NB * 128;
    constexpr int ACC_COLS = (MODE == TS_FWD) ? 4 * NB : NB;
    constexpr int NACC = (2 * ACC_COLS <= TS_A_COL) ? 2 : 1;
    constexpr int BW = Bt::BW;
    // forward: a landed row is split into bf16 (hi | mid) IN PLACE once and copied by 9 windows.  dgrad: 4 windows per class
    // tile; splitting in the copy costs the same ALU work as the separate pass and saves the pass's shared-memory write-back,
    // its barrier and the proxy fence (tools/ts_trace.py: 2.3 of 8.4 k clocks per tile at 64->32 channels)
    constexpr bool INPLACE = (MODE == TS_FWD);
    static_assert(ACC_COLS <= TS_A_COL, "accumulators overlap the A ring");
    // forward: weight blocks (class rows) of the three batches of a chunk
    // Window order.  With two accumulator buffers (4*NB <= 128 columns) the centre window comes first: its MMA spans all four
    // classes and initialises them.  With ONE buffer (NB = 64) the windows are ordered by the highest class slot they touch
    // and every MMA accumulates: the epilogue zero-fills a class slot as soon as it has read it and releases it on its own
    // barrier, so the next tile's first batch (slots 0, 1) starts while slots 2, 3 are still being drained (the issuer was
    // waiting 5 of 25 k clocks per tile for the whole accumulator, tools/ts_trace.py).
    constexpr bool SLOTREL = (MODE == TS_FWD) && NACC == 1;
    constexpr int TWO = SLOTREL ? 7 : 4;                       // the dx = -1 window: classes in slots 0 and 3, two MMA groups
    constexpr int8_t kRow[9] = {SLOTREL ? 0 : 0, SLOTREL ? 1 : 4, SLOTREL ? 3 : 6, SLOTREL ? 4 : 8, SLOTREL ? 6 : 10,
                                SLOTREL ? 7 : 12, SLOTREL ? 9 : 13, SLOTREL ? 10 : 14, SLOTREL ? 12 : 15};
    constexpr int8_t kSlot[9] = {0, 0, SLOTREL ? 1 : 2, 1, SLOTREL ? 2 : 0, SLOTREL ? 2 : 0, SLOTREL ? 3 : 1, SLOTREL ? 0 : 2,
                                 SLOTREL ? 0 : 3};
    constexpr int8_t kNc[9] = {SLOTREL ? 1 : 4, 2, SLOTREL ? 1 : 2, 2, 1, SLOTREL ? 2 : 1, 1, 1, SLOTREL ? 4 : 1};
    constexpr int8_t kDy[9] = {SLOTREL ? -1 : 0, -1, SLOTREL ? -1 : 1, 0, SLOTREL ? 1 : 0, SLOTREL ? 1 : -1, SLOTREL ? 1 : -1,
                               SLOTREL ? 0 : 1, SLOTREL ? 0 : 1};
    constexpr int8_t kDx[9] = {SLOTREL ? -1 : 0, 0, SLOTREL ? 1 : 0, 1, SLOTREL ? 1 : -1, SLOTREL ? 0 : -1, SLOTREL ? -1 : 1,
                               SLOTREL ? -1 : 1, SLOTREL ? 0 : -1};
    constexpr int8_t kB0[3] = {0, SLOTREL ? 4 : 8, SLOTREL ? 9 : 13}, kBn[3] = {SLOTREL ? 4 : 8, 5, SLOTREL ? 7 : 3};

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* bring = smem + ar.nstg * ar.stage_bytes;
    uint64_t* bars = (uint64_t*)(bring + (size_t)ar.nbs * BLOCK_BYTES);
    uint64_t* full = bars;                 // [8]  halo tile landed
    uint64_t* empty = bars + 8;            // [8]  halo tile read by all converter warps
    uint64_t* a_full = bars + 16;          // [8]  a batch of windows copied into its TMEM slots
    uint64_t* done = bars + 24;            // [8]  MMAs of a batch completed (frees its A slots and weight blocks)
    uint64_t* b_full = bars + 32;          // [16] weight blocks of a batch landed
    uint64_t* tfull = bars + 48;           // [2]
    uint64_t* tempty = bars + 50;          // [2] per accumulator buffer, or [4] per class slot (SLOTREL)
    uint32_t* tmem_slot = (uint32_t*)(bars + 56);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler (uniform registers / branches in the single-warp roles)
    const int H = ar.H, W = ar.W, HW = H * W, IMGS = 128 / HW;
    const int tiles = (ar.M + 127) / 128;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 8; ++s) {
            mbar_init(full + s, 1); mbar_init(empty + s, TS_NCONV / 32); mbar_init(a_full + s, TS_NCONV / 32); mbar_init(done + s, 1);
        }
        for (int s = 0; s < 16; ++s) mbar_init(b_full + s, 1);
        for (int a = 0; a < 2; ++a) mbar_init(tfull + a, 1);
        for (int a = 0; a < 4; ++a) mbar_init(tempty + a, 4);
        mbar_fence_init();
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
    }
    float* bias_s = reinterpret_cast<float*>(bars) + 128;                     // bytes 512..1023 of the barrier block
    if (MODE == TS_FWD && threadIdx.x < NB) bias_s[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
    if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long t_role = TRACE ? clock64() : 0;
    if (TRACE && threadIdx.x == 64) {                       // every CTA: wall-clock (ns) of its start, at [64 + 2*cta]
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        ar.trace[64 + 2 * blockIdx.x] = ns;
    }

    if (warp == 0) {
        // ===== activation producer: one halo tile per (tile, chunk[, class]) =====
        const uint32_t a_ring = smem_u32(smem);
        int sidx = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
            const int n0 = t * IMGS;
            for (int ck = 0; ck < ar.chunks; ++ck)
                for (int st = 0; st < S::STAGES; ++st, ++sidx) {
                    const int s = sidx % ar.nstg, ph = (sidx / ar.nstg) & 1;
                    { TS_T0(); mbar_wait_relaxed(empty + s, ph ^ 1); TS_T1(0); }
                    if (ar.dbg & 16) { mbar_arrive_elect(full + s); continue; }
                    mbar_arrive_expect_tx_elect(full + s, ar.stage_tx);
                    if (MODE == TS_FWD) tma_load_4d_elect(a_ring + s * ar.stage_bytes, &tmap_a, full + s, ck * 32, -1, -1, n0);
                    else tma_load_4d_elect(a_ring + s * ar.stage_bytes, &tmap_a, full + s, ck * 32, -(st & 1), -(st >> 1), n0);
                }
        }
        TS_TRACE_OUT(0);
    } else if (warp == 1) {
        // ===== weight producer: the blocks of one batch per mbarrier =====
        const uint32_t b_ring = smem_u32(bring);
        const int back = (MODE == TS_FWD) ? 3 : ar.nbs / 2;          // batches after which a streaming slot is reused
        int bid = 0, ti = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++ti) {
            if (ar.resident && ti > 0) break;
            for (int ck = 0; ck < ar.chunks; ++ck) {
#pragma unroll
                for (int b = 0; b < Bt::PER_CHUNK; ++b, ++bid) {
                    const int blk0 = (MODE == TS_FWD) ? kB0[b % 3] : 2 * b, nblk = (MODE == TS_FWD) ? kBn[b % 3] : 2;
                    uint64_t* bar;
                    int slot;
                    if (ar.resident) {
                        bar = b_full + ck * Bt::PER_CHUNK + b;
                        slot = ck * 16 + blk0;
                    } else {
                        const int j = bid - back;
                        if (j >= 0) mbar_wait_relaxed(done + (j & 7), (j >> 3) & 1);
                        bar = b_full + (bid & 7);
                        slot = (MODE == TS_FWD) ? blk0 : (2 * bid) % ar.nbs;
                    }
                    mbar_arrive_expect_tx_elect(bar, nblk * BLOCK_BYTES);
#pragma unroll
                    for (int q = 0; q < nblk; ++q)
                        tma_load_2d_elect(b_ring + (slot + q) * BLOCK_BYTES, &tmap_b, bar, ck * 64, (blk0 + q) * NB);
                }
            }
        }
    } else if (warp == 2) {
        // ===== MMA issuer (whole warp converged, one elected lane issues) =====
        constexpr uint32_t idesc1 = umma_idesc_bf16(128, NB, 0, 0);
        constexpr uint32_t idesc2 = umma_idesc_bf16(128, 2 * NB <= 256 ? 2 * NB : 256, 0, 0);
        constexpr uint32_t idesc4 = umma_idesc_bf16(128, 4 * NB <= 256 ? 4 * NB : 256, 0, 0);
        const uint64_t db0 = umma_smem_desc(smem_u32(bring), 16, 1024, 2);
        // The CTA owns all 512 columns, so the allocation starts at column 0, lane 0: with the base a literal the operand
        // addresses of every MMA are functions of uniform loop counters only and stay in uniform registers (with the loaded
        // base the issuing lane spent ~40 clocks per MMA moving them there: tools/ts_trace.py, SASS R2UR chains).
        if (tmem_base != 0) __trap();
        constexpr uint32_t tmem_base_u = 0;
        int bid = 0, ti = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++ti) {
            const int acc = ti % NACC, use = ti / NACC;
            if (!SLOTREL) {
                TS_T0(); mbar_wait(tempty + acc, (use & 1) ^ 1); TS_T1(0);
                tc_fence_after_sync();
            }
            const uint32_t d_acc = tmem_base_u + acc * ACC_COLS;
            for (int ck = 0; ck < ar.chunks; ++ck) {
#pragma unroll
                for (int b = 0; b < Bt::PER_CHUNK; ++b, ++bid) {
                    const int blk0 = (MODE == TS_FWD) ? kB0[b % 3] : 2 * b;
                    if (SLOTREL && ck == 0 && b < 2) {          // batch 0 touches class slots 0, 1; batch 1 adds 2, 3 (zero-filled)
                        TS_T0();
                        mbar_wait(tempty + 2 * b, ti & 1);
                        mbar_wait(tempty + 2 * b + 1, ti & 1);
                        TS_T1(0);
                        tc_fence_after_sync();
                    }
                    { TS_T0(); mbar_wait(a_full + (bid & 7), (bid >> 3) & 1); TS_T1(1); }
                    int slot;
                    const long long _cb = TRACE ? clock64() : 0;
                    if (ar.resident) {
                        mbar_wait(b_full + ck * Bt::PER_CHUNK + b, 0);
                        slot = ck * 16 + blk0;
                    } else {
                        mbar_wait(b_full + (bid & 7), (bid >> 3) & 1);
                        slot = (MODE == TS_FWD) ? blk0 : (2 * bid) % ar.nbs;
                    }
                    if (TRACE) tacc[2] += clock64() - _cb;
                    tc_fence_after_sync();
                    const long long _ci = TRACE ? clock64() : 0;
                    const int w0 = bid * BW;                                 // global window index of the batch's first window
#pragma unroll
                    for (int wi = 0; wi < BW; ++wi) {
                        const int i = b * BW + wi;                           // window within the chunk
                        const uint32_t a_col = tmem_base_u + TS_A_COL + ((w0 + wi) & (TS_NAS - 1)) * 32;
                        const int boff = (MODE == TS_FWD) ? kRow[i % 9] - kB0[b % 3] : wi;       // block within the batch
                        const uint64_t db = db0 + (uint64_t)((slot + boff) * (BLOCK_BYTES >> 4));
                        const uint32_t idesc = (MODE == TS_DGRAD) ? idesc1 : (kNc[i % 9] == 4 ? idesc4 : (kNc[i % 9] == 2 ? idesc2 : idesc1));
#pragma unroll
                        for (int kk = 0; kk < 2; ++kk) {
                            if ((ar.dbg & 8) && (ck | i | kk) != 0) break;
                            const uint32_t first = SLOTREL ? 1u : (uint32_t)((ck | i | kk) != 0);      // the first window of a tile spans every accumulator column
#pragma unroll
                            for (int gi = 0; gi < ((MODE == TS_FWD && i == TWO) ? 2 : 1); ++gi) {
                                const uint32_t col = (MODE == TS_FWD) ? (gi ? 3 : kSlot[i % 9]) * NB : 0;
                                const uint64_t dbg = db + (uint64_t)(gi * (BLOCK_BYTES >> 4) + kk * 2);
                                umma_bf16_ts3_elect(d_acc + col, a_col + kk * 8, a_col + 16 + kk * 8, dbg, idesc, first);   // hi*hi, hi*mid, mid*hi
                            }
                        }
                    }
                    umma_commit_elect(done + (bid & 7));
                    if (TRACE) tacc[3] += clock64() - _ci;
                }
            }
            umma_commit_elect(tfull + acc);
        }
        TS_TRACE_OUT(2);
        if (TRACE && lane == 0) {
            unsigned long long ns;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
            ar.trace[64 + 2 * blockIdx.x + 1] = ns;
        }
    } else if (warp < 11) {
        // ===== converters: every warp takes part in every window; the two warps of a lane quarter split the 32 channels
        // (half h: channels 16h..16h+15 = the MMA k-step h: hi columns 8h.., mid columns 16+8h..) =====
        const int ctid = threadIdx.x - 96, half = (warp - 3) >> 2, quarter = warp & 3;
        const int L = quarter * 32 + lane, n_l = L / HW, rem = L - n_l * HW, yy = rem / W, xx = rem - yy * W;
        const int pitch = (MODE == TS_FWD) ? W + 2 : W + 1;
        const int base_row = (MODE == TS_FWD) ? (n_l * (H + 2) + yy + 1) * pitch + xx + 1 : (n_l * (H + 1) + yy) * pitch + xx;
        const int rows = ar.stage_tx >> 7;
        const uint32_t a_ring = smem_u32(smem);
        const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16) + TS_A_COL + 8 * half;
        int sidx = 0, bid = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
            for (int ck = 0; ck < ar.chunks; ++ck)
                for (int st = 0; st < S::STAGES; ++st, ++sidx) {
                    const int s = sidx % ar.nstg, ph = (sidx / ar.nstg) & 1;
                    const uint32_t base = a_ring + s * ar.stage_bytes;
                    { TS_T0(); mbar_wait(full + s, ph); TS_T1(0); }
                    if (INPLACE) {
                        {
                            TS_T0();
                            if (!(ar.dbg & 4))
                                for (int r = ctid; r < rows; r += TS_NCONV) ts_split_row(base + r * 128, r & 7);
                            TS_T1(1);
                        }
                        { TS_T0(); asm volatile("bar.sync 1, 256;" ::: "memory"); TS_T1(2); }
                    }
#pragma unroll
                    for (int b = 0; b < Bt::PER_STAGE; ++b, ++bid) {
                        const int w0 = bid * BW;
                        const int jw = w0 + BW - 1 - TS_NAS;               // the window whose slot the batch's last window reuses
                        if (jw >= 0) {
                            const int j = Bt::of_window(jw);
                            TS_T0(); mbar_wait(done + (j & 7), (j >> 3) & 1); TS_T1(3);
                        }
                        tc_fence_after_sync();
                        // (issuing the shared-memory reads before this wait, or waiting per window, measured slower: 132 / 152 us
                        // against 111 us for the 64->32 layer, tools/ts_trace.py)
                        const long long _cc = TRACE ? clock64() : 0;
                        if (!(ar.dbg & 2)) {
                            uint32_t v[BW][16];
#pragma unroll
                            for (int wi = 0; wi < BW; ++wi) {
                                const int i = b * BW + wi;                   // window within the stage
                                const int row = base_row + ((MODE == TS_FWD) ? kDy[i % 9] * pitch + kDx[i % 9] : (i >> 1) * pitch + (i & 1));
                                const uint32_t addr = base + row * 128;
                                const int sw = row & 7;
                                if (INPLACE) {
#pragma unroll
                                    for (int c = 0; c < 4; ++c) {
                                        const int lc = (c < 2) ? 2 * half + c : 4 + 2 * half + (c - 2);      // logical 16-byte chunk
                                        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[wi][4 * c]), "=r"(v[wi][4 * c + 1]),
                                                     "=r"(v[wi][4 * c + 2]), "=r"(v[wi][4 * c + 3]) : "r"(addr + ((uint32_t)(lc ^ sw) << 4)));
                                    }
                                } else {
                                    // the warp's 16 channels of the fp32 row, split into bf16 (hi, mid) pairs on the way
                                    float f[16];
#pragma unroll
                                    for (int c = 0; c < 4; ++c)
                                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[4 * c]), "=f"(f[4 * c + 1]), "=f"(f[4 * c + 2]),
                                                     "=f"(f[4 * c + 3]) : "r"(addr + ((uint32_t)((4 * half + c) ^ sw) << 4)));
#pragma unroll
                                    for (int j = 0; j < 8; ++j) {
                                        bf16_split_pair(f[2 * j], f[2 * j + 1], v[wi][j], v[wi][8 + j]);
                                    }
                                    if (ar.dbg & 4) {
#pragma unroll
                                        for (int j = 0; j < 16; ++j) v[wi][j] = __float_as_uint(f[j]);
                                    }
                                }
                            }
#pragma unroll
                            for (int wi = 0; wi < BW; ++wi) {
                                const uint32_t ta = t_lane + ((w0 + wi) & (TS_NAS - 1)) * 32;
                                tmem_st_32x8(ta, v[wi]);
                                tmem_st_32x8(ta + 16, v[wi] + 8);
                            }
                        }
                        if (!(ar.dbg & 2)) tmem_st_wait();
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(a_full + (bid & 7));
                        if (TRACE) tacc[4] += clock64() - _cc;
                    }
                    {
                        TS_T0();
                        if (INPLACE) fence_proxy_async_smem();      // generic WRITES to the tile ordered before the TMA that refills it
                        __syncwarp();
                        if (lane == 0) mbar_arrive(empty + s);
                        TS_T1(5);
                    }
                }
        }
        if (warp == 3) TS_TRACE_OUT(3);
    } else {
        // ===== epilogue (warps 11..14): TMEM -> registers -> (transposition pad) -> global =====
        // A thread reads its pixel's 32 accumulator columns; stored directly that is 32 different 128-byte lines per store
        // instruction (the LSU walks 32 tags: ~40 us of the kernel, tools/ts_time.py bit 1).  Each warp therefore transposes
        // its 32 x 32 block through a private 4 KB pad (128-byte rows, 16-byte chunks XOR-swizzled by the row: conflict-free
        // both ways) so that one store instruction covers 4 rows x 128 contiguous bytes.
        const int lane_base = (warp & 3) * 32;
        const uint32_t pad = smem_u32(bars) + 1024 + (uint32_t)(warp & 3) * 4096;
        const uint32_t my_row = pad + lane * 128, my_sw = lane & 7;
        const int tr_c = lane & 7, tr_r0 = lane >> 3;                    // transposed role: chunk tr_c of rows tr_r0 + 4k
        int ti = 0;
        if (SLOTREL) {                                           // every MMA accumulates: the accumulator starts at zero
            const uint32_t tz = tmem_base + ((uint32_t)lane_base << 16);
#pragma unroll
            for (int c = 0; c < ACC_COLS; c += 32) tmem_zero_32x32(tz + (uint32_t)c);
            tmem_st_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0)
                for (int sl = 0; sl < 4; ++sl) mbar_arrive(tempty + sl);
        }
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++ti) {
            const int acc = ti % NACC, use = ti / NACC;
            const int row0 = t * 128 + lane_base, row = row0 + lane;
            const uint32_t tb = tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(acc * ACC_COLS);
            if (MODE == TS_FWD) {
                // Each 32-channel chunk of a parity class leaves as ONE TMA store: the four warps write their pixels' rows
                // (128 bytes, 16-byte chunks XOR-swizzled by the row = SWIZZLE_128B) into one 16 KB tile, and the tensor map
                // of y scatters it to the class's pixels (2i+py, 2j+px).  No transposition reads, shuffles or global store
                // instructions in the warps (they were the bottleneck of the kernel: tools/ts_trace.py).
                { TS_T0(); mbar_wait(tfull + acc, use & 1); TS_T1(0); }
                tc_fence_after_sync();
                // every warp stores its own 32 pixels (4 KB, one box of the per-warp tensor map): no barrier between the warps
                const int R0 = lane_base, n_w = t * IMGS + R0 / HW, i_w = (R0 % HW) / W;
                // chunk j = 32 accumulator columns: slot j / (NB/32), channels (j % (NB/32))*32.  The load of chunk j+1 is in
                // flight while chunk j is processed (tensor memory reads run at 64 B/clk: 256 clocks per chunk).
                constexpr int CPS = NB / 32, NCH = 4 * CPS;
                auto process = [&](uint32_t* r, int j) {
                    const int slot = j / CPS, cc = (j % CPS) * 32;
                    const int py = slot >> 1, px = (slot & 1) ^ py;                  // Gray slot -> (py, px)
                    float v[32];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float4 b4 = *reinterpret_cast<const float4*>(bias_s + cc + 4 * c);        // broadcast
                        v[4 * c] = __uint_as_float(r[4 * c]) + b4.x; v[4 * c + 1] = __uint_as_float(r[4 * c + 1]) + b4.y;
                        v[4 * c + 2] = __uint_as_float(r[4 * c + 2]) + b4.z; v[4 * c + 3] = __uint_as_float(r[4 * c + 3]) + b4.w;
                    }
                    if (ar.act == MMVAE_ACT_RELU) {
#pragma unroll
                        for (int c = 0; c < 32; ++c) v[c] = fmaxf(v[c], 0.f);
                    } else if (ar.act == MMVAE_ACT_SIGMOID) {
#pragma unroll
                        for (int c = 0; c < 32; ++c) v[c] = 1.f / (1.f + expf(-v[c]));
                    }
                    if (lane == 0) bulk_wait_group_read0();              // the warp's previous store has read its rows
                    __syncwarp();
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((uint32_t)(c ^ my_sw) << 4)), "f"(v[4 * c]),
                                     "f"(v[4 * c + 1]), "f"(v[4 * c + 2]), "f"(v[4 * c + 3]) : "memory");
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && !(ar.dbg & 1)) {
                        tma_store_5d(&tmap_y, pad, px * NB + cc, 0, py, i_w, n_w);
                        bulk_commit_group();
                    }
                };
                uint32_t ra[32], rb[32];
                if (SLOTREL) {
                    // a class slot = CPS = 2 chunks: both into registers, the slot zero-filled and released, then processed
#pragma unroll 1
                    for (int sl = 0; sl < 4; ++sl) {
                        tmem_ld_32x32_nowait(tb + (uint32_t)(sl * NB), ra);
                        tmem_ld_32x32_nowait(tb + (uint32_t)(sl * NB + 32), rb);
                        tmem_ld_wait_regs(ra);
                        tmem_ld_wait_regs(rb);
                        tmem_zero_32x32(tb + (uint32_t)(sl * NB));
                        tmem_zero_32x32(tb + (uint32_t)(sl * NB + 32));
                        tmem_st_wait();
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty + sl);
                        process(ra, 2 * sl);
                        process(rb, 2 * sl + 1);
                    }
                } else {
                tmem_ld_32x32_nowait(tb, ra);
#pragma unroll 1
                for (int j = 0; j < NCH; j += 2) {
                    tmem_ld_wait_regs(ra);
                    tmem_ld_32x32_nowait(tb + (uint32_t)((j + 1) * 32), rb);
                    process(ra, j);
                    tmem_ld_wait_regs(rb);
                    if (j + 2 < NCH) {
                        tmem_ld_32x32_nowait(tb + (uint32_t)((j + 2) * 32), ra);
                    } else {                                             // every column of the accumulator is in registers
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty + acc);
                    }
                    process(rb, j + 1);
                }
                }
            } else {
                const bool use_mask = mask != nullptr;
                float4 mk[8];
                auto load_mask = [&](int cc) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int r = 4 * k + tr_r0;
                        mk[k] = (use_mask && row0 + r < ar.M) ? __ldg(reinterpret_cast<const float4*>(mask + (long)(row0 + r) * NB + cc + 4 * tr_c))
                                                              : make_float4(1.f, 1.f, 1.f, 1.f);
                    }
                };
                load_mask(0);                             // requested before the accumulator is complete
                { TS_T0(); mbar_wait(tfull + acc, use & 1); TS_T1(0); }
                tc_fence_after_sync();
#pragma unroll
                for (int cc = 0; cc < NB; cc += 32) {
                    float v[32];
                    tmem_ld_32x32(tb + (uint32_t)cc, v);
                    if (cc + 32 == NB) {
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty + acc);
                    }
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(my_row + ((uint32_t)(c ^ my_sw) << 4)), "f"(v[4 * c]),
                                     "f"(v[4 * c + 1]), "f"(v[4 * c + 2]), "f"(v[4 * c + 3]) : "memory");
                    __syncwarp();
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int r = 4 * k + tr_r0;
                        float4 o;
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
                                     : "r"(pad + r * 128 + ((uint32_t)(tr_c ^ (r & 7)) << 4)));
                        const float4 m = mk[k];
                        o.x = m.x > 0.f ? o.x : 0.f; o.y = m.y > 0.f ? o.y : 0.f;
                        o.z = m.z > 0.f ? o.z : 0.f; o.w = m.w > 0.f ? o.w : 0.f;
                        if (row0 + r < ar.M && !(ar.dbg & 1))
                            __stcs(reinterpret_cast<float4*>(out + (long)(row0 + r) * NB + cc + 4 * tr_c), o);
                    }
                    __syncwarp();
                    if (cc + 32 < NB) load_mask(cc + 32);
                }
            }
        }
        if (MODE == TS_FWD && lane == 0) bulk_wait_group0();
        if (warp == 11) TS_TRACE_OUT(4);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------------------ host side
#define TS_SMEM_MAX 232448            // 227 KB opt-in limit per CTA

static bool ts_geometry(int H, int W) {
    const int HW = H * W;
    return H >= 2 && W >= 2 && W <= 32 && HW <= 128 && 128 % HW == 0 && W + 2 <= 256 && H + 2 <= 256;
}

// shared-memory plan of convt_ts_kernel: weight slots (resident or streaming) + at least two halo tiles
static bool ts_plan(int mode, int NB, int chunks, int rows, int* nbs, int* resident, int* nstg, int* stage_bytes) {
    const int block = NB * 128;
    *stage_bytes = (rows * 128 + 1023) & ~1023;
    *resident = 16L * chunks * block <= 131072;
    *nbs = *resident ? 16 * chunks : (mode == TS_FWD ? 16 : 8);
    const long fixed = (long)*nbs * block + 1024 /*barriers*/ + 16384 /*epilogue pads*/ + 1024 /*alignment*/;
    long n = (TS_SMEM_MAX - fixed) / *stage_bytes;
    *nstg = (int)(n > 8 ? 8 : n);
    return *nstg >= 2;
}
static bool tw_fits(int H, int W, int Cin, int Cout);

// 1 when forward, data gradient and weight gradient of the layer all run on the kernels of this file
extern "C" int mmvae_convt_ts_supported(int H, int W, int Cin, int Cout) {
    if (!mmvae_has_tcgen05() || !ts_geometry(H, W)) return 0;
    if (Cin % 32 || Cout % 32) return 0;
    const bool fwd_ok = (Cout == 32 || Cout == 64) && Cin <= 256;       // accumulators: 4 classes x Cout <= 256 columns
    const bool dgrad_ok = (Cin == 32 || Cin == 64 || Cin == 128) && Cout <= 256;
    if (!fwd_ok || !dgrad_ok) return 0;
    const int IMGS = 128 / (H * W);
    int a, b, c, d;
    if (!ts_plan(TS_FWD, Cout, Cin / 32, IMGS * (H + 2) * (W + 2), &a, &b, &c, &d)) return 0;
    if (!ts_plan(TS_DGRAD, Cin, Cout / 32, IMGS * (H + 1) * (W + 1), &a, &b, &c, &d)) return 0;
    return tw_fits(H, W, Cin, Cout) ? 1 : 0;
}

// Weights of one layer, w [Cin][Cout][4][4] (torch layout), as the two bf16 operand matrices (row = 64-element K groups of
// [32 hi | 32 mid]):  wf [16*Cout][Cin/32 * 64]: row (kRow-ordered class row r)*Cout + co, contraction ci;
//                     wd [16*Cin ][Cout/32 * 64]: row (cls*4 + ey*2 + ex)*Cin + ci, contraction co.
__global__ void convt_ts_pack_kernel(const float* __restrict__ w, uint16_t* __restrict__ wf, uint16_t* __restrict__ wd, int Cin,
                                     int Cout) {
    const long n = 16L * Cin * Cout;
    for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long)gridDim.x * blockDim.x) {
        const int kx = (int)(idx & 3), ky = (int)((idx >> 2) & 3), co = (int)((idx >> 4) % Cout), ci = (int)(idx / (16L * Cout));
        const float v = w[idx];
        uint32_t hb, mb;
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hb) : "f"(0.f), "f"(v));
        const float r = v - __uint_as_float(hb << 16);
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(mb) : "f"(0.f), "f"(r));
        const uint16_t hi = (uint16_t)(hb & 0xffffu), mid = (uint16_t)(mb & 0xffffu);
        // forward: output class (py,px), window shift (dy,dx): ky = 1 + py - 2*dy
        {
            const int py = (ky == 1 || ky == 3) ? 0 : 1, dy = (ky == 1 || ky == 2) ? 0 : (ky == 3 ? -1 : 1);
            const int px = (kx == 1 || kx == 3) ? 0 : 1, dx = (kx == 1 || kx == 2) ? 0 : (kx == 3 ? -1 : 1);
            // window index in the kernel's order and the rank q of the class among those the window feeds (ascending Gray slot)
            const bool slotrel = Cout > 32;                  // the window order of convt_ts_kernel's single-accumulator form
            const int8_t kDyA[9] = {0, -1, 1, 0, 0, -1, -1, 1, 1}, kDxA[9] = {0, 0, 0, 1, -1, -1, 1, 1, -1};
            const int8_t kRowA[9] = {0, 4, 6, 8, 10, 12, 13, 14, 15};
            const int8_t kDyB[9] = {-1, -1, -1, 0, 1, 1, 1, 0, 0}, kDxB[9] = {-1, 0, 1, 1, 1, 0, -1, -1, 0};
            const int8_t kRowB[9] = {0, 1, 3, 4, 6, 7, 9, 10, 12};
            const int8_t* kDy = slotrel ? kDyB : kDyA;
            const int8_t* kDx = slotrel ? kDxB : kDxA;
            const int8_t* kRow = slotrel ? kRowB : kRowA;
            int sh = 0;
            while (kDy[sh] != dy || kDx[sh] != dx) ++sh;
            const int gslot = py * 2 + (px ^ py);
            int q = 0;
            for (int g = 0; g < gslot; ++g) {
                const int gpy = g >> 1, gpx = (g & 1) ^ gpy;
                const bool oky = dy == 0 || (dy < 0 ? gpy == 0 : gpy == 1), okx = dx == 0 || (dx < 0 ? gpx == 0 : gpx == 1);
                q += (oky && okx);
            }
            const long row = (long)(kRow[sh] + q) * Cout + co;
            const long col = (long)(ci >> 5) * 64 + (ci & 31);
            const long ld = (long)(Cin >> 5) * 64;
            wf[row * ld + col] = hi;
            wf[row * ld + col + 32] = mid;
        }
        // dgrad: class py = (ky even), window ey = ky >> 1 (ky = py ? 2*ey : 1 + 2*ey)
        {
            const int py = (ky & 1) ? 0 : 1, ey = ky >> 1, px = (kx & 1) ? 0 : 1, ex = kx >> 1;
            const long row = (long)((py * 2 + px) * 4 + ey * 2 + ex) * Cin + ci;
            const long col = (long)(co >> 5) * 64 + (co & 31);
            const long ld = (long)(Cout >> 5) * 64;
            wd[row * ld + col] = hi;
            wd[row * ld + col + 32] = mid;
        }
    }
}

extern "C" long mmvae_convt_ts_packed_elems(int Cin, int Cout) { return 16L * Cin * Cout * 2; }   // bf16 elements per matrix

extern "C" int mmvae_convt_ts_pack_weights(const float* w, void* wf, void* wd, int Cin, int Cout, void* stream) {
    MMVAE_REQUIRE(w && wf && wd && Cin > 0 && Cout > 0 && Cin % 32 == 0 && Cout % 32 == 0, "convt_ts_pack_weights: bad arguments");
    const long n = 16L * Cin * Cout;
    const int grid = (int)min((long)MMVAE_NUM_SMS * 4, (n + 255) / 256);
    convt_ts_pack_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(w, (uint16_t*)wf, (uint16_t*)wd, Cin, Cout);
    MMVAE_LAUNCH_CHECK("convt_ts_pack_weights");
    return 0;
}

// tools/ts_trace.py: a device buffer of 5 roles x 8 counters (clocks of CTA 0: producer, -, issuer, converter warp, epilogue warp)
static unsigned long long* g_ts_trace = nullptr;
extern "C" void mmvae_convt_ts_set_trace(void* buf) { g_ts_trace = (unsigned long long*)buf; }

template <int MODE, int NB>
static int ts_launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& ty, const float* bias, const float* mask,
                     float* out, TsArgs ar, int rows, cudaStream_t st) {
    static bool attr_done[2] = {false, false};
    const int tr = g_ts_trace != nullptr;
    auto kern = tr ? convt_ts_kernel<MODE, NB, true> : convt_ts_kernel<MODE, NB, false>;
    ar.trace = g_ts_trace;
    constexpr int BLOCK_BYTES = NB * 128;
    ar.stage_tx = rows * 128;
    ar.dbg = mmvae_tc_dbg_bits();
    if (!ts_plan(MODE, NB, ar.chunks, rows, &ar.nbs, &ar.resident, &ar.nstg, &ar.stage_bytes))
        return mmvae_set_error(1, "convt_ts: shared memory plan does not fit (stage %d bytes)", ar.stage_bytes);
    const int smem = (int)((long)ar.nbs * BLOCK_BYTES + 2048 + 16384 + (long)ar.nstg * ar.stage_bytes);
    if (!attr_done[tr]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM_MAX);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_ts: smem attribute: %s", cudaGetErrorString(e));
        attr_done[tr] = true;
    }
    const int tiles = (ar.M + 127) / 128;
    const int grid = tiles < MMVAE_NUM_SMS ? tiles : MMVAE_NUM_SMS;
    kern<<<grid, TS_THREADS, smem, st>>>(ta, tb, ty, bias, mask, out, ar);
    MMVAE_LAUNCH_CHECK("convt_ts");
    return 0;
}

static int ts_map_b(CUtensorMap* m, const void* wb, int rows, int chunks, int NB) {
    cuuint64_t dims[2] = {(cuuint64_t)chunks * 64, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)chunks * 64 * 2};
    cuuint32_t box[2] = {64u, (cuuint32_t)NB}, estr[2] = {1u, 1u};
    return mmvae_tc_encode(m, 1, 2, wb, dims, strides, box, estr);
}

// forward.  x: [Nimg,H,W,Cin] NHWC fp32; wf from mmvae_convt_ts_pack_weights; y: [Nimg,2H,2W,Cout] NHWC = act(convT(x) + bias)
extern "C" int mmvae_convt_ts_fwd(const float* x, const void* wf, const float* bias, float* y, int Nimg, int H, int W, int Cin,
                                  int Cout, int act, void* stream) {
    MMVAE_REQUIRE(x && wf && y && Nimg > 0, "convt_ts_fwd: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_ts_supported(H, W, Cin, Cout), "convt_ts_fwd: unsupported shape H=%d W=%d Cin=%d Cout=%d", H, W, Cin, Cout);
    const int IMGS = 128 / (H * W);
    CUtensorMap ta, tb;
    cuuint64_t dims[4] = {(cuuint64_t)Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Nimg};
    cuuint64_t strides[3] = {(cuuint64_t)Cin * 4, (cuuint64_t)W * Cin * 4, (cuuint64_t)H * W * Cin * 4};
    cuuint32_t box[4] = {32u, (cuuint32_t)(W + 2), (cuuint32_t)(H + 2), (cuuint32_t)IMGS}, estr[4] = {1u, 1u, 1u, 1u};
    int rc = mmvae_tc_encode(&ta, 0, 4, x, dims, strides, box, estr);
    if (rc) return mmvae_set_error(1, "convt_ts_fwd: tensor map (x) failed with %d", rc);
    rc = ts_map_b(&tb, wf, 16 * Cout, Cin / 32, Cout);
    if (rc) return mmvae_set_error(1, "convt_ts_fwd: tensor map (w) failed with %d", rc);
    TsArgs ar{};
    ar.M = Nimg * H * W; ar.H = H; ar.W = W; ar.chunks = Cin / 32; ar.act = act;
    const int rows = IMGS * (H + 2) * (W + 2);
    cudaStream_t st = (cudaStream_t)stream;
    // y as a rank-5 tensor [Nimg][H][py][W][px*Cout + co]: one parity class x 32 channels of a pixel tile is the box
    // {32, W, 1, rows, images} of ONE WARP's 32 pixels at {px*Cout + c0, 0, py, i0, n}; images past Nimg are clipped by the TMA
    CUtensorMap ty;
    {
        cuuint64_t yd[5] = {(cuuint64_t)2 * Cout, (cuuint64_t)W, 2u, (cuuint64_t)H, (cuuint64_t)Nimg};
        cuuint64_t ys[4] = {(cuuint64_t)2 * Cout * 4, (cuuint64_t)2 * W * Cout * 4, (cuuint64_t)4 * W * Cout * 4,
                            (cuuint64_t)4 * H * W * Cout * 4};
        // per-warp box: 32 consecutive pixels of the tile = 32/W image rows of one image, or 32/(H*W) whole images
        const int HWp = H * W;
        cuuint32_t yb[5] = {32u, (cuuint32_t)W, 1u, (cuuint32_t)(HWp >= 32 ? 32 / W : H), (cuuint32_t)(HWp >= 32 ? 1 : 32 / HWp)};
        cuuint32_t ye[5] = {1u, 1u, 1u, 1u, 1u};
        rc = mmvae_tc_encode(&ty, 0, 5, y, yd, ys, yb, ye);
        if (rc) return mmvae_set_error(1, "convt_ts_fwd: tensor map (y) failed with %d", rc);
    }
    return Cout == 32 ? ts_launch<TS_FWD, 32>(ta, tb, ty, bias, nullptr, y, ar, rows, st)
                      : ts_launch<TS_FWD, 64>(ta, tb, ty, bias, nullptr, y, ar, rows, st);
}

// dgrad.  g: [Nimg,2H,2W,Cout] NHWC gradient w.r.t. the layer's pre-activation; wd from mmvae_convt_ts_pack_weights;
// gx: [Nimg,H,W,Cin]; x_mask: NULL or the layer input (gx zeroed where x <= 0: the ReLU backward of the layer below)
extern "C" int mmvae_convt_ts_dgrad(const float* g, const void* wd, const float* x_mask, float* gx, int Nimg, int H, int W,
                                    int Cin, int Cout, void* stream) {
    MMVAE_REQUIRE(g && wd && gx && Nimg > 0, "convt_ts_dgrad: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_ts_supported(H, W, Cin, Cout), "convt_ts_dgrad: unsupported shape");
    const int IMGS = 128 / (H * W);
    CUtensorMap ta, tb;
    cuuint64_t dims[4] = {(cuuint64_t)Cout, (cuuint64_t)(2 * W), (cuuint64_t)(2 * H), (cuuint64_t)Nimg};
    cuuint64_t strides[3] = {(cuuint64_t)Cout * 4, (cuuint64_t)2 * W * Cout * 4, (cuuint64_t)4 * H * W * Cout * 4};
    cuuint32_t box[4] = {32u, (cuuint32_t)(2 * (W + 1)), (cuuint32_t)(2 * (H + 1)), (cuuint32_t)IMGS}, estr[4] = {1u, 2u, 2u, 1u};
    int rc = mmvae_tc_encode(&ta, 0, 4, g, dims, strides, box, estr);
    if (rc) return mmvae_set_error(1, "convt_ts_dgrad: tensor map (g) failed with %d", rc);
    rc = ts_map_b(&tb, wd, 16 * Cin, Cout / 32, Cin);
    if (rc) return mmvae_set_error(1, "convt_ts_dgrad: tensor map (w) failed with %d", rc);
    TsArgs ar{};
    ar.M = Nimg * H * W; ar.H = H; ar.W = W; ar.chunks = Cout / 32; ar.act = 0;
    const int rows = IMGS * (H + 1) * (W + 1);
    cudaStream_t st = (cudaStream_t)stream;
    if (Cin == 32) return ts_launch<TS_DGRAD, 32>(ta, tb, ta, nullptr, x_mask, gx, ar, rows, st);
    if (Cin == 64) return ts_launch<TS_DGRAD, 64>(ta, tb, ta, nullptr, x_mask, gx, ar, rows, st);
    return ts_launch<TS_DGRAD, 128>(ta, tb, ta, nullptr, x_mask, gx, ar, rows, st);
}

// ============================================================================================================== wgrad
// gw[ci, co, ky, kx] = sum over coarse pixels (n,i,j) of x[n,i,j,ci] * g[n, 2i-1+ky, 2j-1+kx, co]   (the weight gradient of
// encoders.py:113-124's ConvTranspose2d layers), and gb[co] = sum of g, in the same kernel (no extra pass over g).
// As a GEMM the contraction runs over pixels, so both operands are "MN-major" in memory (channels contiguous):
//   * A = g: M rows = (tap, co).  Per output-parity class of g one halo tile lands by a stride-2 TMA box (as in dgrad); a
//     converter thread owns one TMEM lane = one (tap, co) and reads its 16 pixels of a k-block COLUMN-wise (16 conflict-free
//     32-bit shared loads: a warp reads 32 consecutive channels of one pixel), splits them into bf16 (hi, mid) in registers,
//     packs along k and writes 8 + 8 columns of its lane with tcgen05.st: the transposition costs no extra pass;
//   * B = x: the landed [128 pixels x 32 ci] fp32 tile is split in place into [32 hi | 32 mid] bf16 per pixel row, which IS a
//     valid MN-major SWIZZLE_128B operand (rows = k); the mid half is the sub-view 64 bytes into the row (tools/ts_probe.cu);
//   * per (M tile, k-block, ci chunk) three N = 32 MMAs (hi*hi, hi*mid, mid*hi) into the same 32 accumulator columns.
// An M tile = 128 / Cout taps; a CTA ("role") owns the M tiles whose accumulators fit 256 TMEM columns (all 16 taps when
// Cin*Cout = 2048, one parity class when Cin = 128, Cout = 64) and a contiguous range of pixel tiles; every TW_CHAIN tiles the
// accumulators are drained to a partial (the TMEM accumulate truncates: chains stay <= ~312 MMAs) and a second kernel adds the
// partials in a fixed order (deterministic) straight into the torch weight layout.
#define TW_THREADS 480
#define TW_A_SLOTS 16                 // 16-column A slots (8 hi-pair + 8 mid-pair columns = one 16-pixel k-block of one M tile)
#define TW_CHAIN 14                   // pixel tiles per accumulation chain (24 MMAs per tile and accumulator)

struct TwArgs {
    int M, H, W, tiles;
    int gstage_bytes, gstage_tx, ngs, nxs;
    int ctas_per_role, tiles_per_cta, parts_per_cta, dr;
    float* ws;        // [partials][MTR*128][Cin]
    float* wsb;       // [CTAs][2][128] bias partial sums
    int dbg;
};

template <int COUT, int CIN> struct TwCfg {
    static constexpr int TPM = 128 / COUT;                       // taps per M tile
    static constexpr int MT = 16 / TPM;                          // M tiles of the layer
    static constexpr int MTR = (MT < 256 / CIN) ? MT : 256 / CIN;   // M tiles per role
    static constexpr int ROLES = MT / MTR;
    static constexpr int COC = COUT / 32, CIC = CIN / 32;
    static constexpr int MPC = 4 / TPM;                          // M tiles per parity class
    static constexpr int CPR = MTR / MPC;                        // classes per role
    static constexpr int ATILES = MTR * 8;                       // (M tile, k-block) pairs per pixel tile
    static constexpr int NBATCH = ATILES / 4;
    static constexpr int GST = CPR * COC;                        // g stages per pixel tile
    static_assert(MTR >= MPC && MTR * CIN <= 256, "role does not fit");
};

// range of pixel tiles and number of partials of CTA `cr` of a role
__host__ __device__ inline void tw_range(int cr, int tiles, int tiles_per_cta, int dr, int* t0, int* t1, int* nparts) {
    *t0 = cr * tiles_per_cta;
    *t1 = *t0 + tiles_per_cta < tiles ? *t0 + tiles_per_cta : tiles;
    if (*t1 < *t0) *t1 = *t0;
    *nparts = (*t1 - *t0 + dr - 1) / dr;
}

// WT: H == W == WT known at compile time (the rows of a k-block's 16 pixels are then immediates off one base), 0 = any geometry
template <int COUT, int CIN, int WT>
__global__ void __launch_bounds__(TW_THREADS, 1)
convt_ts_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_g, const __grid_constant__ CUtensorMap tmap_x, const TwArgs ar) {
    using C = TwCfg<COUT, CIN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* xring = smem + ar.ngs * ar.gstage_bytes;
    // offtab[v][p]: byte offset of pixel p's row in a class tile (linear rows of 128 B: the g tiles are read by threads only,
    // 32 consecutive channels of one row per warp instruction, so they need no swizzle), shifted by window v = ey*2+ex
    uint32_t* offtab = (uint32_t*)(xring + ar.nxs * 16384);
    uint64_t* bars = (uint64_t*)(offtab + 4 * 128);
    uint64_t* g_full = bars;               // [8]
    uint64_t* g_empty = bars + 8;          // [8] count 8 (converter warps)
    uint64_t* x_full = bars + 16;          // [8]
    uint64_t* x_ready = bars + 24;         // [8] count 8: split in place, visible to the tensor core
    uint64_t* x_empty = bars + 32;         // [8] MMAs that read the stage completed
    uint64_t* a_full = bars + 40;          // [4] count 8
    uint64_t* done = bars + 44;            // [4]
    uint64_t* tfull = bars + 48;
    uint64_t* tempty = bars + 49;          // count 4
    uint32_t* tmem_slot = (uint32_t*)(bars + 50);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;      // warp-uniform for the compiler (uniform registers / branches in the single-warp roles)
    const int H = ar.H, W = ar.W, HW = H * W, IMGS = 128 / HW;
    const int role = blockIdx.x % C::ROLES, cr = blockIdx.x / C::ROLES;
    int t0, t1, nparts;
    tw_range(cr, ar.tiles, ar.tiles_per_cta, ar.dr, &t0, &t1, &nparts);

    for (int e = threadIdx.x; e < 4 * 128; e += TW_THREADS) {
        const int v = e >> 7, p = e & 127, n_l = p / HW, rem = p - n_l * HW, i = rem / W, j = rem - i * W;
        const int row = (n_l * (H + 1) + i + (v >> 1)) * (W + 1) + j + (v & 1);
        offtab[e] = (uint32_t)(row * 128);
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < 8; ++s) {
            mbar_init(g_full + s, 1); mbar_init(g_empty + s, 8); mbar_init(x_full + s, 1); mbar_init(x_ready + s, 8);
            mbar_init(x_empty + s, 1);
        }
        for (int s = 0; s < 4; ++s) { mbar_init(a_full + s, 8); mbar_init(done + s, 1); }
        mbar_init(tfull, 1); mbar_init(tempty, 4);
        mbar_fence_init();
        tma_prefetch_desc(&tmap_g);
        tma_prefetch_desc(&tmap_x);
    }
    if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== g producer: per pixel tile the role's parity classes, 32-channel chunk by chunk =====
        const uint32_t g_ring = smem_u32(smem);
        int gsi = 0;
        for (int t = t0; t < t1; ++t)
            for (int cl = 0; cl < C::CPR; ++cl) {
                const int cls = role * C::CPR + cl;
                for (int cch = 0; cch < C::COC; ++cch, ++gsi) {
                    const int s = gsi % ar.ngs, ph = (gsi / ar.ngs) & 1;
                    mbar_wait_relaxed(g_empty + s, ph ^ 1);
                    mbar_arrive_expect_tx_elect(g_full + s, ar.gstage_tx);
                    tma_load_4d_elect(g_ring + s * ar.gstage_bytes, &tmap_g, g_full + s, cch * 32, -(cls & 1), -(cls >> 1), t * IMGS);
                }
            }
    } else if (warp == 1) {
        // ===== x producer: [128 pixels x 32 ci] per chunk =====
        const uint32_t x_ring = smem_u32(xring);
        int xsi = 0;
        for (int t = t0; t < t1; ++t)
            for (int c = 0; c < C::CIC; ++c, ++xsi) {
                const int s = xsi % ar.nxs, ph = (xsi / ar.nxs) & 1;
                mbar_wait_relaxed(x_empty + s, ph ^ 1);
                mbar_arrive_expect_tx_elect(x_full + s, 16384);
                tma_load_2d_elect(x_ring + s * 16384, &tmap_x, x_full + s, c * 32, t * 128);
            }
    } else if (warp == 2) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = umma_idesc_bf16(128, 32, 0, 1);             // B MN-major
        const uint64_t dx0 = umma_smem_desc(smem_u32(xring), 1024, 1024, 2);
        int gb = 0, xsi = 0, ti = 0, part = 0;
        for (int t = t0; t < t1; ++t, ++ti) {
            const bool chain_start = (ti % ar.dr) == 0;
            if (chain_start && part > 0) {                                    // the epilogue has drained the previous partial
                mbar_wait(tempty, (part - 1) & 1);
                tc_fence_after_sync();
            }
            uint64_t dxs[C::CIC];
#pragma unroll
            for (int c = 0; c < C::CIC; ++c) {
                const int s = (xsi + c) % ar.nxs, ph = ((xsi + c) / ar.nxs) & 1;
                mbar_wait(x_ready + s, ph);
                dxs[c] = dx0 + (uint64_t)(s * (16384 >> 4));
            }
            tc_fence_after_sync();
#pragma unroll 1
            for (int b = 0; b < C::NBATCH; ++b, ++gb) {
                mbar_wait(a_full + (gb & 3), (gb >> 2) & 1);
                tc_fence_after_sync();
#pragma unroll
                for (int at = 0; at < 4; ++at) {
                    const int ai = 4 * b + at, mt = ai >> 3, kb = ai & 7;
                    const uint32_t a_col = tmem_base + 256 + ((4 * gb + at) & (TW_A_SLOTS - 1)) * 16;
                    const uint32_t first = !(chain_start && kb == 0);
#pragma unroll
                    for (int c = 0; c < C::CIC; ++c) {
                        const uint32_t d = tmem_base + mt * CIN + c * 32;
                        const uint64_t db = dxs[c] + (uint64_t)(kb * (2048 >> 4));
                        if (ar.dbg & 8) continue;
                        umma_bf16_ts3_elect(d, a_col, a_col + 8, db, idesc, first);       // g_hi*x_hi, g_hi*x_mid, g_mid*x_hi
                    }
                }
                umma_commit_elect(done + (gb & 3));
            }
#pragma unroll
            for (int c = 0; c < C::CIC; ++c) umma_commit_elect(x_empty + ((xsi + c) % ar.nxs));
            xsi += C::CIC;
            if (((ti + 1) % ar.dr) == 0 || t + 1 == t1) { umma_commit_elect(tfull); ++part; }
        }
    } else if (warp < 11) {
        // ===== converters =====
        const int ctid = threadIdx.x - 96, wset = (warp - 3) >> 2, quarter = warp & 3;
        const int L = quarter * 32 + lane;
        const int tap_l = L / COUT, co = L % COUT, cch = co >> 5, w = co & 31;
        const uint32_t g_ring = smem_u32(smem), x_ring = smem_u32(xring), tab = smem_u32(offtab);
        const uint32_t addc = (uint32_t)w << 2;
        const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16) + 256;
        float bsum = 0.f;
        int gs = 0, gph = 0;                 // g stage / phase of the current class's first chunk
        int xs = 0, xph = 0;                 // x stage / phase of the current pixel tile's first chunk
        int gb = 0, pending = -1;
        for (int t = t0; t < t1; ++t) {
            // x: split in place, then hand to the tensor core
            {
                int s = xs, ph = xph;
#pragma unroll 1
                for (int c = 0; c < C::CIC; ++c) {
                    mbar_wait(x_full + s, ph);
                    if (!(ar.dbg & 4))
                        for (int r = ctid; r < 128; r += TS_NCONV) ts_split_row(x_ring + s * 16384 + r * 128, r & 7);
                    if (++s == ar.nxs) { s = 0; ph ^= 1; }
                }
                fence_proxy_async_smem();
                __syncwarp();
                s = xs;
                for (int c = 0; c < C::CIC; ++c) {
                    if (lane == 0) mbar_arrive(x_ready + s);
                    if (++s == ar.nxs) { s = 0; xph ^= 1; }
                }
                xs = s;
            }
            // A tiles: per class, per M tile of the class, two batches of four k-blocks; two k-blocks per warp set and batch
#pragma unroll 1
            for (int cl = 0; cl < C::CPR; ++cl) {
                const int cls = role * C::CPR + cl;
                int s = gs + cch, ph = gph;
                if (s >= ar.ngs) { s -= ar.ngs; ph ^= 1; }
                mbar_wait(g_full + s, ph);
                const uint32_t gbase = g_ring + s * ar.gstage_bytes + addc;
#pragma unroll 1
                for (int mi = 0; mi < C::MPC; ++mi) {
                    const int tap = (cls * C::MPC + mi) * C::TPM + tap_l;               // = cls*4 + ey*2 + ex
                    const int v = tap & 3;
                    const bool bias_tap = v == cls;                                       // (ey,ex) == (py,px)
#pragma unroll 1
                    for (int kb4 = 0; kb4 < 2; ++kb4, ++gb) {
                        // loads of this batch are issued before the previous batch's tcgen05.st are waited for and handed
                        // over: the TMEM-store latency and the shared-memory round trips overlap
                        float f[2][16];
#pragma unroll
                        for (int a2 = 0; a2 < 2; ++a2) {
                            const int kb = kb4 * 4 + 2 * a2 + wset;
                            const uint32_t tb = tab + (uint32_t)(v * 128 + kb * 16) * 4;
                            if (!(ar.dbg & 2)) {
                                if (WT) {
                                    uint32_t b0;
                                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(b0) : "r"(tb));
                                    b0 += gbase;
#pragma unroll
                                    for (int k = 0; k < 16; ++k)
                                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(f[a2][k])
                                                     : "r"(b0 + (uint32_t)(((k / (WT ? WT : 1)) * (WT + 1) + (k % (WT ? WT : 1))) * 128)));
                                } else {
                                    uint32_t off[16];
#pragma unroll
                                    for (int k = 0; k < 16; k += 4)
                                        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                                                     : "=r"(off[k]), "=r"(off[k + 1]), "=r"(off[k + 2]), "=r"(off[k + 3]) : "r"(tb + k * 4));
#pragma unroll
                                    for (int k = 0; k < 16; ++k)
                                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(f[a2][k]) : "r"(gbase + off[k]));
                                }
                            } else {
#pragma unroll
                                for (int k = 0; k < 16; ++k) f[a2][k] = 0.f;
                            }
                        }
                        if (pending >= 0) {
                            tmem_st_wait();
                            tc_fence_before_sync();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(a_full + (pending & 3));
                        }
                        if (gb >= 4) mbar_wait(done + (gb & 3), ((gb >> 2) - 1) & 1);
                        tc_fence_after_sync();
#pragma unroll
                        for (int a2 = 0; a2 < 2; ++a2) {
                            const int at = 2 * a2 + wset;
                            if (bias_tap) {
#pragma unroll
                                for (int k = 0; k < 16; ++k) bsum += f[a2][k];
                            }
                            uint32_t hm[16];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                bf16_split_pair(f[a2][2 * j], f[a2][2 * j + 1], hm[j], hm[8 + j]);
                            }
                            const uint32_t ta = t_lane + ((4 * gb + at) & (TW_A_SLOTS - 1)) * 16;
                            tmem_st_32x8(ta, hm);
                            tmem_st_32x8(ta + 8, hm + 8);
                        }
                        pending = gb;
                    }
                }
                // the class's g stages are free once its last k-block has been copied
                __syncwarp();
                s = gs;
                for (int q = 0; q < C::COC; ++q) {
                    if (lane == 0) mbar_arrive(g_empty + s);
                    if (++s == ar.ngs) { s = 0; gph ^= 1; }
                }
                gs = s;
            }
        }
        if (pending >= 0) {
            tmem_st_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(a_full + (pending & 3));
        }
        ar.wsb[((long)blockIdx.x * 2 + wset) * 128 + L] = bsum;
    } else {
        // ===== epilogue: drain the accumulators of a finished chain into its partial =====
        const int lane_base = (warp & 3) * 32;
        for (int part = 0; part < nparts; ++part) {
            mbar_wait_relaxed(tfull, part & 1);
            tc_fence_after_sync();
            float* wp = ar.ws + ((long)blockIdx.x * ar.parts_per_cta + part) * (C::MTR * 128 * CIN);
#pragma unroll 1
            for (int mt = 0; mt < C::MTR; ++mt) {
                float* prow = wp + (long)(mt * 128 + lane_base + lane) * CIN;
#pragma unroll
                for (int cc = 0; cc < CIN; cc += 32) {
                    float v[32];
                    tmem_ld_32x32(tmem_base + ((uint32_t)lane_base << 16) + (uint32_t)(mt * CIN + cc), v);
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<float4*>(prow + cc + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                }
            }
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty);
        }
        // partials this CTA's (shorter) range does not produce are zero, so that the reduction needs no range logic
        for (int part = nparts; part < ar.parts_per_cta; ++part) {
            float4* wp = reinterpret_cast<float4*>(ar.ws + ((long)blockIdx.x * ar.parts_per_cta + part) * (C::MTR * 128 * CIN));
            for (int i = threadIdx.x - 352; i < C::MTR * 128 * CIN / 4; i += 128) wp[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// partials -> gw in the torch layout [Cin][Cout][4][4] and gb[Cout], in a fixed summation order.  One block per row
// m = tap*Cout + co of the lowered gradient (256 / Cin threads share a column and split the partials), one more for the bias.
template <int COUT, int CIN>
__global__ void __launch_bounds__(256) convt_ts_wgrad_reduce_kernel(const TwArgs ar, float* __restrict__ gw, float* __restrict__ gb,
                                                                    int grid_main) {
    using C = TwCfg<COUT, CIN>;
    __shared__ float red[256];
    const int tid = threadIdx.x;
    if (blockIdx.x == 16 * COUT) {
        if (!gb) return;
        const int l = tid & 127, h = tid >> 7;
        float s0 = 0.f, s1 = 0.f;
        int b = h;
        for (; b + 2 < grid_main * 2; b += 4) { s0 += ar.wsb[(long)b * 128 + l]; s1 += ar.wsb[(long)(b + 2) * 128 + l]; }
        for (; b < grid_main * 2; b += 2) s0 += ar.wsb[(long)b * 128 + l];
        red[tid] = s0 + s1;
        __syncthreads();
        if (tid < COUT) {
            float t = 0.f;
            for (int j = tid; j < 256; j += COUT) t += red[j];
            gb[tid] = t;
        }
        return;
    }
    constexpr int NPG = 256 / CIN;
    const int m = blockIdx.x, ci = tid % CIN, pg = tid / CIN;
    const int tap = m / COUT, co = m % COUT;
    const int mt = tap / C::TPM, role = mt / C::MTR, m_local = (mt % C::MTR) * 128 + (tap % C::TPM) * COUT + co;
    const int total = ar.ctas_per_role * ar.parts_per_cta;              // partial q of the role belongs to CTA q / parts_per_cta
    const long pstride = (long)C::MTR * 128 * CIN;
    const float* base = ar.ws + (long)m_local * CIN + ci;
    auto paddr = [&](int q) {
        const int cr = q / ar.parts_per_cta, p = q - cr * ar.parts_per_cta;
        return base + ((long)(cr * C::ROLES + role) * ar.parts_per_cta + p) * pstride;
    };
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    int q = pg;
    for (; q + 3 * NPG < total; q += 4 * NPG) {                          // four independent loads in flight
        const float a0 = __ldg(paddr(q)), a1 = __ldg(paddr(q + NPG)), a2 = __ldg(paddr(q + 2 * NPG)), a3 = __ldg(paddr(q + 3 * NPG));
        acc[0] += a0; acc[1] += a1; acc[2] += a2; acc[3] += a3;
    }
    for (; q < total; q += NPG) acc[0] += __ldg(paddr(q));
    red[tid] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    __syncthreads();
    if (pg == 0) {
        float s = 0.f;
#pragma unroll
        for (int g = 0; g < NPG; ++g) s += red[g * CIN + ci];
        const int cls = tap >> 2, ey = (tap >> 1) & 1, ex = tap & 1, py = cls >> 1, px = cls & 1;
        const int ky = py ? 2 * ey : 1 + 2 * ey, kx = px ? 2 * ex : 1 + 2 * ex;
        gw[(((long)ci * COUT + co) * 4 + ky) * 4 + kx] = s;
    }
}

template <int COUT, int CIN> static void tw_plan(TwArgs* ar, int* grid, int* smem, int rows_g) {
    using C = TwCfg<COUT, CIN>;
    *grid = (MMVAE_NUM_SMS / C::ROLES) * C::ROLES;
    ar->ctas_per_role = *grid / C::ROLES;
    if (ar->ctas_per_role > ar->tiles) { ar->ctas_per_role = ar->tiles; *grid = ar->ctas_per_role * C::ROLES; }
    ar->tiles_per_cta = (ar->tiles + ar->ctas_per_role - 1) / ar->ctas_per_role;
    ar->parts_per_cta = (ar->tiles_per_cta + TW_CHAIN - 1) / TW_CHAIN;
    ar->dr = (ar->tiles_per_cta + ar->parts_per_cta - 1) / ar->parts_per_cta;
    ar->gstage_tx = rows_g * 128;
    ar->gstage_bytes = (ar->gstage_tx + 1023) & ~1023;
    // shared memory: x ring of up to two pixel tiles, g ring from what is left (at least one pixel tile of each)
    const long fixed = 1024 /*align*/ + 2048 /*offtab*/ + 512 /*barriers*/;
    ar->nxs = 2 * C::CIC > 8 ? 8 : 2 * C::CIC;
    for (;;) {
        ar->ngs = (int)((TS_SMEM_MAX - fixed - (long)ar->nxs * 16384) / ar->gstage_bytes);
        if (ar->ngs >= C::GST + 1 || ar->nxs <= C::CIC) break;
        --ar->nxs;
    }
    if (ar->ngs > 8) ar->ngs = 8;
    *smem = (int)(fixed + (long)ar->nxs * 16384 + (long)ar->ngs * ar->gstage_bytes);
}

template <int COUT, int CIN> static long tw_ws_floats(int Nimg, int H, int W, long* bias_off) {
    using C = TwCfg<COUT, CIN>;
    TwArgs ar{};
    ar.tiles = (Nimg * H * W + 127) / 128;
    int grid, smem;
    tw_plan<COUT, CIN>(&ar, &grid, &smem, (128 / (H * W)) * (H + 1) * (W + 1));
    const long nws = (long)grid * ar.parts_per_cta * C::MTR * 128 * CIN;
    if (bias_off) *bias_off = nws;
    return nws + (long)grid * 2 * 128;
}

template <int COUT, int CIN>
static int tw_launch(const float* x, const float* g, float* gw, float* gb, float* ws, long ws_bytes, int Nimg, int H, int W,
                     cudaStream_t st) {
    using C = TwCfg<COUT, CIN>;
    static bool attr_done[2] = {false, false};
    constexpr int WF = (COUT == 32 && CIN == 64) ? 8 : ((COUT == 64 && CIN == 128) ? 4 : 0);      // the layers of the SVHN decoder
    const int fast = WF && H == WF && W == WF;
    auto kern = fast ? convt_ts_wgrad_kernel<COUT, CIN, WF> : convt_ts_wgrad_kernel<COUT, CIN, 0>;
    const int IMGS = 128 / (H * W);
    TwArgs ar{};
    ar.M = Nimg * H * W; ar.H = H; ar.W = W; ar.tiles = (ar.M + 127) / 128;
    ar.dbg = mmvae_tc_dbg_bits();
    int grid, smem;
    const int rows_g = IMGS * (H + 1) * (W + 1);
    tw_plan<COUT, CIN>(&ar, &grid, &smem, rows_g);
    if (ar.ngs < C::GST || ar.nxs < C::CIC) return mmvae_set_error(1, "convt_ts_wgrad: shared memory plan does not fit");
    long bias_off;
    const long need = tw_ws_floats<COUT, CIN>(Nimg, H, W, &bias_off) * (long)sizeof(float);
    if (!ws || ws_bytes < need) return mmvae_set_error(1, "convt_ts_wgrad: workspace %ld bytes, need %ld", ws_bytes, need);
    ar.ws = ws; ar.wsb = ws + bias_off;
    CUtensorMap tg, tx;
    cuuint64_t dims[4] = {(cuuint64_t)COUT, (cuuint64_t)(2 * W), (cuuint64_t)(2 * H), (cuuint64_t)Nimg};
    cuuint64_t strides[3] = {(cuuint64_t)COUT * 4, (cuuint64_t)2 * W * COUT * 4, (cuuint64_t)4 * H * W * COUT * 4};
    cuuint32_t box[4] = {32u, (cuuint32_t)(2 * (W + 1)), (cuuint32_t)(2 * (H + 1)), (cuuint32_t)IMGS}, estr[4] = {1u, 2u, 2u, 1u};
    int rc = mmvae_tc_encode_linear(&tg, 0, 4, g, dims, strides, box, estr);
    if (rc) return mmvae_set_error(1, "convt_ts_wgrad: tensor map (g) failed with %d", rc);
    cuuint64_t xd[2] = {(cuuint64_t)CIN, (cuuint64_t)ar.M};
    cuuint64_t xs[1] = {(cuuint64_t)CIN * 4};
    cuuint32_t xb[2] = {32u, 128u}, xe[2] = {1u, 1u};
    rc = mmvae_tc_encode(&tx, 0, 2, x, xd, xs, xb, xe);
    if (rc) return mmvae_set_error(1, "convt_ts_wgrad: tensor map (x) failed with %d", rc);
    if (!attr_done[fast]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TS_SMEM_MAX);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_ts_wgrad: smem attribute: %s", cudaGetErrorString(e));
        attr_done[fast] = true;
    }
    kern<<<grid, TW_THREADS, smem, st>>>(tg, tx, ar);
    MMVAE_LAUNCH_CHECK("convt_ts_wgrad");
    convt_ts_wgrad_reduce_kernel<COUT, CIN><<<16 * COUT + 1, 256, 0, st>>>(ar, gw, gb, grid);
    MMVAE_LAUNCH_CHECK("convt_ts_wgrad_reduce");
    return 0;
}

template <int COUT, int CIN> static int tw_fits_t(int H, int W) {
    using C = TwCfg<COUT, CIN>;
    TwArgs ar{};
    ar.tiles = 1;
    int grid, smem;
    tw_plan<COUT, CIN>(&ar, &grid, &smem, (128 / (H * W)) * (H + 1) * (W + 1));
    return ar.ngs >= C::GST && ar.nxs >= C::CIC;
}
static bool tw_fits(int H, int W, int Cin, int Cout) {
    if (Cin > 128 || (Cout != 32 && Cout != 64)) return false;
    return (Cout == 32 ? (Cin == 32 ? tw_fits_t<32, 32>(H, W) : Cin == 64 ? tw_fits_t<32, 64>(H, W) : tw_fits_t<32, 128>(H, W))
                       : (Cin == 32 ? tw_fits_t<64, 32>(H, W) : Cin == 64 ? tw_fits_t<64, 64>(H, W) : tw_fits_t<64, 128>(H, W))) != 0;
}

#define TW_DISPATCH(FN, ...)                                                             \
    (Cout == 32 ? (Cin == 32 ? FN<32, 32>(__VA_ARGS__) : Cin == 64 ? FN<32, 64>(__VA_ARGS__) : FN<32, 128>(__VA_ARGS__)) \
                : (Cin == 32 ? FN<64, 32>(__VA_ARGS__) : Cin == 64 ? FN<64, 64>(__VA_ARGS__) : FN<64, 128>(__VA_ARGS__)))

extern "C" long mmvae_convt_ts_wgrad_workspace_bytes(int Nimg, int H, int W, int Cin, int Cout) {
    if (!mmvae_convt_ts_supported(H, W, Cin, Cout) || Cin > 128) return 0;
    return TW_DISPATCH(tw_ws_floats, Nimg, H, W, nullptr) * (long)sizeof(float);
}

// wgrad.  x: [Nimg,H,W,Cin]; g: [Nimg,2H,2W,Cout] (gradient w.r.t. the pre-activation), NHWC fp32; gw: torch layout
// (Cin,Cout,4,4); gb: [Cout] or NULL.  workspace >= mmvae_convt_ts_wgrad_workspace_bytes().
extern "C" int mmvae_convt_ts_wgrad(const float* x, const float* g, float* gw, float* gb, float* workspace, long workspace_bytes,
                                    int Nimg, int H, int W, int Cin, int Cout, void* stream) {
    MMVAE_REQUIRE(x && g && gw && Nimg > 0, "convt_ts_wgrad: bad arguments");
    MMVAE_REQUIRE(mmvae_convt_ts_supported(H, W, Cin, Cout) && Cin <= 128, "convt_ts_wgrad: unsupported shape");
    cudaStream_t st = (cudaStream_t)stream;
    return TW_DISPATCH(tw_launch, x, g, gw, gb, workspace, workspace_bytes, Nimg, H, W, st);
}

// ============================================================================================== last-layer weight gradient
// ConvTranspose2d(32, Cout <= 8, 4, 2, 1) on a 16x16 grid (encoders.py:124-128, the image-producing layer): weight + bias
// gradient with the same operand scheme as convt_ts_wgrad_kernel -- g transposed into TENSOR MEMORY by converter threads (one
// TMEM lane = one (tap, co) row of the lowered gradient, 16*Cout = 48 of 128 rows used, 12 per lane quarter so that every
// converter warp has work), x split in place into a bf16 MN-major operand, three N = 32 MMAs per 16-pixel k-block -- but with g
// as the NCHW image gradient: per image g (and the layer output y) arrive by two bulk copies, the converter warps form the
// pre-activation gradient g * act'(y) once into a zero-bordered shared-memory image (and the bias gradient from it), and a lane
// reads its 16 pixels of a k-block (one row of the 16x16 grid) at stride 2 from that image.  Replaces the fp32-FMA kernel of
// convt_small.cu (154 us at 7680 images, the largest kernel of the C2 step) for this shape; partials are written in the
// layout of convt_small_wgrad_reduce (one per CTA: chains of 12 pixel tiles are drained into registers, the TMEM accumulate
// truncates), so the fixed-order reduction is shared.
#define L4W_NXS 6                      // x tiles (128 pixels x 32 ci, 16 KB) in flight
#define L4W_GP 36                      // pitch of the bordered pre-activation-gradient image (34 rows x 36 floats per channel)
#define L4W_CHAIN 12                   // pixel tiles per TMEM accumulation chain (24 MMAs each)
#define L4W_NGB 4                      // g / y images in flight

struct L4wArgs {
    const float* g;
    const float* y;
    float* part;
    int n_img, act, imgs_per_cta, dbg;
};

__device__ __forceinline__ float l4w_dact(float y, int act) {
    if (act == MMVAE_ACT_RELU) return y > 0.f ? 1.f : 0.f;
    if (act == MMVAE_ACT_SIGMOID) return y * (1.f - y);
    return 1.f;
}

// Warp roles (480 threads).  A TMEM lane belongs to the warps with (warp % 4) == lane / 32, and the 48 rows sit in lanes 0..23 and
// 32..55 (24 per quarter), so the EIGHT warps with warp % 4 in {0, 1} are the converters: four sets of two, one 16-pixel k-block
// per set and batch.  2 = TMEM owner + MMA issuer, 3 = g / y producer, 6 = x producer, 7 / 10 / 11 / 14 = builders: they form the
// pre-activation-gradient image of the next image (and the bias gradient) and split every landed x tile in place.  The accumulation chains are drained by converter set 0 (its warps own the lanes of the rows).
template <int COUT>
__global__ void __launch_bounds__(TW_THREADS, 1)
convt_l4w_ts_kernel(const __grid_constant__ CUtensorMap tmap_x, const L4wArgs ar) {
    constexpr int ROWS = 16 * COUT, RPQ = (ROWS + 1) / 2, IMG = COUT * 1024, PLANE = 34 * L4W_GP, TOT = 32 * COUT * 16;
    static_assert(RPQ <= 32, "too many output channels");
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* xring = smem;                                           // [L4W_NXS][16384]
    float* graw = (float*)(xring + L4W_NXS * 16384);                 // [L4W_NGB][g IMG | y IMG]
    float* gs = graw + L4W_NGB * 2 * IMG;                            // [2][COUT][34][L4W_GP], borders stay zero
    constexpr int GSZ = (COUT * PLANE + 3) & ~3;
    uint64_t* bars = (uint64_t*)(gs + 2 * GSZ);
    uint64_t* graw_full = bars;            // [4]
    uint64_t* graw_empty = bars + 4;       // [4] count 4 (builder warps)
    uint64_t* x_full = bars + 8;           // [8]
    uint64_t* x_ready = bars + 16;         // [8] count 4: split in place by the splitter warps
    uint64_t* x_empty = bars + 24;         // [8]
    uint64_t* a_full = bars + 32;          // [4] count 8
    uint64_t* done = bars + 36;            // [4]
    uint64_t* tfull = bars + 40;           // [2]
    uint64_t* tempty = bars + 42;          // [2] count 2 (the two warps of converter set 0)
    uint64_t* gs_full = bars + 44;         // [2] count 4: image built
    uint64_t* gs_empty = bars + 46;        // [2] count 8: every converter warp has read it
    uint32_t* tmem_slot = (uint32_t*)(bars + 48);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int n0 = min(ar.n_img, (int)blockIdx.x * ar.imgs_per_cta), n1 = min(ar.n_img, n0 + ar.imgs_per_cta);
    const int ntiles = 2 * (n1 - n0);

    for (int e = threadIdx.x; e < 2 * GSZ; e += TW_THREADS) gs[e] = 0.f;
    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) { mbar_init(tfull + s, 1); mbar_init(tempty + s, 2); mbar_init(gs_full + s, 4); mbar_init(gs_empty + s, 8); }
        for (int s = 0; s < L4W_NGB; ++s) { mbar_init(graw_full + s, 1); mbar_init(graw_empty + s, 4); }
        for (int s = 0; s < 8; ++s) { mbar_init(x_full + s, 1); mbar_init(x_ready + s, 4); mbar_init(x_empty + s, 1); }
        for (int s = 0; s < 4; ++s) { mbar_init(a_full + s, 8); mbar_init(done + s, 1); }
        mbar_fence_init();
        tma_prefetch_desc(&tmap_x);
    }
    if (warp == 2) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const bool is_conv = (warp & 3) < 2 && warp < 14;                // warps 0,1,4,5,8,9,12,13

    if (warp == 3) {
        // ===== g / y producer: one image = two bulk copies =====
        const uint32_t bytes = (uint32_t)(IMG * 4 * (ar.y ? 2 : 1));
        for (int n = n0, it = 0; n < n1; ++n, ++it) {
            const int buf = it % L4W_NGB;
            mbar_wait_relaxed(graw_empty + buf, ((it / L4W_NGB) & 1) ^ 1);
            if (ar.dbg & 16) { if (lane == 0) mbar_arrive(graw_full + buf); __syncwarp(); continue; }
            if (lane == 0) {
                mbar_arrive_expect_tx(graw_full + buf, bytes);
                bulk_load_1d(graw + buf * 2 * IMG, ar.g + (long)n * IMG, IMG * 4, graw_full + buf);
                if (ar.y) bulk_load_1d(graw + buf * 2 * IMG + IMG, ar.y + (long)n * IMG, IMG * 4, graw_full + buf);
            }
            __syncwarp();
        }
    } else if (warp == 6) {
        // ===== x producer: [128 pixels x 32 ci] per tile, two tiles per image =====
        const uint32_t x_ring = smem_u32(xring);
        for (int ti = 0; ti < ntiles; ++ti) {
            const int s = ti % L4W_NXS, ph = (ti / L4W_NXS) & 1;
            mbar_wait_relaxed(x_empty + s, ph ^ 1);
            mbar_arrive_expect_tx_elect(x_full + s, 16384);
            tma_load_2d_elect(x_ring + s * 16384, &tmap_x, x_full + s, 0, (n0 + (ti >> 1)) * 256 + (ti & 1) * 128);
        }
    } else if (warp == 2) {
        // ===== MMA issuer (uniform datapath, literal TMEM base: see convt_ts_kernel) =====
        if (tmem_base != 0) __trap();
        constexpr uint32_t idesc = umma_idesc_bf16(128, 32, 0, 1);             // B MN-major
        const uint64_t dx0 = umma_smem_desc(smem_u32(xring), 1024, 1024, 2);
        int gb = 0, part = 0;
        for (int ti = 0; ti < ntiles; ++ti) {
            const bool chain_start = (ti % L4W_CHAIN) == 0;
            const int acc = part & 1;
            if (chain_start && part >= 2) {                                    // this buffer's previous chain has been drained
                mbar_wait(tempty + acc, ((part >> 1) - 1) & 1);
                tc_fence_after_sync();
            }
            const int s = ti % L4W_NXS, ph = (ti / L4W_NXS) & 1;
            mbar_wait(x_ready + s, ph);
            tc_fence_after_sync();
            const uint64_t dx = dx0 + (uint64_t)(s * (16384 >> 4));
#pragma unroll 1
            for (int b = 0; b < 2; ++b, ++gb) {
                mbar_wait(a_full + (gb & 3), (gb >> 2) & 1);
                tc_fence_after_sync();
#pragma unroll
                for (int at = 0; at < 4; ++at) {
                    const int kb = 4 * b + at;
                    const uint32_t a_col = 256u + (uint32_t)(((4 * gb + at) & (TW_A_SLOTS - 1)) * 16);
                    const uint32_t first = !(chain_start && kb == 0);
                    if (ar.dbg & 8) continue;
                    umma_bf16_ts3_elect((uint32_t)(acc * 32), a_col, a_col + 8, dx + (uint64_t)(kb * (2048 >> 4)), idesc, first);
                }
                umma_commit_elect(done + (gb & 3));
            }
            umma_commit_elect(x_empty + s);
            if (((ti + 1) % L4W_CHAIN) == 0 || ti + 1 == ntiles) { umma_commit_elect(tfull + acc); ++part; }
        }
    } else if (is_conv) {
        // ===== converters: build the pre-activation gradient image, transpose g into tensor memory; set 0 drains the chains =====
        const int quarter = warp & 3, wset = warp >> 2;
        const int r = quarter * RPQ + lane;                                    // row of the lowered gradient owned by this lane
        const bool live = lane < RPQ && r < ROWS;
        const int tap = live ? r / COUT : 0, co = live ? r % COUT : 0, ky = tap >> 2, kx = tap & 3;
        const uint32_t g_lane = smem_u32(gs) + (uint32_t)((co * PLANE + ky * L4W_GP + kx) * 4);
        const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        float accr[32];                                                        // set 0: this row's weight gradient over all chains
#pragma unroll
        for (int j = 0; j < 32; ++j) accr[j] = 0.f;
        int gb = 0, part = 0;
        for (int n = n0, it = 0; n < n1; ++n, ++it) {
            mbar_wait(gs_full + (it & 1), (it >> 1) & 1);
            const uint32_t g_img = g_lane + (uint32_t)((it & 1) * GSZ * 4);
#pragma unroll 1
            for (int b4 = 0; b4 < 4; ++b4, ++gb) {                          // batch = 4 k-blocks (one per set) of tile b4 / 2
                const int irow = 4 * b4 + wset;                              // row i of the 16x16 grid = this set's k-block
                const uint32_t base = g_img + (uint32_t)(2 * irow * L4W_GP * 4);
                float f[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    if (live && !(ar.dbg & 2)) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(f[k]) : "r"(base + (uint32_t)(8 * k)));
                    else f[k] = 0.f;
                }
                if (gb >= 4) mbar_wait(done + (gb & 3), ((gb >> 2) - 1) & 1);
                tc_fence_after_sync();
                uint32_t hm[16];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    bf16_split_pair(f[2 * j], f[2 * j + 1], hm[j], hm[8 + j]);
                }
                const uint32_t ta = t_lane + 256 + ((4 * gb + wset) & (TW_A_SLOTS - 1)) * 16;
                tmem_st_32x8(ta, hm);
                tmem_st_32x8(ta + 8, hm + 8);
                tmem_st_wait();
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full + (gb & 3));
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(gs_empty + (it & 1));                  // the last reads of this image are in registers
            // chain ends after this image's second tile?  set 0 drains it (the MMAs of the chain complete a little later)
            const int ti_last = 2 * it + 1;
            if (wset == 0 && (((ti_last + 1) % L4W_CHAIN) == 0 || ti_last + 1 == ntiles)) {
                const int acc = part & 1;
                mbar_wait_relaxed(tfull + acc, (part >> 1) & 1);
                tc_fence_after_sync();
                float v[32];
                tmem_ld_32x32(t_lane + (uint32_t)(acc * 32), v);
#pragma unroll
                for (int j = 0; j < 32; ++j) accr[j] += v[j];
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(tempty + acc);
                ++part;
            }
        }
        if (wset == 0 && live) {
            float* out = ar.part + (long)blockIdx.x * (TOT + COUT);
#pragma unroll
            for (int ci = 0; ci < 32; ++ci) out[(ci * COUT + co) * 16 + tap] = accr[ci];
        }
    } else if (warp == 7 || warp == 10 || warp == 11 || warp == 14) {
        // ===== builders: g * act'(y) of the NEXT image into its zero-bordered buffer (+ bias sums), x tiles split in place =====
        const uint32_t x_ring = smem_u32(xring);
        const int sidx = ((warp == 7 ? 0 : warp == 10 ? 1 : warp == 11 ? 2 : 3) << 5) + lane;
        float bsum[COUT];
#pragma unroll
        for (int c = 0; c < COUT; ++c) bsum[c] = 0.f;
        auto build = [&](int it) {
            const int buf = it % L4W_NGB;
            mbar_wait(gs_empty + (it & 1), ((it >> 1) & 1) ^ 1);              // the converters have read image it - 2
            mbar_wait(graw_full + buf, (it / L4W_NGB) & 1);
            const float4* g4 = reinterpret_cast<const float4*>(graw + buf * 2 * IMG);
            const float4* y4 = reinterpret_cast<const float4*>(graw + buf * 2 * IMG + IMG);
            float* gd = gs + (it & 1) * GSZ;
            for (int i = sidx; i < IMG / 4 && !(ar.dbg & 64); i += 128) {
                const int e = 4 * i, xx = e & 31, yy = (e >> 5) & 31, c = e >> 10;
                float4 v = g4[i];
                if (ar.y) {
                    const float4 yv = y4[i];
                    v.x *= l4w_dact(yv.x, ar.act); v.y *= l4w_dact(yv.y, ar.act); v.z *= l4w_dact(yv.z, ar.act); v.w *= l4w_dact(yv.w, ar.act);
                }
                float* d = gd + c * PLANE + (yy + 1) * L4W_GP + xx + 1;
                d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
                const float sv = (v.x + v.y) + (v.z + v.w);
#pragma unroll
                for (int cc = 0; cc < COUT; ++cc) bsum[cc] += (cc == c) ? sv : 0.f;
            }
            __syncwarp();
            if (lane == 0) { mbar_arrive(gs_full + (it & 1)); mbar_arrive(graw_empty + buf); }
        };
        if (n0 < n1) build(0);
        for (int n = n0, it = 0; n < n1; ++n, ++it) {
            for (int h = 0; h < 2; ++h) {
                const int ti = 2 * it + h, s = ti % L4W_NXS, ph = (ti / L4W_NXS) & 1;
                mbar_wait(x_full + s, ph);
                if (!(ar.dbg & 4)) ts_split_row(x_ring + s * 16384 + sidx * 128, sidx & 7);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(x_ready + s);
            }
            if (n + 1 < n1) build(it + 1);
        }
        // bias partials of the 128 builder threads -> shared memory (the g staging area is free once EVERY builder warp has
        // read the last image)
        asm volatile("bar.sync 2, 128;" ::: "memory");
#pragma unroll
        for (int c = 0; c < COUT; ++c) graw[c * 128 + sidx] = bsum[c];
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
    if (n0 >= n1 && threadIdx.x < TOT) {                                       // a CTA without images still owns a partial
        for (int i = threadIdx.x; i < TOT; i += TW_THREADS) ar.part[(long)blockIdx.x * (TOT + COUT) + i] = 0.f;
    }
    if (threadIdx.x < COUT) {                                                  // bias: fixed order over the converter threads
        float s = 0.f;
        for (int i = 0; i < 128; ++i) s += graw[threadIdx.x * 128 + i];
        ar.part[(long)blockIdx.x * (TOT + COUT) + TOT + threadIdx.x] = s;
    }
}

// partials [P][Cin*Cout*16 + Cout] in the layout of convt_small_wgrad_reduce.  -1: not this kernel's shape.
int mmvae_convt_l4_wgrad_ts(const float* x, const float* g, const float* y_act, float* part, int P, int N, int H, int W, int Cin,
                            int Cout, int act, cudaStream_t st) {
    if (!mmvae_has_tcgen05() || H != 16 || W != 16 || Cin != 32 || Cout != 3 || P < 1) return -1;
    if ((((uintptr_t)g | (uintptr_t)x | (uintptr_t)y_act) & 15) != 0) return -1;
    constexpr int COUT = 3;
    CUtensorMap tx;
    cuuint64_t xd[2] = {32u, (cuuint64_t)N * 256};
    cuuint64_t xs[1] = {128u};
    cuuint32_t xb[2] = {32u, 128u}, xe[2] = {1u, 1u};
    int rc = mmvae_tc_encode(&tx, 0, 2, x, xd, xs, xb, xe);
    if (rc) return mmvae_set_error(1, "convt_l4_wgrad_ts: tensor map (x) failed with %d", rc);
    L4wArgs ar{};
    ar.g = g; ar.y = y_act; ar.part = part; ar.n_img = N; ar.act = act; ar.imgs_per_cta = (N + P - 1) / P;
    ar.dbg = mmvae_tc_dbg_bits();
    const int smem = 1024 + L4W_NXS * 16384 + L4W_NGB * 2 * COUT * 1024 * 4 + 2 * ((COUT * 34 * L4W_GP + 3) & ~3) * 4 + 512;
    static bool attr_done = false;
    auto kern = convt_l4w_ts_kernel<COUT>;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return mmvae_set_error(1000 + (int)e, "convt_l4_wgrad_ts: smem attribute: %s", cudaGetErrorString(e));
        attr_done = true;
    }
    kern<<<P, TW_THREADS, smem, st>>>(tx, ar);
    MMVAE_LAUNCH_CHECK("convt_l4_wgrad_ts");
    return 0;
}
