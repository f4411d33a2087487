// Fused ResBlock iterations, tcgen05 / TMEM implicit GEMM for sm_100a (the VO2W_PREC_BF16 hot loop).
//
//   y = sum_s [ x_s + conv2_s(lrelu(conv1_s(xa_s) + b1_s, 0.1)) + b2_s ],   xa_s = lrelu(x_s, 0.1)
//
// One segment s is one iteration of ResBlock.forward (scripts/hifigan/models.py:96-103); three segments
// (the last iteration of the three ResBlocks of a stage) give the MRF sum of models.py:154-160 in one launch.
//
// Orientation: time is the UMMA M dimension (TMEM lane = sample), output channels N, input channels K;
// a convolution tap is the SAME shared-memory tile read at a row offset (SWIZZLE_NONE K-major core
// matrices, any 16-byte aligned start address), so dilation halos cost no copies.
//
// One persistent CTA (CG=1) or CTA pair (CG=2, tcgen05 cta_group::2, M=256 = 2 x 128 rows, the weight
// operand split N/2 per CTA) walks tiles of NB x 128 rows of one utterance.  Warp roles (pair_umma.cuh), per tile and segment:
//   warp 12 lane 0: bulk-copies the input tile (128*NB + 2*h1 rows, all channels) into one of up to 4 shared-memory
//                   buffers and prefetches later tiles into L2
//   warp 13 lane 0: packed weights: copied once per kernel where a launch's convolutions fit shared memory (C <= 64),
//                   else streamed through an mbarrier ring (every stage is used by all NB row blocks)
//   warps 14-15   : leader CTA: one elected thread per warp issues tcgen05.mma (split issue: warp 14 every conv1, warp 15
//                   every conv2; else both walk the same schedule and split the row blocks);
//                   peer CTA: warp 14 relays its "bytes have landed" barriers to the leader
//   warps 8-11    : e1: D1 (TMEM) + b1 -> lrelu -> bf16 -> shared memory (the conv2 operand never leaves the SM)
//   warps 0-7     : e2 (after the last segment): D2 + sum b2 + sum residuals -> bf16 stream in HBM (or conv_post + tanh
//                   -> waveform); single-segment launches: the residual is written INTO D2 before conv2 (RINIT)
// Accumulators: ND copies of D1[NB], D2[NB] in TMEM (2*ND*NB*C <= 512 columns; ND = 2 for C <= 64: conv1 of unit n+1
// is issued before conv2 of unit n); conv2 of all segments accumulates into D2.  e2 of a tile overlaps the MMAs of the
// next tile; e1 is on the critical path of its segment where ND = 1.
#include "pair_umma.cuh"
#include "umma_cg.cuh"

namespace {

// barrier slots inside the static array
enum {
    // "landed" barriers of the LEADER collect two arrivals per phase: its own producer's (with the byte count) and the
    // peer CTA's relay (the peer's tile / weights landed), so the issuer waits once per operand, not twice
    BAR_XA_FULL = 0,                                  // [4]  input tile landed (complete_tx)
    BAR_XA_EMPTY = 4,                                 // [4]  conv1 finished reading it (tcgen05.commit)
    BAR_XA_FULL_PEER = 8,                             // [4]  leader only: peer's input tile landed
    BAR_W_FULL = 12,                                  // [PAIR_W_BARS] ring slot (or, weights resident: convolution) landed
    BAR_W_EMPTY = BAR_W_FULL + PAIR_W_BARS,
    BAR_W_FULL_PEER = BAR_W_EMPTY + PAIR_W_BARS,      // leader only
    // the accumulators and the conv2 operand buffer exist ND times (ND = 2 where TMEM has the room, C <= 64):
    // the barriers below are indexed by the buffer, unit n uses buffer n % ND with phase (n / ND) & 1
    BAR_D1_FULL = BAR_W_FULL_PEER + PAIR_W_BARS,      // [2] conv1 accumulators complete
    BAR_T_FULL = BAR_D1_FULL + 2,                     // [2][4] leader only: 64-channel block kb of the conv2 operand
    BAR_D2_FULL = BAR_T_FULL + 8,                     //        written by every e1 warp of the pair
    BAR_D2_EMPTY = BAR_D2_FULL + 2,                   // [2] leader only: every e2 warp of the pair has read D2
    // split issue (one thread per convolution, see p.split): the orderings the single issuer gets from in-order execution
    BAR_D1_EMPTY = BAR_D2_EMPTY + 2,                  // [2] leader only: every e1 warp of the pair has read D1
    BAR_T_EMPTY = BAR_D1_EMPTY + 2,                   // [2] conv2 has finished reading the operand buffer (tcgen05.commit)
    BAR_COUNT = BAR_T_EMPTY + 2
};

struct TileCoord {
    int b;          // utterance
    int o0;         // first sample this tile is responsible for
    int l0;         // first sample of the tile's rows (= o0, or o0 - 3 where conv_post rides the launch: its 3-sample halo)
    bool live;      // false: padding iteration (the pair must stay in lock step), nothing is stored
};
__device__ __forceinline__ TileCoord tile_of(const PairP& p, int it) {
    int g = (int)blockIdx.x + it * (int)gridDim.x;
    TileCoord t;
    t.live = g < p.n_tiles;
    if (!t.live) g = p.n_tiles - 1;
    // g / tiles_per_utt without the ~35-instruction integer division (every role calls this once or twice per tile):
    // float estimate, exact after one correction step (g < 2^31, the estimate is off by at most 1)
    int b = __float2int_rd(__int2float_rn(g) * p.inv_tiles_per_utt);
    int r = g - b * p.tiles_per_utt;
    if (r < 0) { --b; r += p.tiles_per_utt; }
    else if (r >= p.tiles_per_utt) { ++b; r -= p.tiles_per_utt; }
    t.b = b;
    t.o0 = r * p.Vs;
    t.l0 = p.post ? max(t.o0 - PAIR_POST_HALO, 0) : t.o0;   // the first tile starts at sample 0: rows before it are conv_post's zero padding
    return t;
}

// Cycle accounting and event traces exist only in builds made with -DVO2W_PAIR_PROFILE: the product kernel (and the
// experiments build with its what-if switches) carries no clock reads and no counters -- they cost registers.
#ifdef VO2W_PAIR_PROFILE
constexpr bool PAIR_PROF = true;
#else
constexpr bool PAIR_PROF = false;
#endif
__device__ __forceinline__ long long prof_clock() { return PAIR_PROF ? clock64() : 0LL; }
// VO2W_PAIR_PROF=1: event trace of CTA 0 (the leader of pair 0): trace[role][unit][2] timestamps (clock64)
#define PAIR_TR_UNITS 48
#define PAIR_TR_ROLES 10     // 0 input producer, 1 conv1 issue, 2 conv2 issue, 3 e1, 4 e2, 5 weight producer
__device__ __forceinline__ void trace(const PairP& p, int role, int unit, int slot) {
    if (PAIR_PROF && p.prof && blockIdx.x == 0 && unit < PAIR_TR_UNITS && (threadIdx.x & 31) == 0)
        p.prof[(size_t)gridDim.x * 16 + ((size_t)role * PAIR_TR_UNITS + unit) * 2 + slot] = (unsigned long long)prof_clock();
}

// One convolution = the bias item followed by the (tap, 64-channel block) weight items, each applied to all
// NB row blocks.
// second: reads the conv2 operand buffer (dilation 1; block c > 0 starts at row c*128 - 2*h2max).
// keep:   accumulate on top of what the accumulators hold (conv2 of segments after the first).
// The whole warp walks the (warp-uniform) loops so that addresses stay in uniform registers; one elected
// lane issues.  NI issuer warps split the row blocks (issue, not execution, limits the small-N stages).  All descriptor arithmetic is in 16-byte units on the low descriptor word.
// Runs in ONE thread (the elected lane of an issuer warp): the tensor pipe queues ~24 MMAs, so a thread that issues
// them a few cycles apart runs ahead of the pipe and its barrier waits / commits overlap queued work.
template <int CG, int C, int NB, int KS, int NI>
__device__ __forceinline__ void issue_conv(const PairP& p, const bool prof, int issuer, uint32_t bar0, uint32_t a_addr, uint32_t a_rows,
                                           int n_items, int dil, bool second, bool keep, uint32_t d_tmem, uint32_t ring_addr,
                                           uint32_t idesc, uint32_t ones_lo, uint32_t& wit, long long& w_wait,
                                           int k_taps, uint32_t t_bar, uint32_t t_parity, long long& t_wait, int widx, uint32_t& w_seen,
                                           int tr_unit = -1) {
    constexpr uint32_t n_loc = (uint32_t)C / CG;
    constexpr uint32_t KB = C < 64 ? C : 64;
    constexpr uint32_t item_units = n_loc * KB * 2u / 16u;                    // 16-byte units of one weight item
    constexpr uint32_t b_lbo_field = n_loc << 16;                             // LBO = n_loc * 16 B
    constexpr int MY = NB / NI;                                               // row blocks of this issuer
    const uint32_t a_base = ((a_addr >> 4) & 0x3FFFu) | (a_rows << 16);      // LBO = a_rows * 16 B
    const uint32_t kb_rows = (KB >> 3) * a_rows;                              // 16-byte units per 64-channel block
    const uint32_t blk_back = second ? 2u * (uint32_t)p.h2max : 0u;
    const uint32_t slot_mask = (uint32_t)p.n_slots - 1u, slot_shift = (uint32_t)p.slots_log2;
    const int per_stage = p.items_per_stage;
    const bool no_mma = (p.dbg & 4) != 0;
    // per (row block, K step) offsets of the A descriptor and the accumulators of this issuer: loop invariants
    uint32_t a_off[MY][KS], d_blk[MY];
#pragma unroll
    for (int bi = 0; bi < MY; ++bi) {
        const int blk = issuer + bi * NI;
        d_blk[bi] = d_tmem + (uint32_t)(blk * C);
#pragma unroll
        for (int s = 0; s < KS; ++s) a_off[bi][s] = (uint32_t)(blk * 128) - (blk > 0 ? blk_back : 0u) + (uint32_t)s * 2u * a_rows;
    }
    uint32_t tap = 0, kb = 0;                                                 // items 1.. run kb-major: (kb, tap)
    uint32_t a_lo = a_base;                                                   // descriptor low word of (kb, tap), block 0, step 0
    int item = 0;
    while (item < n_items) {
        // weights resident (p.resident): convolution `widx` was copied to shared memory once, at kernel start, and its
        // barrier stays complete (phase 0) for good: no ring slot, no per-stage wait / commit / relay
        const uint32_t slot = p.resident ? (uint32_t)widx : (wit & slot_mask), par = p.resident ? 0u : ((wit >> slot_shift) & 1u);
        long long tw = 0;
        if (prof) tw = prof_clock();
        if (!p.resident || !((w_seen >> widx) & 1u)) {   // resident weights are awaited once per convolution and kernel
            mbar_wait_cluster(bar0 + 8u * (BAR_W_FULL + slot), par);
            w_seen |= 1u << widx;
        }
        if (prof) { w_wait += prof_clock() - tw; if (tr_unit >= 0 && item == 0) trace(p, 6, tr_unit, 0); }
        const int cnt = p.resident ? n_items : min(per_stage, n_items - item);
        uint32_t b_lo = (((ring_addr + (p.resident ? p.wres_off[widx] : slot * p.stage_bytes)) >> 4) & 0x3FFFu) | b_lbo_field;
        int i = 0;
        if (item == 0) {
            // item 0 is the bias: D = ones[128 x 16] . (b_hi, b_lo, 0, ...)^T initialises the accumulators
#pragma unroll
            for (int bi = 0; bi < MY; ++bi) mma_lo<CG>(d_blk[bi], ones_lo, b_lo, idesc, keep ? 1u : 0u);
            b_lo += item_units;
            i = 1;
        }
        for (; i < cnt; ++i) {
            if (second && tap == 0) {
                // conv2 reads channel block kb of the operand buffer: wait until e1 has written it (the later
                // blocks are still being converted while these MMAs run)
                long long tt = 0;
                if (prof) tt = prof_clock();
                mbar_wait_cluster(t_bar + 8u * kb, t_parity);
                if (prof) t_wait += prof_clock() - tt;
                tc_fence_after();
            }
            if (!no_mma) {
#pragma unroll
                for (int bi = 0; bi < MY; ++bi) {
#pragma unroll
                    for (int s = 0; s < KS; ++s)
                        mma_lo<CG>(d_blk[bi], a_lo + a_off[bi][s], b_lo + (uint32_t)s * 2u * n_loc, idesc, 1u);
                }
            }
            b_lo += item_units;
            a_lo += (uint32_t)dil;
            if (++tap == (uint32_t)k_taps) { tap = 0; ++kb; a_lo = a_base + kb * kb_rows; }
        }
        if (prof && tr_unit >= 0 && item == 0) trace(p, 6, tr_unit, 1);
        if (!p.resident) {
            mma_commit<CG>(bar0 + 8u * (BAR_W_EMPTY + slot)); // ring slot reusable once these MMAs retire
            ++wit;
        }
        item += cnt;
    }
}

// RINIT (= p.rinit, single-segment launches): the residual enters D2 by initialisation, see e2.  A template parameter so
// that each launch kind carries only its own e2 code: the generic drain (three residual tiles in registers) sets the
// register pressure of the whole kernel, and its spills were reloaded in every role's loop -- with 227 KB of the SM's
// L1 configured as shared memory a spill reload is an L2 round trip (ncu: ~23 % of e2's stall samples at C = 32).
template <int CG, int C, bool POST, bool RINIT>
__global__ void __launch_bounds__(PAIR_THREADS, 1) pair_umma_kernel(const __grid_constant__ PairP p) {
    static_assert(!POST || C == 32, "conv_post rides the last stage only (32 channels: one TMEM load holds a whole row)");
    constexpr int NB = C >= 256 ? 1 : (C >= 64 ? 2 : 4);     // 128-row accumulator blocks per tile (2*NB*C <= 512 TMEM columns)
    constexpr int KS = C < 64 ? 2 : 4;                       // K=16 steps per weight item (64 input channels, 32 for C=32)
    constexpr int GROUPS = C / 32;                           // 32 accumulator columns per tcgen05.ld
    constexpr int CHUNKS = C / 8;
    constexpr int NI = NB >= 2 ? 2 : 1;                      // MMA issuer warps (they split the row blocks)
    constexpr int ND = PAIR_ND(C);                           // copies of D1 / D2 / conv2 operand buffer (skewed schedule if 2)
    static_assert(2 * ND * NB * C <= 512, "accumulators exceed TMEM");
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[BAR_COUNT];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
    const uint32_t sm0 = smem_u32(smem);
    const uint32_t bar0 = smem_u32(bars);
    const int n_seg = p.n_seg;
    const int n_units = p.iters * n_seg;             // (tile, segment) units this CTA walks

    if (threadIdx.x == 0) {
        for (int i = 0; i < BAR_COUNT; ++i)
            {
            const bool by_issuers = (i >= BAR_XA_EMPTY && i < BAR_XA_EMPTY + 4) || (i >= BAR_W_EMPTY && i < BAR_W_EMPTY + PAIR_W_BARS) ||
                                    (i >= BAR_D1_FULL && i < BAR_D1_FULL + 2) || (i >= BAR_D2_FULL && i < BAR_D2_FULL + 2);   // one tcgen05.commit per issuer warp
            const bool landed = (i >= BAR_XA_FULL && i < BAR_XA_FULL + 4) || (i >= BAR_W_FULL && i < BAR_W_FULL + PAIR_W_BARS);
            uint32_t cnt = 1;
            if ((i >= BAR_T_FULL && i < BAR_T_FULL + 8) || (i >= BAR_D1_EMPTY && i < BAR_D1_EMPTY + 2)) cnt = CG * PAIR_E1_WARPS;
            else if (i >= BAR_D2_EMPTY && i < BAR_D2_EMPTY + 2) cnt = CG * PAIR_E2_WARPS;
            else if (by_issuers) cnt = p.split ? 1 : NI;
            if (RINIT && i >= BAR_XA_EMPTY && i < BAR_XA_EMPTY + 4) cnt += PAIR_E2_WARPS;   // + this CTA's e2 warps (they read the tile)
            else if (landed && CG == 2 && rank == 0) cnt = 2;
            mbar_init(bar0 + 8u * i, cnt);
        }
        fence_mbar_init();
    }
    // the next kernel of the stream may be scheduled as soon as SMs free up (its set-up overlaps this grid's tail)
    if (threadIdx.x == 0) griddep_launch_dependents();
    if (warp == PAIR_WARP_TMEM) tmem_alloc_cg<CG>(smem_u32(&tmem_base_s), 512u);
    // A operand of the bias MMA: 128 rows x K=16 with ones in the first two K columns (bias = hi + lo in bf16)
    if (threadIdx.x < 256)
        st_shared_v4(sm0 + p.ones_off + threadIdx.x * 16u, make_uint4(threadIdx.x < 128 ? 0x3F803F80u : 0u, 0u, 0u, 0u));
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    if (warp == PAIR_WARP_XA) {
        // ================= input tile producer =================
        if (lane == 0) {
            constexpr int chunks = CHUNKS;
            // The smem tile of unit n + 2 can only be requested once conv1 of unit n has retired, which leaves one
            // DRAM round trip per two units on the critical path.  So the rows of unit n + pf are pulled into L2
            // (no shared memory needed) well ahead, and the bulk copy into shared memory is an L2 hit.
            auto prefetch_unit = [&](int m) {
                if (m >= n_units) return;
                const int it = m / n_seg, s = m - it * n_seg;
                const TileCoord t = tile_of(p, it);
                if (!t.live) return;
                const PairSegP& sg = p.seg[s];
                const int start = VO2W_PAD_ROWS + t.l0 - sg.h2 - sg.h1;
                const int rows = min(sg.xa_rows, p.Lp - start);
                const __nv_bfloat16* src = sg.xa + ((size_t)t.b * chunks * p.Lp + start) * 8;
                for (int c = 0; c < chunks; ++c) prefetch_l2(src + (size_t)c * p.Lp * 8, (uint32_t)rows * 16u);
            };
            griddep_wait();                              // the input is the previous launch's output
            for (int m = 2; m < 2 + p.pf; ++m) prefetch_unit(m);
            int n = 0;                                   // running (tile, segment) counter -> buffer and phase
            for (int it = 0; it < p.iters; ++it) {
                const TileCoord t = tile_of(p, it);
                for (int s = 0; s < n_seg; ++s, ++n) {
                    const PairSegP& sg = p.seg[s];
                    const int use = n / p.xa_bufs, buf = n - use * p.xa_bufs;
                    if (p.pf) prefetch_unit(n + 2 + p.pf);
                    trace(p, 0, n, 0);
                    if (use > 0) mbar_wait(bar0 + 8u * (BAR_XA_EMPTY + buf), (uint32_t)((use - 1) & 1));
                    trace(p, 0, n, 1);
                    const int start = VO2W_PAD_ROWS + t.l0 - sg.h2 - sg.h1;   // first staged row inside the padded axis
                    const int rows = min(sg.xa_rows, p.Lp - start);
                    const uint32_t row_bytes = (uint32_t)rows * 16u;
                    const uint32_t full = bar0 + 8u * (BAR_XA_FULL + buf);
                    mbar_arrive_expect_tx(full, row_bytes * (uint32_t)chunks);
                    const __nv_bfloat16* src = sg.xa + ((size_t)t.b * chunks * p.Lp + start) * 8;
                    const uint32_t dst = sm0 + p.xa_off[buf];
                    for (int c = 0; c < chunks; ++c)
                        bulk_g2s(dst + (uint32_t)(c * sg.xa_rows) * 16u, src + (size_t)c * p.Lp * 8, row_bytes, full);
                }
            }
        }
    } else if (warp == PAIR_WARP_W) {
        // ================= weight producer (same conv order as the MMA issuers) =================
        if (lane == 0) {
            uint32_t wit = 0;
            const uint32_t slot_mask = (uint32_t)p.n_slots - 1u, slot_shift = (uint32_t)p.slots_log2;
            auto feed = [&](int n, int s, int conv) {
                const PairSegP& sg = p.seg[s];
                const uint8_t* base = (conv ? sg.w2 : sg.w1) + (size_t)rank * sg.conv_bytes_rank;
                for (int item = 0; item < sg.n_items; item += p.items_per_stage, ++wit) {
                    const int cnt = min(p.items_per_stage, sg.n_items - item);
                    const uint32_t slot = wit & slot_mask, use = wit >> slot_shift;
                    if (use > 0) mbar_wait(bar0 + 8u * (BAR_W_EMPTY + slot), (use - 1) & 1u);
                    if (item == 0) trace(p, 5, n, conv);     // slot free seen = request time of the first stage
                    const uint32_t bytes = (uint32_t)cnt * p.item_bytes;
                    const uint32_t full = bar0 + 8u * (BAR_W_FULL + slot);
                    mbar_arrive_expect_tx(full, bytes);
                    bulk_g2s(sm0 + p.ring_off + slot * p.stage_bytes, base + (size_t)item * p.item_bytes, bytes, full);
                }
            };
            int s2 = 0;
            if (p.resident) {
                // every convolution of the launch fits shared memory at once: one bulk copy each, for the whole kernel
                for (int s = 0; s < n_seg; ++s)
                    for (int conv = 0; conv < 2; ++conv) {
                        const PairSegP& sg = p.seg[s];
                        const uint32_t full = bar0 + 8u * (BAR_W_FULL + 2 * s + conv);
                        mbar_arrive_expect_tx(full, sg.conv_bytes_rank);
                        bulk_g2s(sm0 + p.ring_off + p.wres_off[2 * s + conv], (conv ? sg.w2 : sg.w1) + (size_t)rank * sg.conv_bytes_rank,
                                 sg.conv_bytes_rank, full);
                    }
            } else if (ND == 2) {
                int s1 = 0;
                feed(0, 0, 0);
                for (int n = 0; n < n_units; ++n) {
                    if (++s1 == n_seg) s1 = 0;
                    if (n + 1 < n_units) feed(n + 1, s1, 0);
                    feed(n, s2, 1);
                    if (++s2 == n_seg) s2 = 0;
                }
            } else {
                for (int n = 0; n < n_units; ++n) {
                    feed(n, s2, 0); feed(n, s2, 1);
                    if (++s2 == n_seg) s2 = 0;
                }
            }
        }
    } else if (warp >= PAIR_WARP_MMA0 && warp < PAIR_WARP_MMA0 + NI && rank == 0) {
        // ================= MMA issuers (whole warp loops, one elected lane issues) =================
        // The NI issuer warps walk the same schedule in lock step and split the row blocks of every weight item:
        // a single thread cannot issue the small-N MMAs as fast as the tensor pipe retires them.
        // ND == 2: skewed schedule  c1(0), { c1(n+1), c2(n) }  -- e1 of unit n runs under conv1 of unit n+1.
        // In-order execution of one thread's MMAs gives the implicit orderings: conv1(n+ND) is issued after the
        // T_FULL wait of conv2(n) (D1 free), and e1(n+ND) starts after conv1(n+ND), issued after conv2(n) (t free).
        const int issuer = warp - PAIR_WARP_MMA0;
        if (elect_one()) {
        const uint32_t idesc = umma_idesc_bf16(128u * CG, (uint32_t)C);
        const uint32_t ones_lo = (((sm0 + p.ones_off) >> 4) & 0x3FFFu) | (128u << 16);   // [2 chunks][128 rows][8], LBO = 2 KB
        const uint32_t d1 = tmem_base, d2 = tmem_base + (uint32_t)(ND * NB * C);
        uint32_t wit = 0, w_seen = 0;
        long long c_xa = 0, c_w = 0, c_t = 0, c_d2 = 0;
        const bool prof = PAIR_PROF && p.prof != nullptr;
        const long long c_start = prof_clock();
        const int xa_bufs = p.xa_bufs;                                   // 1 .. 4
        // (n, s): unit and its segment; no divisions on this path
        auto conv1 = [&](int n, int s) {
            const PairSegP& sg = p.seg[s];
            const int use = n / xa_bufs, buf = n - use * xa_bufs;
            long long c0 = 0;
            if (prof) c0 = prof_clock();
            if (prof && issuer == 0) trace(p, 8, n, 0);
            mbar_wait_cluster(bar0 + 8u * (BAR_XA_FULL + buf), (uint32_t)(use & 1));   // both CTAs' tiles
            if (prof && issuer == 0) trace(p, 8, n, 1);
            if (prof) { c_xa += prof_clock() - c0; if (issuer == 0) trace(p, 1, n, 0); }
            tc_fence_after();
            issue_conv<CG, C, NB, KS, NI>(p, prof, issuer, bar0, sm0 + p.xa_off[buf], (uint32_t)sg.xa_rows, sg.n_items, sg.dil, false, false,
                           d1 + (uint32_t)((n % ND) * NB * C), sm0 + p.ring_off, idesc, ones_lo, wit, c_w, sg.k, 0u, 0u, c_t, 2 * s, w_seen, issuer == 0 ? n : -1);
            mma_commit<CG>(bar0 + 8u * (BAR_XA_EMPTY + buf));            // input tile reusable
            mma_commit<CG>(bar0 + 8u * (BAR_D1_FULL + n % ND));          // e1 may start
            if (prof && issuer == 0) trace(p, 1, n, 1);
        };
        auto conv2 = [&](int n, int it, int s) {
            const PairSegP& sg = p.seg[s];
            // D2[it % ND] drained by the e2 of tile it - ND (the operand buffer is awaited per channel block inside)
            long long c0 = 0;
            if (prof) c0 = prof_clock();
            // rinit: D2 arrives INITIALISED with the residual by the e2 warps (every tile, the first ones too)
            if (RINIT) mbar_wait_cluster(bar0 + 8u * (BAR_D2_EMPTY + it % ND), (uint32_t)((it / ND) & 1));
            else if (s == 0 && it >= ND) mbar_wait_cluster(bar0 + 8u * (BAR_D2_EMPTY + it % ND), (uint32_t)((it / ND - 1) & 1));
            if (prof) { c_d2 += prof_clock() - c0; if (issuer == 0) trace(p, 2, n, 0); }
            tc_fence_after();
            issue_conv<CG, C, NB, KS, NI>(p, prof, issuer, bar0, sm0 + p.t_off + (uint32_t)(n % ND) * p.t_bytes, (uint32_t)p.t_rows, sg.n_items, 1,
                           true, s > 0 || RINIT, d2 + (uint32_t)((it % ND) * NB * C), sm0 + p.ring_off, idesc, ones_lo, wit, c_w, sg.k,
                           bar0 + 8u * (BAR_T_FULL + 4 * (n % ND)), (uint32_t)((n / ND) & 1), c_t, 2 * s + 1, w_seen);
            if (s == n_seg - 1) mma_commit<CG>(bar0 + 8u * (BAR_D2_FULL + it % ND));
            if (prof && issuer == 0) trace(p, 2, n, 1);
        };
        // Split issue (p.split: weights resident, C <= 64): issuer warp 0 runs every conv1, issuer warp 1 every conv2,
        // each over ALL row blocks and on its own schedule.  One thread issuing both convolutions was the bottleneck of
        // the small-k launches: its waits (~230 cycles per mbarrier wait even when the phase is long complete) and MMA
        // bursts are serial, 5100 cycles per k = 3 tile at C = 64 for 1660 cycles of tensor-pipe work (event trace).
        // What in-order execution of one thread's MMAs gave for free is explicit here: D1[n % ND] is free once e1 of
        // unit n - ND has read it (BAR_D1_EMPTY), the conv2 operand buffer once conv2 of unit n - ND has retired
        // (BAR_T_EMPTY, awaited by e1).
        auto conv1s = [&](int n, int s) {
            const PairSegP& sg = p.seg[s];
            const int use = n / xa_bufs, buf = n - use * xa_bufs;
            long long c0 = 0;
            if (prof) { c0 = prof_clock(); trace(p, 8, n, 0); }
            mbar_wait_cluster(bar0 + 8u * (BAR_XA_FULL + buf), (uint32_t)(use & 1));
            if (n >= ND) mbar_wait_cluster(bar0 + 8u * (BAR_D1_EMPTY + n % ND), (uint32_t)((n / ND - 1) & 1));
            if (prof) { c_xa += prof_clock() - c0; trace(p, 1, n, 0); }
            tc_fence_after();
            issue_conv<CG, C, NB, KS, 1>(p, prof, 0, bar0, sm0 + p.xa_off[buf], (uint32_t)sg.xa_rows, sg.n_items, sg.dil, false, false,
                           d1 + (uint32_t)((n % ND) * NB * C), sm0 + p.ring_off, idesc, ones_lo, wit, c_w, sg.k, 0u, 0u, c_t, 2 * s, w_seen, n);
            mma_commit<CG>(bar0 + 8u * (BAR_XA_EMPTY + buf));
            mma_commit<CG>(bar0 + 8u * (BAR_D1_FULL + n % ND));
            if (prof) trace(p, 1, n, 1);
        };
        auto conv2s = [&](int n, int it, int s) {
            const PairSegP& sg = p.seg[s];
            long long c0 = 0;
            if (prof) c0 = prof_clock();
            if (RINIT) mbar_wait_cluster(bar0 + 8u * (BAR_D2_EMPTY + it % ND), (uint32_t)((it / ND) & 1));
            else if (s == 0 && it >= ND) mbar_wait_cluster(bar0 + 8u * (BAR_D2_EMPTY + it % ND), (uint32_t)((it / ND - 1) & 1));
            if (prof) { c_d2 += prof_clock() - c0; trace(p, 2, n, 0); }
            tc_fence_after();
            issue_conv<CG, C, NB, KS, 1>(p, prof, 0, bar0, sm0 + p.t_off + (uint32_t)(n % ND) * p.t_bytes, (uint32_t)p.t_rows, sg.n_items, 1,
                           true, s > 0 || RINIT, d2 + (uint32_t)((it % ND) * NB * C), sm0 + p.ring_off, idesc, ones_lo, wit, c_w, sg.k,
                           bar0 + 8u * (BAR_T_FULL + 4 * (n % ND)), (uint32_t)((n / ND) & 1), c_t, 2 * s + 1, w_seen);
            mma_commit<CG>(bar0 + 8u * (BAR_T_EMPTY + n % ND));          // operand buffer reusable once these MMAs retire
            if (s == n_seg - 1) mma_commit<CG>(bar0 + 8u * (BAR_D2_FULL + it % ND));
            if (prof) trace(p, 2, n, 1);
        };
        int it2 = 0, s2 = 0;                          // (tile, segment) of the unit conv2 works on
        if (p.split) {
            if (issuer == 0) {
                for (int n = 0; n < n_units; ++n) {
                    conv1s(n, s2);
                    if (++s2 == n_seg) s2 = 0;
                }
            } else {
                for (int n = 0; n < n_units; ++n) {
                    conv2s(n, it2, s2);
                    if (++s2 == n_seg) { s2 = 0; ++it2; }
                }
            }
        } else if (ND == 2) {
            int s1 = 0;                               // segment of the unit conv1 works on (one unit ahead)
            conv1(0, 0);
            for (int n = 0; n < n_units; ++n) {
                if (++s1 == n_seg) s1 = 0;
                if (n + 1 < n_units) conv1(n + 1, s1);
                conv2(n, it2, s2);
                if (++s2 == n_seg) { s2 = 0; ++it2; }
            }
        } else {
            for (int n = 0; n < n_units; ++n) {
                conv1(n, s2);
                conv2(n, it2, s2);
                if (++s2 == n_seg) { s2 = 0; ++it2; }
            }
        }
        if (PAIR_PROF && p.prof && (issuer == 0 || p.split)) {
            unsigned long long* o = p.prof + (size_t)blockIdx.x * 16;
            if (issuer == 0) {
                o[0] = (unsigned long long)(prof_clock() - c_start);
                o[1] = (unsigned long long)c_xa; o[2] = (unsigned long long)c_w;
            }
            if (issuer == 1 || !p.split) { o[3] = (unsigned long long)c_t; o[4] = (unsigned long long)c_d2; }
        }
        }   // elected thread
        __syncwarp();
    } else if (warp == PAIR_WARP_MMA0 && rank != 0) {
        if (lane == 0) {
            // ================= peer relay: forward "landed" events to the leader's barriers =================
            uint32_t wit = 0;
            const int xa_bufs = p.xa_bufs;
            const uint32_t slot_mask = (uint32_t)p.n_slots - 1u, slot_shift = (uint32_t)p.slots_log2;
            auto relay_xa = [&](int n) {
                const int use = n / xa_bufs, buf = n - use * xa_bufs;
                mbar_wait(bar0 + 8u * (BAR_XA_FULL + buf), (uint32_t)(use & 1));
                mbar_arrive_remote(mapa_u32(bar0 + 8u * (BAR_XA_FULL + buf), 0));
            };
            if (p.resident)
                for (int i = 0; i < 2 * n_seg; ++i) {
                    mbar_wait(bar0 + 8u * (BAR_W_FULL + i), 0u);
                    mbar_arrive_remote(mapa_u32(bar0 + 8u * (BAR_W_FULL + i), 0));
                }
            auto relay_w = [&](int s) {       // the ring stages of one convolution of segment s
                const int stages = p.resident ? 0 : p.seg[s].n_stages;
                for (int q = 0; q < stages; ++q, ++wit) {
                    const uint32_t slot = wit & slot_mask, par = (wit >> slot_shift) & 1u;
                    mbar_wait(bar0 + 8u * (BAR_W_FULL + slot), par);
                    mbar_arrive_remote(mapa_u32(bar0 + 8u * (BAR_W_FULL + slot), 0));
                }
            };
            int s2 = 0;
            if (ND == 2) {
                int s1 = 0;
                relay_xa(0); relay_w(0);
                for (int n = 0; n < n_units; ++n) {
                    if (++s1 == n_seg) s1 = 0;
                    if (n + 1 < n_units) { relay_xa(n + 1); relay_w(s1); }
                    relay_w(s2);
                    if (++s2 == n_seg) s2 = 0;
                }
            } else {
                for (int n = 0; n < n_units; ++n) {
                    relay_xa(n); relay_w(s2); relay_w(s2);
                    if (++s2 == n_seg) s2 = 0;
                }
            }
        }
    }
    else if (warp >= PAIR_E1_WARP0 && warp < PAIR_E1_WARP0 + PAIR_E1_WARPS) {
        // ================= e1: conv1 accumulators -> bf16 conv2 operand in shared memory =================
        // On the critical path of every segment (conv2 cannot start before it ends): fully unrolled, two TMEM
        // loads in flight per warp.
        constexpr int UNITS = GROUPS * NB;               // (32-column group, row block) units per warp, two in flight
        constexpr int EU = ND == 2 ? 1 : UNITS / 2;      // flat where e1 is on the critical path of every segment (ND = 1)
        const int q4 = warp & 3;                         // TMEM lane quarter this warp may read
        const int q = q4 * 32 + lane;                    // row inside a 128-row block
        const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const uint32_t t_full_addr = CG == 2 ? mapa_u32(bar0 + 8u * BAR_T_FULL, 0) : bar0 + 8u * BAR_T_FULL;
        const uint32_t d1_empty_addr = CG == 2 ? mapa_u32(bar0 + 8u * BAR_D1_EMPTY, 0) : bar0 + 8u * BAR_D1_EMPTY;
        constexpr int UNITS_PER_KB = (C < 64 ? 1 : 2) * NB;   // 32-column groups per weight channel block x row blocks
        const __nv_bfloat162 slope2 = __floats2bfloat162_rn(0.1f, 0.1f);
        const uint32_t t_row0 = sm0 + p.t_off + (uint32_t)q * 16u;
        // convert 32 accumulator columns of one row block: bf16, leaky-relu, 4 x 16-byte stores
        auto convert = [&](const uint32_t (&r)[32], uint32_t t_row, int g, int blk, bool inside) {
#pragma unroll
            for (int h = 0; h < 4; ++h) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(r[h * 8 + 0]), __uint_as_float(r[h * 8 + 1]));
                o.y = pack_bf16x2(__uint_as_float(r[h * 8 + 2]), __uint_as_float(r[h * 8 + 3]));
                o.z = pack_bf16x2(__uint_as_float(r[h * 8 + 4]), __uint_as_float(r[h * 8 + 5]));
                o.w = pack_bf16x2(__uint_as_float(r[h * 8 + 6]), __uint_as_float(r[h * 8 + 7]));
                o.x = lrelu_bf16x2(o.x, slope2); o.y = lrelu_bf16x2(o.y, slope2);
                o.z = lrelu_bf16x2(o.z, slope2); o.w = lrelu_bf16x2(o.w, slope2);
                // rows outside the utterance are conv2's zero padding: a predicated second store instead of four selects
                const uint32_t addr = t_row + (uint32_t)(((g * 4 + h) * NB * 128 + blk * 128) * 16);
                if (inside) st_shared_v4(addr, o);
                else st_shared_v4(addr, make_uint4(0u, 0u, 0u, 0u));
            }
        };
        int n = 0;
        long long c_wait = 0, c_work = 0, c_fence = 0, c_ld = 0;
        for (int it = 0; it < p.iters; ++it) {
            const TileCoord t = tile_of(p, it);
            for (int s = 0; s < n_seg; ++s, ++n) {
                const int t0 = t.l0 - p.seg[s].h2 + q;   // sample index of this thread's operand row in block 0
                const long long c0 = prof_clock();
                const int nd = n % ND;                   // accumulator / operand buffer of this unit
                mbar_wait(bar0 + 8u * (BAR_D1_FULL + nd), (uint32_t)((n / ND) & 1));
                // split issue: the operand buffer of unit n - ND must have been read by its conv2 before it is rewritten
                if (p.split && n >= ND) mbar_wait(bar0 + 8u * (BAR_T_EMPTY + nd), (uint32_t)((n / ND - 1) & 1));
                const long long c1 = prof_clock();
                c_wait += c1 - c0;
                if (warp == PAIR_E1_WARP0) trace(p, 3, n, 0);
                tc_fence_after();
                const uint32_t t_acc = t_lane + (uint32_t)(nd * NB * C);
                const uint32_t t_row = t_row0 + (uint32_t)nd * p.t_bytes;
#pragma unroll EU
                for (int u = 0; u < UNITS; u += 2) {             // g-major unit order; both loads precede the wait
                    const int g0 = u / NB, b0 = u % NB, g1 = (u + 1) / NB, b1 = (u + 1) % NB;   // not unrolled: compact code (see e2)
                    uint32_t ra[32], rb[32];
                    const long long cl0 = prof_clock();
                    tmem_ld32(t_acc + (uint32_t)(b0 * C + g0 * 32), ra);
                    tmem_ld32(t_acc + (uint32_t)(b1 * C + g1 * 32), rb);
                    tmem_ld_wait();
                    c_ld += prof_clock() - cl0;
                    const int r0 = t0 + b0 * 128, r1 = t0 + b1 * 128;      // conv2 zero-pads outside the utterance
                    if (!(p.dbg & 2)) {
                        convert(ra, t_row, g0, b0, r0 >= 0 && r0 < p.L);
                        convert(rb, t_row, g1, b1, r1 >= 0 && r1 < p.L);
                    }
                    if ((u + 2) % UNITS_PER_KB == 0) {
                        // channel block (u + 1) / UNITS_PER_KB of the operand is complete for this warp's rows
                        fence_proxy_async();             // generic-proxy stores -> visible to the tensor core (async proxy)
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            const uint32_t bar = t_full_addr + 8u * (uint32_t)(4 * nd + (u + 1) / UNITS_PER_KB);
                            if (CG == 2) mbar_arrive_remote(bar);
                            else mbar_arrive(bar);
                        }
                    }
                }
                if (p.split) {                      // D1[nd] has been read: conv1 of unit n + ND may overwrite it
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        const uint32_t bar = d1_empty_addr + 8u * (uint32_t)nd;
                        if (CG == 2) mbar_arrive_remote(bar);
                        else mbar_arrive(bar);
                    }
                }
                const long long c2 = prof_clock();
                if (warp == PAIR_E1_WARP0) trace(p, 3, n, 1);
                c_work += prof_clock() - c1;
                c_fence += prof_clock() - c2;
            }
        }
        if (PAIR_PROF && p.prof && warp == PAIR_E1_WARP0 && lane == 0) {
            p.prof[(size_t)blockIdx.x * 16 + 5] = (unsigned long long)c_wait;
            p.prof[(size_t)blockIdx.x * 16 + 6] = (unsigned long long)c_work;
            p.prof[(size_t)blockIdx.x * 16 + 9] = (unsigned long long)c_ld;
        }
    } else if (warp < PAIR_E2_WARPS) {
        // ================= e2: conv2 accumulators + residuals -> global =================
        // Runs behind the MMAs of the next tile.  Two warps per lane quarter split the (block, 32-column group) items.
        // The loops over items / 16-column chunks are NOT unrolled: with every loop flat this role alone was ~13 KB of
        // straight-line code per tile, the five roles of a CTA run different code on the same four schedulers, and the
        // ncu source page showed a third of e2's stall samples as instruction fetch (stall_no_inst) -- in launches
        // where e2 is the busiest role (C <= 64).  Addresses are computed per chunk instead.
        constexpr int ITEMS = NB * GROUPS / 2;           // per warp
        constexpr int JU = ITEMS > 2 ? 2 * ITEMS : 1;    // generic drain: flat where e2 is on the critical path between tiles (ND = 1)
        const int q4 = warp & 3;
        const int q = q4 * 32 + lane;
        const int half = warp >> 2;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const uint32_t d2_empty_addr = CG == 2 ? mapa_u32(bar0 + 8u * BAR_D2_EMPTY, 0) : bar0 + 8u * BAR_D2_EMPTY;
        const size_t e_step = (size_t)p.Lp * 8;
        const float s_pos = p.scale, s_neg = p.scale * p.out_slope;
        const int back = 2 * p.h2max;
        float* const part = reinterpret_cast<float*>(smem + p.part_off);   // POST: [2][PAIR_POST_K][PAIR_POST_ROWS] partial sums
        long long c_wait = 0, c_work = 0;
        // Generic drain (MRF launches, conv_post on board), in 16-column half-items: cc = half (cc & 1) of item cc >> 1.  The
        // residual rows of the segments of ONE half-item live in R and are fetched one use ahead -- the registers are
        // refilled with the next half-item's rows (the next tile's first after the last) right after their last use, so
        // the L2 round trip runs under the rest of the half-item.  With the loads issued at the top of a tile e2 waited
        // ~1000 cycles per tile for them (ncu source page).
        uint4 R[PAIR_MAX_SEG][2];                        // [segment][chunk] of the current half-item
        auto item_geom = [&](const TileCoord& t, int j, int& blk, int& g, int& rel, bool& ok) {
            const int i = half + 2 * j;
            blk = i / GROUPS;
            g = i % GROUPS;
            rel = blk == 0 ? 0 : blk * 128 - back;
            ok = t.live && (t.l0 + rel + q) < p.L && (blk > 0 || q < 128 - back) && !(p.dbg & 1);
        };
        auto load_res = [&](const TileCoord& t, int cc) {   // half-item cc, all segments, into R[s]
            int blk, g, rel; bool ok;
            item_geom(t, cc >> 1, blk, g, rel, ok);
            if (!ok) return;
            const size_t e = ((size_t)t.b * CHUNKS * p.Lp + VO2W_PAD_ROWS + t.l0 + q) * 8 + (size_t)rel * 8 + (size_t)(g * 4 + (cc & 1) * 2) * e_step;
#pragma unroll
            for (int s = 0; s < PAIR_MAX_SEG; ++s)
                if (s < n_seg) {
                    R[s][0] = __ldg((const uint4*)(p.seg[s].xa + e));
                    R[s][1] = __ldg((const uint4*)(p.seg[s].xa + e + e_step));
                }
        };
        // Residual by initialisation (p.rinit: single-segment launches).  y = x + conv2(...): instead of fetching x again
        // from HBM after the convolutions (the re-read mostly MISSED L2 -- ncu: 23 % read hit rate, 505 MB read from DRAM
        // for a 361 MB tensor -- and with the stores it cost 150 of the 316 us of a k = 3 launch), the e2 warps write x
        // into D2 BEFORE conv2 accumulates onto it: x is recovered from the staged input tile in shared memory, which holds
        // exactly the stream rows the residual needs, and goes to TMEM with tcgen05.st.  Each warp re-initialises the
        // (row block, column group) items it has just drained, so there is no cross-warp hazard on D2; the input tile is
        // released (XA_EMPTY) by conv1's commit plus these warps' arrivals.
        // Both directions walk this warp's columns in 16-column chunks: chunk cc = half (cc & 1) of item cc >> 1.
        const uint32_t xa_empty_local = bar0 + 8u * BAR_XA_EMPTY;
        auto init_d2 = [&](int u) {
            const int use = u / p.xa_bufs, buf = u - use * p.xa_bufs, ndu = u % ND;
            mbar_wait_cluster(bar0 + 8u * (BAR_XA_FULL + buf), (uint32_t)(use & 1));
            const PairSegP& sg = p.seg[0];
            const uint32_t xa_row0 = sm0 + p.xa_off[buf] + (uint32_t)(q + sg.h1 + sg.h2) * 16u;
            const uint32_t d2u = t_lane + (uint32_t)((ND + ndu) * NB * C);
#pragma unroll 1
            for (int cc = 0; cc < 2 * ITEMS; ++cc) {
                const int i = half + 2 * (cc >> 1), blk = i / GROUPS, hc = (i % GROUPS) * 4 + (cc & 1) * 2;
                const int rel = blk == 0 ? 0 : blk * 128 - back;
                uint32_t r[16];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const uint4 w4 = ld_shared_v4(xa_row0 + (uint32_t)(((hc + h) * sg.xa_rows + rel) * 16));
                    const uint32_t w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                    for (int k2 = 0; k2 < 4; ++k2) {      // x = s >= 0 ? s : 10 s, exact in fp32
                        const float2 f = unpack_bf16x2(w[k2]);
                        float x0, x1;
                        fma2(x0, x1, fminf(f.x, 0.f), fminf(f.y, 0.f), 9.f, f.x, f.y);
                        r[h * 8 + 2 * k2] = __float_as_uint(x0);
                        r[h * 8 + 2 * k2 + 1] = __float_as_uint(x1);
                    }
                }
                tmem_st16(d2u + (uint32_t)(blk * C + hc * 8), r);
            }
            tmem_st_wait();
            if (warp == 0 && u >= ND) trace(p, 9, u - ND, 1);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(xa_empty_local + 8u * (uint32_t)buf);
                if (CG == 2) mbar_arrive_remote(d2_empty_addr + 8u * (uint32_t)ndu);
                else mbar_arrive(d2_empty_addr + 8u * (uint32_t)ndu);
            }
        };
        if constexpr (RINIT) {
            for (int u = 0; u < ND && u < p.iters; ++u) init_d2(u);
        } else if (p.iters > 0) {
            griddep_wait();                              // the residual rows are read straight from the previous launch's output
            load_res(tile_of(p, 0), 0);
        }
        for (int it = 0; it < p.iters; ++it) {
            if (warp == 0) trace(p, 7, it * n_seg + n_seg - 1, 0);
            const TileCoord t = tile_of(p, it);
            const bool has_next = it + 1 < p.iters;
            const size_t e_tile = ((size_t)t.b * CHUNKS * p.Lp + VO2W_PAD_ROWS + t.l0 + q) * 8;   // chunk 0, block 0 row of this thread
            const long long c0 = prof_clock();
            const int nd = it % ND;
            mbar_wait(bar0 + 8u * (BAR_D2_FULL + nd), (uint32_t)((it / ND) & 1));
            const long long c1 = prof_clock();
            c_wait += c1 - c0;
            if (warp == 0) trace(p, 4, it * n_seg + n_seg - 1, 0);
            tc_fence_after();
            const uint32_t d2n = t_lane + (uint32_t)((ND + nd) * NB * C);
            if constexpr (RINIT) {
                // residual already inside D2: drain = activation, bf16, store.  The load of chunk cc + 1 is in flight
                // while chunk cc is converted and stored.
                auto chunk_geom = [&](int cc, uint32_t& col, size_t& e, bool& ok) {
                    const int i = half + 2 * (cc >> 1), blk = i / GROUPS, hc = (i % GROUPS) * 4 + (cc & 1) * 2;
                    const int rel = blk == 0 ? 0 : blk * 128 - back;
                    col = (uint32_t)(blk * C + hc * 8);
                    e = e_tile + (size_t)rel * 8 + (size_t)hc * e_step;
                    ok = t.live && (t.l0 + rel + q) < p.L && (blk > 0 || q < 128 - back) && !(p.dbg & 1);
                };
                auto store_chunk = [&](const uint32_t (&r)[16], size_t e) {
                    float v[16];
#pragma unroll
                    for (int c = 0; c < 16; c += 2) {      // scale = 1 in these launches (pair_launch): max(y, y * slope)
                        const float y0 = __uint_as_float(r[c]), y1 = __uint_as_float(r[c + 1]);
                        float n0, n1;
                        mul2(n0, n1, y0, y1, s_neg);
                        v[c] = fmaxf(y0, n0);
                        v[c + 1] = fmaxf(y1, n1);
                    }
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint4 o;
                        o.x = pack_bf16x2(v[h * 8 + 0], v[h * 8 + 1]); o.y = pack_bf16x2(v[h * 8 + 2], v[h * 8 + 3]);
                        o.z = pack_bf16x2(v[h * 8 + 4], v[h * 8 + 5]); o.w = pack_bf16x2(v[h * 8 + 6], v[h * 8 + 7]);
                        *(uint4*)(p.out_act + e + (size_t)h * e_step) = o;
                    }
                };
                uint32_t ra[16], rb[16];
                uint32_t col;
                size_t ea, eb;
                bool oka, okb;
                chunk_geom(0, col, ea, oka);
                tmem_ld16(d2n + col, ra);
#pragma unroll 1
                for (int cc = 0; cc < 2 * ITEMS; cc += 2) {
                    chunk_geom(cc + 1, col, eb, okb);
                    tmem_ld_wait();
                    if (warp == 0 && cc == 0) trace(p, 7, it, 1);
                    tmem_ld16(d2n + col, rb);
                    if (oka) store_chunk(ra, ea);
                    tmem_ld_wait();
                    if (cc + 2 < 2 * ITEMS) {
                        chunk_geom(cc + 2, col, ea, oka);
                        tmem_ld16(d2n + col, ra);
                    }
                    if (okb) store_chunk(rb, eb);
                }
                // this warp's part of D2[nd] is drained: initialise it for tile it + ND (init_d2 signals the issuer)
                if (warp == 0) trace(p, 9, it, 0);
                if (it + ND < p.iters) init_d2(it + ND);
            } else {
                const TileCoord tn = tile_of(p, has_next ? it + 1 : it);
#pragma unroll JU
                for (int cc = 0; cc < 2 * ITEMS; ++cc) {
                    // 16 live accumulator registers next to the 24 of R: with a 32-column load and a whole item's residuals
                    // in registers this path spilled, and the spills were reloaded in every loop of the kernel
                    int blk, g, rel; bool ok;
                    item_geom(t, cc >> 1, blk, g, rel, ok);
                    const int hh = cc & 1;
                    uint32_t r[16];
                    tmem_ld16(d2n + (uint32_t)(blk * C + g * 32 + hh * 16), r);
                    tmem_ld_wait();
                    float v[16];
#pragma unroll
                    for (int c = 0; c < 16; ++c) v[c] = __uint_as_float(r[c]);
                    // the stream holds s = bf16(leaky_relu(x, 0.1)); the residual is x = s >= 0 ? s : 10 s  (exact in fp32)
                    auto add_res = [&](const uint4& x0, const uint4& x1) {
                        const uint32_t w[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float2 f = unpack_bf16x2(w[i]);
                            float t0, t1;
                            add2(t0, t1, v[2 * i], v[2 * i + 1], f.x, f.y);
                            fma2(v[2 * i], v[2 * i + 1], fminf(f.x, 0.f), fminf(f.y, 0.f), 9.f, t0, t1);
                        }
                    };
                    if (ok) {
#pragma unroll
                        for (int s = 0; s < PAIR_MAX_SEG; ++s)
                            if (s < n_seg) add_res(R[s][0], R[s][1]);
                    }
                    // the registers are free again: fetch what uses them next while this half-item is scaled and stored
                    if (cc + 1 < 2 * ITEMS) load_res(t, cc + 1);
                    else if (has_next) load_res(tn, 0);
                    if (ok) {
                        // leaky_relu(v * scale, out_slope); max(a, a * slope) for slope < 1
#pragma unroll
                        for (int c = 0; c < 16; c += 2) {
                            float a0, a1, n0, n1;
                            mul2(a0, a1, v[c], v[c + 1], s_pos);
                            mul2(n0, n1, v[c], v[c + 1], s_neg);
                            v[c] = fmaxf(a0, n0);
                            v[c + 1] = fmaxf(a1, n1);
                        }
                    }
                    if constexpr (POST) {
                        // conv_post + tanh (models.py:161-163) ride this launch: the row is leaky_relu(xs / num_kernels, 0.01)
                        // in fp32 registers, so each thread reduces its row against the 7 taps (weights are kernel
                        // parameters: constant-bank operands of the FFMAs; the 7 sums advance together, so no FFMA waits
                        // for the one before it) and leaves 7 partial sums in shared memory (the second half adds to the
                        // first's); the taps are summed across rows after the tile's barrier.  Rows outside the utterance
                        // are conv_post's zero padding.  x is never written to HBM.
                        if (blk > 0 || q < 128 - back) {
                            float* pr = part + (it & 1) * (PAIR_POST_K * PAIR_POST_ROWS) + rel + q;
                            float a[PAIR_POST_K];
#pragma unroll
                            for (int jj = 0; jj < PAIR_POST_K; ++jj) a[jj] = hh == 0 ? 0.f : pr[jj * PAIR_POST_ROWS];
                            if (ok) {
                                const float* pw = p.post_w + hh * 16;
#pragma unroll
                                for (int c = 0; c < 16; ++c) {
#pragma unroll
                                    for (int jj = 0; jj < PAIR_POST_K; ++jj) a[jj] = fmaf(v[c], pw[jj * 32 + c], a[jj]);
                                }
                            }
#pragma unroll
                            for (int jj = 0; jj < PAIR_POST_K; ++jj) pr[jj * PAIR_POST_ROWS] = a[jj];
                        }
                    } else if (ok) {
                        const size_t e = e_tile + (size_t)rel * 8 + (size_t)(g * 4 + hh * 2) * e_step;
                        // out = bf16(...), one rounding
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            uint4 o;
                            o.x = pack_bf16x2(v[h * 8 + 0], v[h * 8 + 1]); o.y = pack_bf16x2(v[h * 8 + 2], v[h * 8 + 3]);
                            o.z = pack_bf16x2(v[h * 8 + 4], v[h * 8 + 5]); o.w = pack_bf16x2(v[h * 8 + 6], v[h * 8 + 7]);
                            *(uint4*)(p.out_act + e + (size_t)h * e_step) = o;
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (CG == 2) mbar_arrive_remote(d2_empty_addr + 8u * (uint32_t)nd);
                    else mbar_arrive(d2_empty_addr + 8u * (uint32_t)nd);
                }
            }
            if constexpr (POST) {
                // all 8 e2 warps have left their partial sums (the buffers alternate per tile, so one barrier per tile
                // is enough: nobody can write this buffer again before everybody has passed the next tile's barrier)
                asm volatile("bar.sync 1, 256;" ::: "memory");
                const float* pb = part + (it & 1) * (PAIR_POST_K * PAIR_POST_ROWS);
                const int d0 = t.o0 - t.l0;                              // 3, or 0 for the first tile of an utterance
                for (int lr = (int)threadIdx.x; lr < p.Vs; lr += 32 * PAIR_E2_WARPS) {
                    const int l = t.o0 + lr;
                    if (t.live && l < p.L) {
                        float a = p.post_b;
#pragma unroll
                        for (int jj = 0; jj < PAIR_POST_K; ++jj) {
                            const int rr = d0 + lr + jj - PAIR_POST_HALO;
                            if (rr >= 0) a += pb[jj * PAIR_POST_ROWS + rr];
                        }
                        const float y = tanhf(a);
                        const size_t o = (size_t)t.b * p.L + l;
                        if (p.wav) p.wav[o] = y;
                        if (p.wav_i16) p.wav_i16[o] = (short)(int)(y * p.i16_scale);   // numpy astype("int16") truncates toward zero
                    }
                }
            }
            c_work += prof_clock() - c1;
            if (warp == 0) trace(p, 4, it * n_seg + n_seg - 1, 1);
        }
        if (PAIR_PROF && p.prof && warp == 0 && lane == 0) {
            p.prof[(size_t)blockIdx.x * 16 + 7] = (unsigned long long)c_wait;
            p.prof[(size_t)blockIdx.x * 16 + 8] = (unsigned long long)c_work;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();
    if (warp == PAIR_WARP_TMEM) tmem_dealloc_cg<CG>(tmem_base, 512u);
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// host side: weight packing, tile plan, launch
// ---------------------------------------------------------------------------------------------------
void pack_pair_conv_host(std::vector<uint16_t>& out, const float* w, const float* bias, int C, int k, int cg) {
    const int KB = C < 64 ? C : 64, kbs = C / KB, n_loc = C / cg;
    const size_t item = (size_t)n_loc * KB;                      // elements of one item of one rank
    out.assign((size_t)cg * (k * kbs + 1) * item, 0);
    size_t o = 0;
    for (int rank = 0; rank < cg; ++rank) {
        // item 0: the bias as a K=16 operand, columns 0/1 = bf16 hi/lo parts (the A operand has ones there)
        for (int n = 0; n < n_loc; ++n) {
            const float b = bias[rank * n_loc + n];
            const uint16_t hi = f32_to_bf16_rn(b);
            uint32_t hb = (uint32_t)hi << 16;
            float hf;
            memcpy(&hf, &hb, 4);
            out[o + (size_t)n * 8 + 0] = hi;
            out[o + (size_t)n * 8 + 1] = f32_to_bf16_rn(b - hf);
        }
        o += item;
        for (int kb = 0; kb < kbs; ++kb)
            for (int tap = 0; tap < k; ++tap)
                for (int chunk = 0; chunk < KB / 8; ++chunk)
                    for (int n = 0; n < n_loc; ++n)
                        for (int e = 0; e < 8; ++e) {
                            const int co = rank * n_loc + n, ci = kb * KB + chunk * 8 + e;
                            out[o++] = f32_to_bf16_rn(w[((size_t)co * C + ci) * k + tap]);
                        }
    }
}

int pack_pair(PackedPair& P, const float* w1, const float* b1, const float* w2, const float* b2, int C, int k, int dil,
              int cg) {
    VO2W_REQUIRE(C == 32 || C == 64 || C == 128 || C == 256, "pack_pair: channels %d unsupported (32, 64, 128 or 256)", C);
    VO2W_REQUIRE((k & 1) && k >= 1 && k <= 11, "pack_pair: kernel size %d unsupported", k);
    VO2W_REQUIRE(cg == 1 || cg == 2, "pack_pair: cluster size %d", cg);
    VO2W_REQUIRE((dil + 1) * (k - 1) / 2 <= VO2W_PAD_ROWS - 2, "pack_pair: halo exceeds the pad rows (k=%d dil=%d)", k, dil);
    P.C = C; P.k = k; P.dil = dil; P.cg = cg;
    std::vector<uint16_t> h;
    pack_pair_conv_host(h, w1, b1, C, k, cg);
    VO2W_CUDA_OK(cudaMalloc(&P.w1_dev, h.size() * 2));
    VO2W_CUDA_OK(cudaMemcpy(P.w1_dev, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    pack_pair_conv_host(h, w2, b2, C, k, cg);
    VO2W_CUDA_OK(cudaMalloc(&P.w2_dev, h.size() * 2));
    VO2W_CUDA_OK(cudaMemcpy(P.w2_dev, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    return VO2W_OK;
}

void free_pair(PackedPair& P) {
    if (P.w1_dev) cudaFree(P.w1_dev);
    if (P.w2_dev) cudaFree(P.w2_dev);
    P.w1_dev = P.w2_dev = nullptr;
}

int make_group(PackedGroup& G, const PackedPair* const* pairs, int n_seg) {
    VO2W_REQUIRE(n_seg >= 1 && n_seg <= PAIR_MAX_SEG, "make_group: %d segments", n_seg);
    for (int s = 0; s < n_seg; ++s) {
        VO2W_REQUIRE(pairs[s]->C == pairs[0]->C && pairs[s]->cg == pairs[0]->cg, "make_group: segments must share channels and cluster size");
        G.seg[s] = pairs[s];
    }
    G.n_seg = n_seg;
    return VO2W_OK;
}

void free_group(PackedGroup& G) { G.n_seg = 0; }

int pair_plan(PairP& p, int C, int n_seg, const int* ks, const int* dils, int cg, int B, int L, int n_sms, bool post) {
    memset(&p, 0, sizeof(p));
    p.C = C; p.L = L; p.Lp = (int)vo2w_padded_len(L);
    p.n_seg = n_seg;
    p.NB = C >= 256 ? 1 : (C >= 64 ? 2 : 4);
    VO2W_REQUIRE(2 * p.NB * C <= 512, "pair_plan: accumulators exceed TMEM");
    p.KB = C < 64 ? C : 64;
    p.ksteps = p.KB / 16;
    const int kbs = C / p.KB;
    p.kbs_log2 = kbs == 1 ? 0 : (kbs == 2 ? 1 : 2);
    int h1max = 0, kmax = 0;
    for (int s = 0; s < n_seg; ++s) {
        PairSegP& sg = p.seg[s];
        sg.k = ks[s]; sg.dil = dils[s];
        sg.h1 = dils[s] * (ks[s] - 1) / 2;
        sg.h2 = (ks[s] - 1) / 2;
        sg.n_items = ks[s] * kbs + 1;                 // bias item + (tap, channel block) items
        sg.xa_rows = 128 * p.NB + 2 * sg.h1;

        if (sg.h1 > h1max) h1max = sg.h1;
        if (sg.h2 > p.h2max) p.h2max = sg.h2;
        if (ks[s] > kmax) kmax = ks[s];
    }
    p.V = 128 * p.NB - 2 * p.h2max;
    VO2W_REQUIRE(!post || (C == 32 && cg == 2), "pair_plan: conv_post rides the 32-channel stage on CTA pairs only");
    p.post = post ? 1 : 0;
    p.Vs = post ? p.V - 2 * PAIR_POST_HALO : p.V;      // conv_post needs 3 more rows of x on either side
    p.tiles_per_utt = vo2w_ceil_div(L, p.Vs);
    p.n_tiles = B * p.tiles_per_utt;
    p.inv_tiles_per_utt = 1.0f / (float)p.tiles_per_utt;
    p.item_bytes = (uint32_t)(C / cg) * p.KB * 2;
    for (int s = 0; s < n_seg; ++s) p.seg[s].conv_bytes_rank = (uint32_t)p.seg[s].n_items * p.item_bytes;
    p.items_per_stage = (int)(16384u / p.item_bytes);
    if (p.items_per_stage < 1) p.items_per_stage = 1;
    if (p.items_per_stage > kmax * kbs + 1) p.items_per_stage = kmax * kbs + 1;
    p.stage_bytes = (uint32_t)p.items_per_stage * p.item_bytes;
    p.n_slots = p.stage_bytes > 16384u ? 2 : PAIR_MAX_SLOTS;
    p.slots_log2 = p.n_slots == 2 ? 1 : 2;
    for (int s = 0; s < n_seg; ++s) p.seg[s].n_stages = (p.seg[s].n_items + p.items_per_stage - 1) / p.items_per_stage;
    p.t_rows = 128 * p.NB;
    const uint32_t xa_bytes = (uint32_t)(128 * p.NB + 2 * h1max) * C * 2, t_bytes = (uint32_t)p.t_rows * C * 2;
    const uint32_t ones_bytes = 2u * 128u * 16u;
    p.t_bufs = PAIR_ND(C);
    p.t_bytes = t_bytes;
    // Weights resident: where all convolutions of the launch fit next to the tiles (every launch of the 32-channel
    // stage, the k = 3 launches of the 64-channel stage) they are copied once per CTA instead of streaming through the
    // ring for every tile: the ring's per-stage wait / commit / peer relay was ~1000 idle cycles per convolution, half
    // of a k = 3 unit at C = 32.
    uint32_t res_total = 0;
    for (int s = 0; s < n_seg; ++s)
        for (int conv = 0; conv < 2; ++conv) {
            p.wres_off[2 * s + conv] = res_total;
            res_total += (p.seg[s].conv_bytes_rank + 127u) & ~127u;
        }
    p.resident = res_total <= PAIR_RESIDENT_MAX ? 1 : 0;
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_PAIR_RESIDENT") && getenv("VO2W_PAIR_RESIDENT")[0] == '0') p.resident = 0;
#endif
    p.split = (p.resident && p.NB >= 2) ? 1 : 0;       // two issuer warps exist where NB >= 2
    p.rinit = (n_seg == 1 && !post) ? 1 : 0;           // residual by initialisation of D2 (single-segment launches; the
                                                       // conv_post epilogue lives in the generic drain loop)
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_PAIR_RINIT") && getenv("VO2W_PAIR_RINIT")[0] == '0') p.rinit = 0;
#endif
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_PAIR_SPLIT") && getenv("VO2W_PAIR_SPLIT")[0] == '0') p.split = 0;
#endif
    const uint32_t ring_bytes = p.resident ? res_total : (uint32_t)p.n_slots * p.stage_bytes;
    const uint32_t fixed = (uint32_t)p.t_bufs * t_bytes + ring_bytes + ones_bytes;
    // Input tiles in flight: the small-k launches are streams (a k = 3 launch moves 722 MB for 69-277 GFLOP), and what
    // bounds a stream is the bytes each SM has in flight (HBM bandwidth x the latency of the load -> barrier -> relay ->
    // MMA -> commit chain, ~5 us): with 2 tiles of 36 KB the 32-channel stage sat at 2.4 TB/s.  So take every buffer that fits.
    const uint32_t part_bytes = post ? 2u * PAIR_POST_K * PAIR_POST_ROWS * 4u : 0u;
    int max_bufs = 4;
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_PAIR_XABUFS")) max_bufs = atoi(getenv("VO2W_PAIR_XABUFS"));
#endif
    p.xa_bufs = 1;
    for (int nb = 2; nb <= max_bufs; ++nb)
        if (fixed + part_bytes + (uint32_t)nb * xa_bytes <= PAIR_SMEM_LIMIT) p.xa_bufs = nb;
    VO2W_REQUIRE(2 * p.t_bufs * p.NB * C <= 512, "pair_plan: accumulators exceed TMEM");
    uint32_t off = 0;
    for (int i = 0; i < 4; ++i) p.xa_off[i] = 0;
    for (int i = 0; i < p.xa_bufs; ++i) { p.xa_off[i] = off; off += xa_bytes; }
    p.t_off = off; off += (uint32_t)p.t_bufs * t_bytes;
    p.ring_off = off; off += ring_bytes;
    p.ones_off = off; off += ones_bytes;
    if (post) { p.part_off = off; off += 2u * PAIR_POST_K * PAIR_POST_ROWS * 4u; }
    p.smem_bytes = off;
    VO2W_REQUIRE(off <= PAIR_SMEM_LIMIT, "pair_plan: %u bytes of shared memory (C=%d)", off, C);
    int grid = n_sms / cg * cg;
    const int need = (p.n_tiles + cg - 1) / cg * cg;
    if (grid > need) grid = need;
    p.iters = (p.n_tiles + grid - 1) / grid;
    return grid;
}

typedef void (*pair_kernel_t)(const PairP);
template <int CG, bool RINIT>
static pair_kernel_t pair_kernel_c(int C) {
    return C == 256 ? pair_umma_kernel<CG, 256, false, RINIT> : (C == 128 ? pair_umma_kernel<CG, 128, false, RINIT>
         : (C == 64 ? pair_umma_kernel<CG, 64, false, RINIT> : pair_umma_kernel<CG, 32, false, RINIT>));
}
static pair_kernel_t pair_kernel_for(int cg, int C, bool post, bool rinit) {
    if (post) return pair_umma_kernel<2, 32, true, false>;
    if (cg == 2) return rinit ? pair_kernel_c<2, true>(C) : pair_kernel_c<2, false>(C);
    return rinit ? pair_kernel_c<1, true>(C) : pair_kernel_c<1, false>(C);
}

int pair_launch(const PackedGroup& G, const PairIO& io, int B, int L, cudaStream_t st) {
    static bool attr_set[VO2W_MAX_DEV];
    const int cur = vo2w_cur_dev();
    const int n_sms = vo2w_sm_count();
    if (!attr_set[cur]) {
        for (int cgi = 1; cgi <= 2; ++cgi)
            for (int c = 32; c <= 256; c *= 2)
                for (int ri = 0; ri < 2; ++ri)
                    VO2W_CUDA_OK(cudaFuncSetAttribute(pair_kernel_for(cgi, c, false, ri != 0), cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_LIMIT));
        VO2W_CUDA_OK(cudaFuncSetAttribute(pair_kernel_for(2, 32, true, false), cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_LIMIT));
        attr_set[cur] = true;
    }
    const bool post = io.post_w != nullptr;
    VO2W_REQUIRE(G.n_seg >= 1 && (post ? (io.wav || io.wav_i16) : io.out_act != nullptr), "pair_launch: incomplete group or missing output");
    const int cg = G.seg[0]->cg;
    int ks[PAIR_MAX_SEG], dils[PAIR_MAX_SEG];
    for (int s = 0; s < G.n_seg; ++s) { ks[s] = G.seg[s]->k; dils[s] = G.seg[s]->dil; }
    PairP p;
    const int grid = pair_plan(p, G.seg[0]->C, G.n_seg, ks, dils, cg, B, L, n_sms, post);
    if (grid < 0) return grid;
    if (post) {
        memcpy(p.post_w, io.post_w, sizeof(p.post_w));
        p.post_b = io.post_b;
        p.wav = io.wav; p.wav_i16 = io.wav_i16; p.i16_scale = io.i16_scale;
    }
    for (int s = 0; s < G.n_seg; ++s) {
        VO2W_REQUIRE(io.xa[s], "pair_launch: null input of segment %d", s);
        p.seg[s].xa = io.xa[s];
        p.seg[s].w1 = G.seg[s]->w1_dev; p.seg[s].w2 = G.seg[s]->w2_dev;
    }
    p.out_act = io.out_act;
    p.out_slope = io.out_slope; p.scale = io.scale;
    if (io.scale != 1.0f) p.rinit = 0;            // the residual-initialised drain assumes scale = 1 (every ResBlock iteration)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(PAIR_THREADS);
    cfg.dynamicSmemBytes = p.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)cg; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;     // see griddep_wait(); VO2W_PDL=0 turns it off
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = vo2w_pdl_enabled() ? 2 : 1;
    static const bool prof_on = PAIR_PROF && getenv("VO2W_PAIR_PROF") && getenv("VO2W_PAIR_PROF")[0] == '1';   // -DVO2W_PAIR_PROFILE builds only
#ifdef VO2W_EXPERIMENTS
    // what-if timing switches (results are WRONG under VO2W_PAIR_DBG): only in builds made with -DVO2W_EXPERIMENTS, so a
    // stray environment variable can never corrupt the audio of the shipped library
    static const int dbg = getenv("VO2W_PAIR_DBG") ? atoi(getenv("VO2W_PAIR_DBG")) : 0;
    p.dbg = dbg;
    static const int pf = getenv("VO2W_PAIR_PF") ? atoi(getenv("VO2W_PAIR_PF")) : 0;
    p.pf = pf;
#endif
    if (prof_on) {
        VO2W_CUDA_OK(cudaMalloc(&p.prof, ((size_t)grid * 16 + PAIR_TR_ROLES * PAIR_TR_UNITS * 2) * 8));
        VO2W_CUDA_OK(cudaMemsetAsync(p.prof, 0, ((size_t)grid * 16 + PAIR_TR_ROLES * PAIR_TR_UNITS * 2) * 8, st));
    }
    VO2W_CUDA_OK(cudaLaunchKernelEx(&cfg, pair_kernel_for(cg, p.C, post, p.rinit != 0), p));
    if (prof_on) {   // debug aid: per-role cycle accounting, averaged over the CTAs that recorded it
        std::vector<unsigned long long> h((size_t)grid * 16 + PAIR_TR_ROLES * PAIR_TR_UNITS * 2);
        VO2W_CUDA_OK(cudaStreamSynchronize(st));
        VO2W_CUDA_OK(cudaMemcpy(h.data(), p.prof, h.size() * 8, cudaMemcpyDeviceToHost));
        cudaFree(p.prof);
        double a[11] = {0};
        int nl = 0;
        for (int b = 0; b < grid; ++b) {
            if (h[(size_t)b * 16]) ++nl;
            for (int i = 0; i < 11; ++i) a[i] += (double)h[(size_t)b * 16 + i];
        }
        const double dl = nl ? nl : 1, dg = grid;
        fprintf(stderr, "[pair prof] C=%d segs=%d k0=%d NB=%d iters=%d tiles=%d | mma total %.0f: wait xa %.0f w %.0f t %.0f d2e %.0f | "
                        "e1 wait %.0f work %.0f (tmem ld+wait %.0f) | e2 wait %.0f work %.0f (cycles per CTA)\n",
                p.C, p.n_seg, p.seg[0].k, p.NB, p.iters, p.n_tiles, a[0] / dl, a[1] / dl, a[2] / dl, a[3] / dl, a[4] / dl,
                a[5] / dg, a[6] / dg, a[9] / dg, a[7] / dg, a[8] / dg);
        static const bool trace_on = getenv("VO2W_PAIR_TRACE") && getenv("VO2W_PAIR_TRACE")[0] == '1';
        if (trace_on) {   // CTA 0, units 16..27: cycles since conv1 of unit 16 was ready to issue
            const unsigned long long* tr = h.data() + (size_t)grid * 16;
            auto at = [&](int role, int u, int sl) { return (long long)tr[((size_t)role * PAIR_TR_UNITS + u) * 2 + sl]; };
            const long long t0 = at(1, 16, 0);
            static const char* names[PAIR_TR_ROLES] = {"xa producer (idle, EMPTY seen)", "conv1 issue (start, done)", "conv2 issue (start, done)",
                                                       "e1 (D1 full seen, done)", "e2 (D2 full seen, done)", "w producer (request c1, c2)",
                                                       "conv1 (W seen, items issued)", "e2 (loop top, first tmem ld done)",
                                                       "conv1 (before XA wait, XA local seen)", "e2 (drained, D2 init stored)"};
            for (int role = 0; role < PAIR_TR_ROLES; ++role) {
                fprintf(stderr, "  [trace] %-32s", names[role]);
                for (int u = 16; u < 28 && u < p.iters * p.n_seg; ++u)
                    fprintf(stderr, " u%d:%lld,%lld", u, at(role, u, 0) ? at(role, u, 0) - t0 : -1, at(role, u, 1) ? at(role, u, 1) - t0 : -1);
                fprintf(stderr, "\n");
            }
        }
    }
    return VO2W_OK;
}

// ---------------------------------------------------------------------------------------------------
// Known-answer probe: 1..3 ResBlock iterations (summed) through the fused kernel, NCL fp32 in / out.
// ---------------------------------------------------------------------------------------------------
#include "conv_umma.cuh"

namespace {
// bf16 operand buffer [B][C/8][Lp][8] -> fp32 NCL
__global__ void unchunk_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int C, int L, int Lp, float inv_slope) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y, b = blockIdx.z;
    if (l >= L) return;
    const float v = __bfloat162float(x[(((size_t)b * (C >> 3) + (c >> 3)) * Lp + VO2W_PAD_ROWS + l) * 8 + (c & 7)]);
    out[((size_t)b * C + c) * L + l] = v < 0.f ? v * inv_slope : v;
}
}  // namespace

int unchunk_bf16(const __nv_bfloat16* x, float* out, int B, int C, int L, int Lp, cudaStream_t st, float inv_slope) {
    dim3 grid((unsigned)vo2w_ceil_div(L, 128), (unsigned)C, (unsigned)B);
    unchunk_bf16_kernel<<<grid, 128, 0, st>>>(x, out, C, L, Lp, inv_slope);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

static int debug_pair_impl(int n_seg, const float* const* w1_host, const float* const* b1_host,
                           const float* const* w2_host, const float* const* b2_host, int C, const int* ks,
                           const int* dils, int cg, const float* const* x_dev, int B, int L, float scale,
                           float out_slope, float* out_dev, float* act_dev, const float* post_w_host, float post_b,
                           float* wav_dev, int16_t* wav_i16_dev, float i16_scale, void* stream) {
    const bool post = post_w_host != nullptr;
    VO2W_REQUIRE(w1_host && b1_host && w2_host && b2_host && ks && dils && x_dev && (post ? (wav_dev || wav_i16_dev) : act_dev != nullptr),
                 "debug_pair: null argument");
    VO2W_REQUIRE(!post || (C == 32 && cg == 2), "debug_pair_post: conv_post rides the 32-channel CTA-pair kernel only");
    VO2W_REQUIRE(n_seg >= 1 && n_seg <= PAIR_MAX_SEG, "debug_pair: %d segments", n_seg);
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    PackedPair P[PAIR_MAX_SEG];
    PackedGroup G;
    const PackedPair* pp[PAIR_MAX_SEG];
    for (int s = 0; s < n_seg && rc == VO2W_OK; ++s) {
        rc = pack_pair(P[s], w1_host[s], b1_host[s], w2_host[s], b2_host[s], C, ks[s], dils[s], cg);
        pp[s] = &P[s];
    }
    if (rc == VO2W_OK) rc = make_group(G, pp, n_seg);
    const long long Lp = vo2w_padded_len(L);
    const size_t bytes = (size_t)B * C * Lp * 2;
    __nv_bfloat16* buf[PAIR_MAX_SEG + 1] = {};   // the stream of every segment, out_act
    const int n_buf = n_seg + 1;
    for (int i = 0; i < n_buf && rc == VO2W_OK; ++i) {
        if (cudaMalloc(&buf[i], bytes) != cudaSuccess || cudaMemsetAsync(buf[i], 0, bytes, st) != cudaSuccess) {
            vo2w_set_error("debug_pair: allocation failed");
            rc = VO2W_E_CUDA;
        }
    }
    PairIO io;
    for (int s = 0; s < n_seg && rc == VO2W_OK; ++s) {
        rc = pack_ncl_to_operand(x_dev[s], buf[s], B, C, L, (int)Lp, 0.1f, st);
        io.xa[s] = buf[s];
    }
    float pw[PAIR_POST_K * 32];
    if (rc == VO2W_OK) {
        io.out_act = buf[n_seg];
        io.scale = scale;
        io.out_slope = out_slope;
        if (post) {                                   // torch layout [1, 32, 7] -> [tap][channel]
            for (int c = 0; c < 32; ++c)
                for (int j = 0; j < PAIR_POST_K; ++j) pw[j * 32 + c] = post_w_host[c * PAIR_POST_K + j];
            io.out_act = nullptr;
            io.post_w = pw; io.post_b = post_b;
            io.wav = wav_dev; io.wav_i16 = wav_i16_dev; io.i16_scale = i16_scale;
        }
        rc = pair_launch(G, io, B, L, st);
    }
    // out = the value a consumer recovers from the stream (activation undone), act = the stream itself
    if (rc == VO2W_OK && out_dev && !post) rc = unchunk_bf16(buf[n_seg], out_dev, B, C, L, (int)Lp, st, 1.0f / out_slope);
    if (rc == VO2W_OK && !post) rc = unchunk_bf16(buf[n_seg], act_dev, B, C, L, (int)Lp, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (rc == VO2W_OK && e != cudaSuccess) {
        vo2w_set_error("debug_pair: kernel failed: %s", cudaGetErrorString(e));
        rc = VO2W_E_CUDA;
    }
    for (int i = 0; i < n_buf; ++i) if (buf[i]) cudaFree(buf[i]);
    free_group(G);
    for (int s = 0; s < n_seg; ++s) free_pair(P[s]);
    return rc;
}

extern "C" int vo2w_debug_pair(int n_seg, const float* const* w1_host, const float* const* b1_host,
                               const float* const* w2_host, const float* const* b2_host, int C, const int* ks,
                               const int* dils, int cg, const float* const* x_dev, int B, int L, float scale,
                               float out_slope, float* out_dev, float* act_dev, void* stream) {
    VO2W_REQUIRE(act_dev, "debug_pair: null argument");
    return debug_pair_impl(n_seg, w1_host, b1_host, w2_host, b2_host, C, ks, dils, cg, x_dev, B, L, scale, out_slope, out_dev, act_dev,
                           nullptr, 0.f, nullptr, nullptr, 0.f, stream);
}

extern "C" int vo2w_debug_pair_post(int n_seg, const float* const* w1_host, const float* const* b1_host,
                                    const float* const* w2_host, const float* const* b2_host, const int* ks, const int* dils,
                                    const float* const* x_dev, int B, int L, float scale, float out_slope,
                                    const float* post_w_host, float post_b, float* wav_dev, int16_t* wav_i16_dev,
                                    float i16_scale, void* stream) {
    VO2W_REQUIRE(post_w_host, "debug_pair_post: null conv_post weights");
    return debug_pair_impl(n_seg, w1_host, b1_host, w2_host, b2_host, 32, ks, dils, 2, x_dev, B, L, scale, out_slope, nullptr, nullptr,
                           post_w_host, post_b, wav_dev, wav_i16_dev, i16_scale, stream);
}
