// Fused ResBlock conv pair, tcgen05 / TMEM implicit GEMM for sm_100a (the VO2W_PREC_BF16 hot loop).
//
//   r = x + conv2(lrelu(conv1(xa) + b1, 0.1)) + b2,   xa = lrelu(x, 0.1)          (models.py:96-103)
//
// Orientation: time is the UMMA M dimension (TMEM lane = sample), output channels N, input channels K;
// a convolution tap is the SAME shared-memory tile read at a row offset (SWIZZLE_NONE K-major core
// matrices, any 16-byte aligned start address), so dilation halos cost no copies.
//
// One persistent CTA (CG=1) or CTA pair (CG=2, tcgen05 cta_group::2, M=256 = 2 x 128 rows, the weight
// operand split N/2 per CTA) walks tiles of NB x 128 rows of one utterance:
//   warp 0  lane 0: bulk-copies the input tile (128*NB + 2*h1 rows, all channels) into shared memory
//   warp 2  lane 0: streams packed weights through an mbarrier ring (re-used by all NB row blocks)
//   warp 1        : leader CTA: one elected lane issues every tcgen05.mma of both convolutions;
//                   peer CTA: relays its "bytes have landed" barriers to the leader
//   warps 4-7     : epilogue.  e1: D1 (TMEM) + b1 -> lrelu -> bf16 -> shared memory (the conv2 operand,
//                   never leaves the SM);  e2: D2 + b2 + residual -> bf16 stream / fp32 MRF sum in HBM
// Accumulators: D1[NB], D2[NB] in TMEM (2*NB*C <= 512 columns), so e2 of a tile overlaps conv1 of the next.
#include "pair_umma.cuh"

namespace {

// ---------------------------------------------------------------------------------------------------
// PTX helpers beyond common.cuh
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// barrier that receives arrivals from the peer CTA; bounded like mbar_wait (trap, never hang)
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait_cluster(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait_cluster(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) __trap();
    }
}

template <int CG>
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
    if (CG == 1)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc),
                     "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc),
                     "r"(idesc), "r"(acc) : "memory");
}
// all tcgen05.mma issued so far by this thread arrive on `bar` (in every CTA of the pair for CG=2) when done
template <int CG>
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    if (CG == 1)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    else
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                     "h"((uint16_t)3) : "memory");
}
template <int CG>
__device__ __forceinline__ void tmem_alloc_cg(uint32_t dst_smem, uint32_t ncols) {
    if (CG == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
}
template <int CG>
__device__ __forceinline__ void tmem_dealloc_cg(uint32_t taddr, uint32_t ncols) {
    if (CG == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// TMEM -> registers: lane i reads TMEM lane (base + i), 32 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
}

// barrier slots inside the static array
enum {
    BAR_XA_FULL = 0,                                  // [2]  input tile landed (complete_tx)
    BAR_XA_EMPTY = 2,                                 // [2]  conv1 finished reading it (tcgen05.commit)
    BAR_XA_FULL_PEER = 4,                             // [2]  leader only: peer's input tile landed
    BAR_W_FULL = 6,                                   // [PAIR_MAX_SLOTS]
    BAR_W_EMPTY = BAR_W_FULL + PAIR_MAX_SLOTS,
    BAR_W_FULL_PEER = BAR_W_EMPTY + PAIR_MAX_SLOTS,   // leader only
    BAR_D1_FULL = BAR_W_FULL_PEER + PAIR_MAX_SLOTS,   // conv1 accumulators complete
    BAR_T_FULL,                                       // leader only: conv2 operand written by every epilogue warp
    BAR_D2_FULL,
    BAR_D2_EMPTY,                                     // leader only: every epilogue warp has read D2
    BAR_COUNT
};

struct TileCoord {
    int b;          // utterance
    int l0;         // first output sample of the tile
    bool live;      // false: padding iteration (the pair must stay in lock step), nothing is stored
};
__device__ __forceinline__ TileCoord tile_of(const PairP& p, int it) {
    int g = (int)blockIdx.x + it * (int)gridDim.x;
    TileCoord t;
    t.live = g < p.n_tiles;
    if (!t.live) g = p.n_tiles - 1;
    t.b = g / p.tiles_per_utt;
    t.l0 = (g - t.b * p.tiles_per_utt) * p.V;
    return t;
}

// One convolution = n_items weight items (tap, 64-channel block), each applied to all NB row blocks.
// `second`: reads the conv2 operand buffer (dilation 1, block c starts at row c*128 - 2*h2 for c > 0).
template <int CG>
__device__ __forceinline__ void issue_conv(const PairP& p, uint32_t bar0, uint32_t a_addr, uint32_t a_rows, bool second,
                                           uint32_t d_tmem, uint32_t ring_addr, uint32_t idesc, uint32_t& wit) {
    const uint32_t n_loc = (uint32_t)p.C / CG;
    const uint32_t b_lbo = n_loc * 16u, a_lbo = a_rows * 16u;
    const uint64_t a_hi = ((uint64_t)(a_lbo >> 4) << 16) | ((uint64_t)(128u >> 4) << 32) | ((uint64_t)1 << 46);
    const uint64_t b_hi = ((uint64_t)(b_lbo >> 4) << 16) | ((uint64_t)(128u >> 4) << 32) | ((uint64_t)1 << 46);
    const int dil = second ? 1 : p.dil;
    const int kbs_mask = (1 << p.kbs_log2) - 1;
    const uint32_t chunks_per_kb = (uint32_t)p.KB >> 3;
    int item = 0;
    while (item < p.n_items) {
        const uint32_t slot = wit % (uint32_t)p.n_slots, par = (wit / (uint32_t)p.n_slots) & 1u;
        mbar_wait(bar0 + 8u * (BAR_W_FULL + slot), par);
        if (CG == 2) mbar_wait_cluster(bar0 + 8u * (BAR_W_FULL_PEER + slot), par);
        tc_fence_after();
        const int cnt = min(p.items_per_stage, p.n_items - item);
        if (elect_one()) {
            const uint32_t stage = ring_addr + slot * p.stage_bytes;
            for (int i = 0; i < cnt; ++i) {
                const int idx = item + i;
                const int tap = idx >> p.kbs_log2, kb = idx & kbs_mask;
                const uint32_t b_item = stage + (uint32_t)i * p.item_bytes;
                for (int blk = 0; blk < p.NB; ++blk) {
                    const int row0 = second ? (blk == 0 ? 0 : blk * 128 - 2 * p.h2) + tap : blk * 128 + tap * dil;
                    const uint32_t a_blk = a_addr + ((uint32_t)kb * chunks_per_kb * a_rows + (uint32_t)row0) * 16u;
                    const uint32_t d = d_tmem + (uint32_t)(blk * p.C);
                    for (int s = 0; s < p.ksteps; ++s) {
                        const uint64_t ad = a_hi | (uint64_t)(((a_blk + (uint32_t)s * 2u * a_lbo) >> 4) & 0x3FFFu);
                        const uint64_t bd = b_hi | (uint64_t)(((b_item + (uint32_t)s * 2u * b_lbo) >> 4) & 0x3FFFu);
                        mma_bf16<CG>(d, ad, bd, idesc, (idx | s) != 0 ? 1u : 0u);
                    }
                }
            }
            mma_commit<CG>(bar0 + 8u * (BAR_W_EMPTY + slot));     // ring slot reusable once these MMAs retire
        }
        __syncwarp();
        item += cnt;
        ++wit;
    }
}

template <int CG>
__global__ void __launch_bounds__(PAIR_THREADS, 1) pair_umma_kernel(const __grid_constant__ PairP p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[BAR_COUNT];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
    const uint32_t sm0 = smem_u32(smem);
    const uint32_t bar0 = smem_u32(bars);
    const int C = p.C, NB = p.NB;

    if (threadIdx.x == 0) {
        for (int i = 0; i < BAR_COUNT; ++i)
            mbar_init(bar0 + 8u * i, (i == BAR_T_FULL || i == BAR_D2_EMPTY) ? CG * PAIR_EPI_WARPS : 1);
        fence_mbar_init();
    }
    if (warp == 3) tmem_alloc_cg<CG>(smem_u32(&tmem_base_s), 512u);
    // biases -> shared memory
    for (int i = threadIdx.x; i < 2 * C; i += PAIR_THREADS) ((float*)(smem + p.bias_off))[i] = __ldg(p.bias + i);
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    if (warp == 0) {
        // ================= input tile producer =================
        if (lane == 0) {
            const int chunks = C >> 3;
            for (int it = 0; it < p.iters; ++it) {
                const TileCoord t = tile_of(p, it);
                const int buf = it % p.xa_bufs, use = it / p.xa_bufs;
                if (use > 0) mbar_wait(bar0 + 8u * (BAR_XA_EMPTY + buf), (uint32_t)((use - 1) & 1));
                const int start = VO2W_PAD_ROWS + t.l0 - p.h2 - p.h1;          // first staged row inside the padded axis
                const int rows = min(p.xa_rows, p.Lp - start);
                const uint32_t row_bytes = (uint32_t)rows * 16u;
                const uint32_t full = bar0 + 8u * (BAR_XA_FULL + buf);
                mbar_arrive_expect_tx(full, row_bytes * (uint32_t)chunks);
                const __nv_bfloat16* src = p.xa + ((size_t)t.b * chunks * p.Lp + start) * 8;
                const uint32_t dst = sm0 + p.xa_off[buf];
                for (int c = 0; c < chunks; ++c)
                    bulk_g2s(dst + (uint32_t)(c * p.xa_rows) * 16u, src + (size_t)c * p.Lp * 8, row_bytes, full);
            }
        }
    } else if (warp == 2) {
        // ================= weight producer =================
        if (lane == 0) {
            uint32_t wit = 0;
            for (int it = 0; it < p.iters; ++it) {
                for (int conv = 0; conv < 2; ++conv) {
                    const uint8_t* base = (conv ? p.w2 : p.w1) + (size_t)rank * p.conv_bytes_rank;
                    for (int item = 0; item < p.n_items; item += p.items_per_stage, ++wit) {
                        const int cnt = min(p.items_per_stage, p.n_items - item);
                        const uint32_t slot = wit % (uint32_t)p.n_slots, use = wit / (uint32_t)p.n_slots;
                        if (use > 0) mbar_wait(bar0 + 8u * (BAR_W_EMPTY + slot), (use - 1) & 1u);
                        const uint32_t bytes = (uint32_t)cnt * p.item_bytes;
                        const uint32_t full = bar0 + 8u * (BAR_W_FULL + slot);
                        mbar_arrive_expect_tx(full, bytes);
                        bulk_g2s(sm0 + p.ring_off + slot * p.stage_bytes, base + (size_t)item * p.item_bytes, bytes, full);
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (rank == 0) {
            // ================= MMA issuer (whole warp loops, one elected lane issues) =================
            const uint32_t idesc = umma_idesc_bf16(128u * CG, (uint32_t)C);
            const uint32_t d1 = tmem_base, d2 = tmem_base + (uint32_t)(NB * C);
            uint32_t wit = 0;
            for (int it = 0; it < p.iters; ++it) {
                const int buf = it % p.xa_bufs, use = it / p.xa_bufs;
                mbar_wait(bar0 + 8u * (BAR_XA_FULL + buf), (uint32_t)(use & 1));
                if (CG == 2) mbar_wait_cluster(bar0 + 8u * (BAR_XA_FULL_PEER + buf), (uint32_t)(use & 1));
                tc_fence_after();
                issue_conv<CG>(p, bar0, sm0 + p.xa_off[buf], (uint32_t)p.xa_rows, false, d1, sm0 + p.ring_off, idesc, wit);
                if (elect_one()) {
                    mma_commit<CG>(bar0 + 8u * (BAR_XA_EMPTY + buf));   // input tile reusable
                    mma_commit<CG>(bar0 + 8u * BAR_D1_FULL);            // e1 may start
                }
                __syncwarp();
                // conv2 operand written by all epilogue warps of the pair; D2 drained by the previous tile's e2
                mbar_wait_cluster(bar0 + 8u * BAR_T_FULL, (uint32_t)(it & 1));
                if (it > 0) mbar_wait_cluster(bar0 + 8u * BAR_D2_EMPTY, (uint32_t)((it - 1) & 1));
                tc_fence_after();
                issue_conv<CG>(p, bar0, sm0 + p.t_off, (uint32_t)p.t_rows, true, d2, sm0 + p.ring_off, idesc, wit);
                if (elect_one()) mma_commit<CG>(bar0 + 8u * BAR_D2_FULL);
                __syncwarp();
            }
        } else if (lane == 0) {
            // ================= peer relay: forward "landed" events to the leader's barriers =================
            const int stages_per_conv = (p.n_items + p.items_per_stage - 1) / p.items_per_stage;
            uint32_t wit = 0;
            for (int it = 0; it < p.iters; ++it) {
                const int buf = it % p.xa_bufs, use = it / p.xa_bufs;
                mbar_wait(bar0 + 8u * (BAR_XA_FULL + buf), (uint32_t)(use & 1));
                mbar_arrive_remote(mapa_u32(bar0 + 8u * (BAR_XA_FULL_PEER + buf), 0));
                for (int s = 0; s < 2 * stages_per_conv; ++s, ++wit) {
                    const uint32_t slot = wit % (uint32_t)p.n_slots, par = (wit / (uint32_t)p.n_slots) & 1u;
                    mbar_wait(bar0 + 8u * (BAR_W_FULL + slot), par);
                    mbar_arrive_remote(mapa_u32(bar0 + 8u * (BAR_W_FULL_PEER + slot), 0));
                }
            }
        }
    }
    else if (warp >= 4) {
        // ================= epilogue warps =================
        const int q4 = warp & 3;                         // TMEM lane quarter this warp may read
        const int q = q4 * 32 + lane;                    // row inside a 128-row block
        const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const float* sb1 = (const float*)(smem + p.bias_off);
        const float* sb2 = sb1 + C;
        const uint32_t t_full_addr = CG == 2 ? mapa_u32(bar0 + 8u * BAR_T_FULL, 0) : 0u;
        const uint32_t d2_empty_addr = CG == 2 ? mapa_u32(bar0 + 8u * BAR_D2_EMPTY, 0) : 0u;
        const int groups = C >> 5;                       // 32 accumulator columns per tcgen05.ld
        const int chunks = C >> 3;
        for (int it = 0; it < p.iters; ++it) {
            const TileCoord t = tile_of(p, it);
            // ---- e1: conv1 accumulators -> bf16 conv2 operand in shared memory ----
            mbar_wait(bar0 + 8u * BAR_D1_FULL, (uint32_t)(it & 1));
            tc_fence_after();
            for (int blk = 0; blk < NB; ++blk) {
                const int grow = t.l0 - p.h2 + blk * 128 + q;              // sample index of this operand row
                const bool inside = grow >= 0 && grow < p.L;               // conv2 zero-pads outside the utterance
                const uint32_t t_row = sm0 + p.t_off + (uint32_t)(blk * 128 + q) * 16u;
                for (int g = 0; g < groups; ++g) {
                    uint32_t r[32];
                    tmem_ld32(t_lane + (uint32_t)(blk * C + g * 32), r);
                    tmem_ld_wait();
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const float4 ba = *(const float4*)(sb1 + g * 32 + h * 8), bb = *(const float4*)(sb1 + g * 32 + h * 8 + 4);
                        float v[8];
                        v[0] = __uint_as_float(r[h * 8 + 0]) + ba.x; v[1] = __uint_as_float(r[h * 8 + 1]) + ba.y;
                        v[2] = __uint_as_float(r[h * 8 + 2]) + ba.z; v[3] = __uint_as_float(r[h * 8 + 3]) + ba.w;
                        v[4] = __uint_as_float(r[h * 8 + 4]) + bb.x; v[5] = __uint_as_float(r[h * 8 + 5]) + bb.y;
                        v[6] = __uint_as_float(r[h * 8 + 6]) + bb.z; v[7] = __uint_as_float(r[h * 8 + 7]) + bb.w;
                        uint4 o;
                        o.x = pack_bf16x2(lrelu(v[0], 0.1f), lrelu(v[1], 0.1f));
                        o.y = pack_bf16x2(lrelu(v[2], 0.1f), lrelu(v[3], 0.1f));
                        o.z = pack_bf16x2(lrelu(v[4], 0.1f), lrelu(v[5], 0.1f));
                        o.w = pack_bf16x2(lrelu(v[6], 0.1f), lrelu(v[7], 0.1f));
                        if (!inside) o = make_uint4(0u, 0u, 0u, 0u);
                        st_shared_v4(t_row + (uint32_t)((g * 4 + h) * p.t_rows) * 16u, o);
                    }
                }
            }
            fence_proxy_async();                         // generic-proxy stores -> visible to the tensor core (async proxy)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (CG == 2) mbar_arrive_remote(t_full_addr);
                else mbar_arrive(bar0 + 8u * BAR_T_FULL);
            }
            // ---- e2: conv2 accumulators + bias + residual -> global ----
            mbar_wait(bar0 + 8u * BAR_D2_FULL, (uint32_t)(it & 1));
            tc_fence_after();
            for (int blk = 0; blk < NB; ++blk) {
                const int rel = (blk == 0 ? 0 : blk * 128 - 2 * p.h2) + q;
                const int orow = t.l0 + rel;
                const bool ok = t.live && orow < p.L && (blk > 0 || q < 128 - 2 * p.h2);
                const size_t e_row = ((size_t)t.b * chunks * p.Lp + VO2W_PAD_ROWS + orow) * 8;       // chunk 0 of this row (bf16 elements)
                const size_t e_step = (size_t)p.Lp * 8;
                const size_t s_row = ((size_t)t.b * (C >> 2) * p.L + orow) * 4;                       // fp32 sum buffer
                const size_t s_step = (size_t)p.L * 4;
                for (int g = 0; g < groups; ++g) {
                    uint32_t r[32];
                    uint4 res[4];
                    if (ok) {
#pragma unroll
                        for (int h = 0; h < 4; ++h) res[h] = __ldg((const uint4*)(p.x + e_row + (size_t)(g * 4 + h) * e_step));
                    }
                    tmem_ld32(t_lane + (uint32_t)((NB + blk) * C + g * 32), r);
                    tmem_ld_wait();
                    if (ok) {
#pragma unroll
                        for (int h = 0; h < 4; ++h) {
                            const int chunk = g * 4 + h;
                            const float4 ba = *(const float4*)(sb2 + chunk * 8), bb = *(const float4*)(sb2 + chunk * 8 + 4);
                            const float2 x0 = unpack_bf16x2(res[h].x), x1 = unpack_bf16x2(res[h].y);
                            const float2 x2 = unpack_bf16x2(res[h].z), x3 = unpack_bf16x2(res[h].w);
                            float v[8];
                            v[0] = __uint_as_float(r[h * 8 + 0]) + ba.x + x0.x; v[1] = __uint_as_float(r[h * 8 + 1]) + ba.y + x0.y;
                            v[2] = __uint_as_float(r[h * 8 + 2]) + ba.z + x1.x; v[3] = __uint_as_float(r[h * 8 + 3]) + ba.w + x1.y;
                            v[4] = __uint_as_float(r[h * 8 + 4]) + bb.x + x2.x; v[5] = __uint_as_float(r[h * 8 + 5]) + bb.y + x2.y;
                            v[6] = __uint_as_float(r[h * 8 + 6]) + bb.z + x3.x; v[7] = __uint_as_float(r[h * 8 + 7]) + bb.w + x3.y;
                            const size_t e = e_row + (size_t)chunk * e_step;
                            if (p.mode == PAIR_OUT_STREAM) {
                                uint4 o;
                                o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
                                o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
                                *(uint4*)(p.out_raw + e) = o;
                                // the activation is taken of the ROUNDED stream value (what the next layer would compute from HBM)
                                const float2 y0 = unpack_bf16x2(o.x), y1 = unpack_bf16x2(o.y), y2 = unpack_bf16x2(o.z), y3 = unpack_bf16x2(o.w);
                                uint4 a;
                                a.x = pack_bf16x2(lrelu(y0.x, p.out_slope), lrelu(y0.y, p.out_slope));
                                a.y = pack_bf16x2(lrelu(y1.x, p.out_slope), lrelu(y1.y, p.out_slope));
                                a.z = pack_bf16x2(lrelu(y2.x, p.out_slope), lrelu(y2.y, p.out_slope));
                                a.w = pack_bf16x2(lrelu(y3.x, p.out_slope), lrelu(y3.y, p.out_slope));
                                *(uint4*)(p.out_act + e) = a;
                            } else {
                                float* s0 = p.S + s_row + (size_t)(chunk * 2) * s_step;
                                float* s1 = s0 + s_step;
                                if (p.mode == PAIR_OUT_SUM_ADD || p.mode == PAIR_OUT_SUM_LAST) {
                                    const float4 a0 = *(const float4*)s0, a1 = *(const float4*)s1;
                                    v[0] += a0.x; v[1] += a0.y; v[2] += a0.z; v[3] += a0.w;
                                    v[4] += a1.x; v[5] += a1.y; v[6] += a1.z; v[7] += a1.w;
                                }
                                if (p.mode == PAIR_OUT_SUM_FIRST || p.mode == PAIR_OUT_SUM_ADD) {
                                    *(float4*)s0 = make_float4(v[0], v[1], v[2], v[3]);
                                    *(float4*)s1 = make_float4(v[4], v[5], v[6], v[7]);
                                } else {
                                    uint4 a;
#pragma unroll
                                    for (int i = 0; i < 8; ++i) {
                                        // x = xs / num_kernels, stored as bf16, then the next layer's leaky-relu on the bf16 value
                                        v[i] = __bfloat162float(__float2bfloat16_rn(v[i] * p.inv_div));
                                        v[i] = lrelu(v[i], p.out_slope);
                                    }
                                    a.x = pack_bf16x2(v[0], v[1]); a.y = pack_bf16x2(v[2], v[3]);
                                    a.z = pack_bf16x2(v[4], v[5]); a.w = pack_bf16x2(v[6], v[7]);
                                    *(uint4*)(p.out_act + e) = a;
                                }
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (CG == 2) mbar_arrive_remote(d2_empty_addr);
                else mbar_arrive(bar0 + 8u * BAR_D2_EMPTY);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();
    if (warp == 3) tmem_dealloc_cg<CG>(tmem_base, 512u);
}

}  // namespace

// ---------------------------------------------------------------------------------------------------
// host side: weight packing, tile plan, launch
// ---------------------------------------------------------------------------------------------------


void pack_pair_conv_host(std::vector<uint16_t>& out, const float* w, int C, int k, int cg) {
    const int KB = C < 64 ? C : 64, kbs = C / KB, n_loc = C / cg;
    out.assign((size_t)C * C * k, 0);
    size_t o = 0;
    for (int rank = 0; rank < cg; ++rank)
        for (int tap = 0; tap < k; ++tap)
            for (int kb = 0; kb < kbs; ++kb)
                for (int chunk = 0; chunk < KB / 8; ++chunk)
                    for (int n = 0; n < n_loc; ++n)
                        for (int e = 0; e < 8; ++e) {
                            const int co = rank * n_loc + n, ci = kb * KB + chunk * 8 + e;
                            out[o++] = f32_to_bf16_rn(w[((size_t)co * C + ci) * k + tap]);
                        }
}

int pack_pair(PackedPair& P, const float* w1, const float* b1, const float* w2, const float* b2, int C, int k, int dil,
              int cg) {
    VO2W_REQUIRE(C % 32 == 0 && C >= 32 && C <= 256, "pack_pair: channels %d unsupported", C);
    VO2W_REQUIRE((k & 1) && k >= 1 && k <= 11, "pack_pair: kernel size %d unsupported", k);
    VO2W_REQUIRE(cg == 1 || cg == 2, "pack_pair: cluster size %d", cg);
    VO2W_REQUIRE((dil + 1) * (k - 1) / 2 <= VO2W_PAD_ROWS - 2, "pack_pair: halo exceeds the pad rows (k=%d dil=%d)", k, dil);
    P = PackedPair();
    P.C = C; P.k = k; P.dil = dil; P.cg = cg;
    std::vector<uint16_t> h;
    const size_t bytes = (size_t)C * C * k * 2;
    pack_pair_conv_host(h, w1, C, k, cg);
    VO2W_CUDA_OK(cudaMalloc(&P.w1_dev, bytes));
    VO2W_CUDA_OK(cudaMemcpy(P.w1_dev, h.data(), bytes, cudaMemcpyHostToDevice));
    pack_pair_conv_host(h, w2, C, k, cg);
    VO2W_CUDA_OK(cudaMalloc(&P.w2_dev, bytes));
    VO2W_CUDA_OK(cudaMemcpy(P.w2_dev, h.data(), bytes, cudaMemcpyHostToDevice));
    std::vector<float> b(2 * (size_t)C);
    memcpy(b.data(), b1, (size_t)C * 4);
    memcpy(b.data() + C, b2, (size_t)C * 4);
    VO2W_CUDA_OK(cudaMalloc(&P.bias_dev, b.size() * 4));
    VO2W_CUDA_OK(cudaMemcpy(P.bias_dev, b.data(), b.size() * 4, cudaMemcpyHostToDevice));
    return VO2W_OK;
}

void free_pair(PackedPair& P) {
    if (P.w1_dev) cudaFree(P.w1_dev);
    if (P.w2_dev) cudaFree(P.w2_dev);
    if (P.bias_dev) cudaFree(P.bias_dev);
    P.w1_dev = P.w2_dev = nullptr;
    P.bias_dev = nullptr;
}

int pair_plan(PairP& p, int C, int k, int dil, int cg, int B, int L, int n_sms) {
    memset(&p, 0, sizeof(p));
    p.C = C; p.L = L; p.Lp = (int)vo2w_padded_len(L);
    p.k = k; p.dil = dil;
    p.h1 = dil * (k - 1) / 2;
    p.h2 = (k - 1) / 2;
    p.NB = C >= 256 ? 1 : (C >= 64 ? 2 : 4);
    VO2W_REQUIRE(2 * p.NB * C <= 512, "pair_plan: accumulators exceed TMEM");
    p.V = 128 * p.NB - 2 * p.h2;
    p.tiles_per_utt = vo2w_ceil_div(L, p.V);
    p.n_tiles = B * p.tiles_per_utt;
    p.KB = C < 64 ? C : 64;
    p.ksteps = p.KB / 16;
    const int kbs = C / p.KB;
    p.kbs_log2 = kbs == 1 ? 0 : (kbs == 2 ? 1 : 2);
    p.n_items = k * kbs;
    p.item_bytes = (uint32_t)(C / cg) * p.KB * 2;
    p.items_per_stage = (int)(16384u / p.item_bytes);
    if (p.items_per_stage < 1) p.items_per_stage = 1;
    if (p.items_per_stage > p.n_items) p.items_per_stage = p.n_items;
    p.stage_bytes = (uint32_t)p.items_per_stage * p.item_bytes;
    p.n_slots = p.stage_bytes > 16384u ? 2 : PAIR_MAX_SLOTS;
    p.conv_bytes_rank = (uint32_t)((size_t)C * C * k * 2 / cg);
    p.xa_rows = 128 * p.NB + 2 * p.h1;
    p.t_rows = 128 * p.NB;
    const uint32_t xa_bytes = (uint32_t)p.xa_rows * C * 2, t_bytes = (uint32_t)p.t_rows * C * 2;
    const uint32_t fixed = t_bytes + (uint32_t)p.n_slots * p.stage_bytes + 2u * C * 4u;
    p.xa_bufs = (fixed + 2 * xa_bytes <= PAIR_SMEM_LIMIT - 1024) ? 2 : 1;
    uint32_t off = 0;
    for (int i = 0; i < p.xa_bufs; ++i) { p.xa_off[i] = off; off += xa_bytes; }
    if (p.xa_bufs == 1) p.xa_off[1] = p.xa_off[0];
    p.t_off = off; off += t_bytes;
    p.ring_off = off; off += (uint32_t)p.n_slots * p.stage_bytes;
    p.bias_off = off; off += 2u * C * 4u;
    p.smem_bytes = off;
    VO2W_REQUIRE(off <= PAIR_SMEM_LIMIT, "pair_plan: %u bytes of shared memory (C=%d k=%d dil=%d)", off, C, k, dil);
    int grid = n_sms / cg * cg;
    const int need = (p.n_tiles + cg - 1) / cg * cg;
    if (grid > need) grid = need;
    p.iters = (p.n_tiles + grid - 1) / grid;
    return grid;
}

int pair_launch(const PackedPair& P, const PairIO& io, int B, int L, cudaStream_t st) {
    static int n_sms = 0;
    static bool attr_set = false;
    if (!attr_set) {
        int dev = 0;
        VO2W_CUDA_OK(cudaGetDevice(&dev));
        VO2W_CUDA_OK(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev));
        VO2W_CUDA_OK(cudaFuncSetAttribute(pair_umma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_LIMIT));
        VO2W_CUDA_OK(cudaFuncSetAttribute(pair_umma_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, PAIR_SMEM_LIMIT));
        attr_set = true;
    }
    VO2W_REQUIRE(io.xa && io.x, "pair_launch: null input");
    VO2W_REQUIRE(io.mode >= PAIR_OUT_STREAM && io.mode <= PAIR_OUT_SUM_ONLY, "pair_launch: bad mode %d", io.mode);
    VO2W_REQUIRE(io.mode != PAIR_OUT_STREAM || (io.out_raw && io.out_act), "pair_launch: missing stream output");
    VO2W_REQUIRE(io.mode == PAIR_OUT_STREAM || io.mode == PAIR_OUT_SUM_ONLY || io.S, "pair_launch: missing sum buffer");
    VO2W_REQUIRE((io.mode != PAIR_OUT_SUM_LAST && io.mode != PAIR_OUT_SUM_ONLY) || io.out_act, "pair_launch: missing activation output");
    PairP p;
    const int grid = pair_plan(p, P.C, P.k, P.dil, P.cg, B, L, n_sms);
    if (grid < 0) return grid;
    p.xa = io.xa; p.x = io.x; p.out_raw = io.out_raw; p.out_act = io.out_act; p.S = io.S;
    p.w1 = P.w1_dev; p.w2 = P.w2_dev; p.bias = P.bias_dev;
    p.mode = io.mode; p.out_slope = io.out_slope; p.inv_div = io.inv_div;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(PAIR_THREADS);
    cfg.dynamicSmemBytes = p.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)P.cg; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (P.cg == 2) VO2W_CUDA_OK(cudaLaunchKernelEx(&cfg, pair_umma_kernel<2>, p));
    else VO2W_CUDA_OK(cudaLaunchKernelEx(&cfg, pair_umma_kernel<1>, p));
    return VO2W_OK;
}

// ---------------------------------------------------------------------------------------------------
// Known-answer probe: ONE ResBlock iteration through the fused kernel, NCL fp32 in / out.
// ---------------------------------------------------------------------------------------------------
#include "conv_umma.cuh"

namespace {
// bf16 operand buffer [B][C/8][Lp][8] -> fp32 NCL
__global__ void unchunk_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int C, int L, int Lp) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y, b = blockIdx.z;
    if (l >= L) return;
    out[((size_t)b * C + c) * L + l] =
        __bfloat162float(x[(((size_t)b * (C >> 3) + (c >> 3)) * Lp + VO2W_PAD_ROWS + l) * 8 + (c & 7)]);
}
}  // namespace

int unchunk_bf16(const __nv_bfloat16* x, float* out, int B, int C, int L, int Lp, cudaStream_t st) {
    dim3 grid((unsigned)vo2w_ceil_div(L, 128), (unsigned)C, (unsigned)B);
    unchunk_bf16_kernel<<<grid, 128, 0, st>>>(x, out, C, L, Lp);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

extern "C" int vo2w_debug_pair(const float* w1_host, const float* b1_host, const float* w2_host, const float* b2_host,
                               int C, int k, int dil, int cg, const float* x_dev, int B, int L, float* out_dev,
                               float* act_dev, void* stream) {
    VO2W_REQUIRE(w1_host && b1_host && w2_host && b2_host && x_dev && out_dev, "debug_pair: null argument");
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    PackedPair P;
    rc = pack_pair(P, w1_host, b1_host, w2_host, b2_host, C, k, dil, cg);
    if (rc != VO2W_OK) { free_pair(P); return rc; }
    const long long Lp = vo2w_padded_len(L);
    const size_t bytes = (size_t)B * C * Lp * 2;
    __nv_bfloat16* buf[4] = {nullptr, nullptr, nullptr, nullptr};   // xa, x, out_raw, out_act
    for (int i = 0; i < 4 && rc == VO2W_OK; ++i) {
        if (cudaMalloc(&buf[i], bytes) != cudaSuccess || cudaMemsetAsync(buf[i], 0, bytes, st) != cudaSuccess) {
            vo2w_set_error("debug_pair: allocation failed");
            rc = VO2W_E_CUDA;
        }
    }
    if (rc == VO2W_OK) rc = pack_ncl_to_operand(x_dev, buf[0], B, C, L, (int)Lp, 0.1f, st);
    if (rc == VO2W_OK) rc = pack_ncl_to_operand(x_dev, buf[1], B, C, L, (int)Lp, 1.0f, st);
    if (rc == VO2W_OK) {
        PairIO io;
        io.xa = buf[0]; io.x = buf[1]; io.out_raw = buf[2]; io.out_act = buf[3];
        rc = pair_launch(P, io, B, L, st);
    }
    if (rc == VO2W_OK) rc = unchunk_bf16(buf[2], out_dev, B, C, L, (int)Lp, st);
    if (rc == VO2W_OK && act_dev) rc = unchunk_bf16(buf[3], act_dev, B, C, L, (int)Lp, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (rc == VO2W_OK && e != cudaSuccess) {
        vo2w_set_error("debug_pair: kernel failed: %s", cudaGetErrorString(e));
        rc = VO2W_E_CUDA;
    }
    for (int i = 0; i < 4; ++i) if (buf[i]) cudaFree(buf[i]);
    free_pair(P);
    return rc;
}
