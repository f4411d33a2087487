// Persistent single-layer implicit-GEMM convolution on tcgen05 CTA pairs (sm_100a).
//
// Runs conv_pre and the four ConvTranspose1d upsamplers of Generator.forward (scripts/hifigan/models.py:150-153) in
// the bf16 mode.  It replaces umma_conv_kernel on the hot path, which spent 14 % of a forward on 3 % of its FLOPs
// (one non-persistent CTA per 128 rows, one accumulator set, every CTA re-streaming its weights from L2).
//
//   D[q, n] = sum over items i, input channels ci of  X[l0 + q + d_i(tile), ci] * W_i[n, ci]
//
// * time is the UMMA M dimension, output channels (x phase for the transposed convolutions: out[s q + ph] is column
//   (ph, co) of row q) the N dimension, input channels the K dimension; a tap / phase dependency is the SAME staged
//   input tile read at a row offset (SWIZZLE_NONE K-major core matrices, any 16-byte aligned start address);
// * one persistent CTA PAIR (cta_group::2, M = 256 = 2 x 128 rows, the weight operand split N/2 per CTA) walks units
//   = (a tile of NB x 128 input rows per CTA, a group of N-tiles).  The input tile is staged once per unit by 1-D bulk
//   copies and stays for all N-tiles of the group; the next unit's tile is prefetched into L2 (or into the second
//   buffer where shared memory has room);
// * weights stream through a 4-stage mbarrier ring, every stage serving all NB row blocks of both CTAs (M = 256 NB
//   rows per stage); where a unit's weights fit the ring (the s = 2 upsamplers) they are loaded once per CTA;
// * accumulators are double-buffered in TMEM (2 x NB x n_cols <= 512 columns): the 8 epilogue warps drain N-tile t
//   (tcgen05.ld.x32, bias, leaky-relu, bf16 pack) under the MMAs of N-tile t + 1;
// * stores: an N-tile of a transposed convolution holds an even number of phases of the same channels, so a thread
//   owns out[s q + 2j] and out[s q + 2j + 1] of a channel chunk = 32 contiguous bytes = one full sector per store
//   (st.global.v8.b32).  Half-sector (16 B) writes from different N-tiles made L2 fetch every output sector from DRAM
//   first: ncu showed 440 MB read / 598 MB written for ups[1]'s 90 MB in / 361 MB out.
// Warp roles (384 threads): 0 input producer, 1 weight producer, 2 MMA issuer (leader) / barrier relay (peer), 3 TMEM
// owner, 4-11 epilogue.
#include "stream_umma.cuh"
#include "umma_cg.cuh"

#include <algorithm>

namespace {

enum {
    SB_A_FULL = 0,                          // [SU_A_BUFS] input tile landed (complete_tx)
    SB_A_EMPTY = SB_A_FULL + SU_A_BUFS,     // every MMA of the unit has read it (tcgen05.commit, both CTAs)
    SB_A_FULL_PEER = SB_A_EMPTY + SU_A_BUFS,   // (unused: the peer's relay arrives on the leader's SB_A_FULL)
    SB_W_FULL = SB_A_FULL_PEER + SU_A_BUFS, // [SU_SLOTS]
    SB_W_EMPTY = SB_W_FULL + SU_SLOTS,
    SB_W_FULL_PEER = SB_W_EMPTY + SU_SLOTS, // (unused, as above)
    SB_ACC_FULL = SB_W_FULL_PEER + SU_SLOTS,   // [2] accumulator buffer complete (commit, both CTAs)
    SB_ACC_EMPTY = SB_ACC_FULL + 2,         // [2] leader only: all epilogue warps of the pair have read the buffer
    SB_COUNT = SB_ACC_EMPTY + 2
};

// 32 bytes to one 32-byte aligned address: a full L2 sector per thread (SASS: STG.E.ENL2.256)
__device__ __forceinline__ void st_global_v8(void* ptr, uint4 a, uint4 b) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(ptr), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x),
                 "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
}

struct SuTile {
    int b, l0, ng;
    bool live;
};
// unit v of the pair -> this CTA's tile: the pair shares the N-tile group (the B operand is split between the CTAs),
// the two CTAs take neighbouring M tiles
__device__ __forceinline__ SuTile su_tile(const SuP& p, int v, int rank) {
    SuTile t;
    t.ng = v / p.pair_m;
    int mi = 2 * (v - t.ng * p.pair_m) + rank;
    t.live = mi < p.bm_tiles;
    if (!t.live) mi = p.bm_tiles - 1;
    t.b = mi / p.m_tiles;
    t.l0 = (mi - t.b * p.m_tiles) * p.NB * 128;
    return t;
}

__global__ void __launch_bounds__(SU_THREADS, 1) stream_umma_kernel(const __grid_constant__ SuP p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[SB_COUNT];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const uint32_t sm0 = smem_u32(smem);
    const uint32_t bar0 = smem_u32(bars);
    const int pair = (int)blockIdx.x >> 1, n_pairs = (int)gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < SB_COUNT; ++i) {
            // "landed" barriers of the LEADER collect two arrivals per phase: its own producer's (with the byte count) and
            // the peer CTA's relay, so the issuer waits once per operand (as pair_umma_kernel does)
            const bool landed = (i >= SB_A_FULL && i < SB_A_FULL + SU_A_BUFS) || (i >= SB_W_FULL && i < SB_W_FULL + SU_SLOTS);
            mbar_init(bar0 + 8u * i, (i >= SB_ACC_EMPTY && i < SB_ACC_EMPTY + 2) ? 2 * SU_EPI_WARPS : ((landed && rank == 0) ? 2 : 1));
        }
        fence_mbar_init();
    }
    if (threadIdx.x == 0) griddep_launch_dependents();   // programmatic dependent launch, see umma_cg.cuh
    if (warp == 3) tmem_alloc_cg<2>(smem_u32(&tmem_base_s), 512u);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const int a_bufs = p.a_bufs;                                      // 1 .. SU_A_BUFS
    const int stages_per_tile = p.n_items * p.n_kb;
    const uint32_t n_slots = (uint32_t)p.n_slots;     // ring of 4 .. 8 weight stages; position = (use << 8) | slot
    auto ring_next = [n_slots](uint32_t wit) { return ((wit & 0xffu) + 1u == n_slots) ? ((wit & ~0xffu) + 0x100u) : wit + 1u; };

    if (warp == 0) {
        // ================= input tile producer =================
        if (lane == 0) {
            griddep_wait();                              // the input is the previous launch's output
            int n = 0;
            for (int v = pair; v < p.n_units; v += n_pairs, ++n) {
                const SuTile t = su_tile(p, v, (int)rank);
                const int use = n / a_bufs, buf = n - use * a_bufs;
                if (use > 0) mbar_wait(bar0 + 8u * (SB_A_EMPTY + buf), (uint32_t)((use - 1) & 1));
                const int start = VO2W_PAD_ROWS + t.l0 - p.halo_l;        // first staged row inside the padded axis
                const int rows = min(p.a_rows, p.Lp_in - start);          // the last tile of an utterance may be short
                const uint32_t row_bytes = (uint32_t)rows * 16u;
                const uint32_t full = bar0 + 8u * (SB_A_FULL + buf);
                mbar_arrive_expect_tx(full, row_bytes * (uint32_t)p.chunks);
                const __nv_bfloat16* src = p.in + ((size_t)t.b * p.chunks * p.Lp_in + start) * 8;
                const uint32_t dst = sm0 + p.a_off[buf];
                for (int c = 0; c < p.chunks; ++c)
                    bulk_g2s(dst + (uint32_t)(c * p.a_rows) * 16u, src + (size_t)c * p.Lp_in * 8, row_bytes, full);
                // the tile after this one: pull its rows into L2 now, so that its bulk copy (which has to wait for a
                // free buffer) is an L2 hit
                const int vn = v + n_pairs * (p.a_bufs + p.pf_extra);
                if (vn < p.n_units) {
                    const SuTile tn = su_tile(p, vn, (int)rank);
                    const int sn = VO2W_PAD_ROWS + tn.l0 - p.halo_l;
                    const int rn = min(p.a_rows, p.Lp_in - sn);
                    const __nv_bfloat16* pn = p.in + ((size_t)tn.b * p.chunks * p.Lp_in + sn) * 8;
                    for (int c = 0; c < p.chunks; ++c) prefetch_l2(pn + (size_t)c * p.Lp_in * 8, (uint32_t)rn * 16u);
                }
            }
        }
    } else if (warp == 1) {
        // ================= weight producer (same stage order as the MMA issuer) =================
        if (lane == 0) {
            uint32_t wit = 0;
            const uint8_t* wr = p.w + (size_t)rank * p.rank_w_bytes;
            for (int v = pair; v < p.n_units; v += n_pairs) {
                const int ng = v / p.pair_m;
                for (int tt = 0; tt < p.tiles_per_group; ++tt) {
                    const uint8_t* base = wr + (size_t)(ng * p.tiles_per_group + tt) * p.tile_w_bytes;
                    for (int s = 0; s < stages_per_tile; ++s, wit = ring_next(wit)) {
                        const uint32_t slot = wit & 0xffu, use = wit >> 8;
                        if (use > 0) mbar_wait(bar0 + 8u * (SB_W_EMPTY + slot), (use - 1) & 1u);
                        const uint32_t full = bar0 + 8u * (SB_W_FULL + slot);
                        mbar_arrive_expect_tx(full, p.stage_bytes);
                        bulk_g2s(sm0 + p.ring_off + slot * p.stage_bytes, base + (size_t)s * p.stage_bytes, p.stage_bytes, full);
                    }
                }
                if (p.resident) break;          // the ring now holds the layer's weights for good
            }
        }
    } else if (warp == 2 && rank == 0) {
        // ================= MMA issuer: ONE elected thread runs the whole schedule =================
        if (elect_one()) {
            const uint32_t idesc = umma_idesc_bf16(256u, (uint32_t)p.n_cols);
            const uint32_t n_half = (uint32_t)p.n_cols >> 1;
            const uint32_t b_lbo_field = n_half << 16;                   // LBO = n_half * 16 B between the K chunks
            const uint32_t a_rows = (uint32_t)p.a_rows, a_step = 2u * a_rows, b_step = 2u * n_half, n_cols = (uint32_t)p.n_cols;
            const int nb = p.NB, ksteps = p.ksteps;
            uint32_t wit = 0, acc_it = 0;
            int n = 0;
            for (int v = pair; v < p.n_units; v += n_pairs, ++n) {
                const int use = n / a_bufs, buf = n - use * a_bufs;
                mbar_wait_cluster(bar0 + 8u * (SB_A_FULL + buf), (uint32_t)(use & 1));   // both CTAs' tiles
                tc_fence_after();
                const uint32_t a_base = (((sm0 + p.a_off[buf]) >> 4) & 0x3FFFu) | (a_rows << 16);   // LBO = a_rows * 16 B
                const int ng = v / p.pair_m;
                const bool w_wait = !(p.resident && n > 0);
                for (int tt = 0; tt < p.tiles_per_group; ++tt, ++acc_it) {
                    const int nt = ng * p.tiles_per_group + tt;
                    const uint32_t ab = acc_it & 1u, au = acc_it >> 1;
                    if (au > 0) {
                        mbar_wait_cluster(bar0 + 8u * (SB_ACC_EMPTY + ab), (au - 1) & 1u);
                        tc_fence_after();
                    }
                    const uint32_t d_tmem = tmem_base + ab * (uint32_t)(p.NB * p.n_cols);
                    int s = 0;
                    for (int i = 0; i < p.n_items; ++i) {
                        const uint32_t row = (uint32_t)(p.item_shift[i] + p.tile_shift[nt]);
                        for (int kb = 0; kb < p.n_kb; ++kb, ++s) {
                            uint32_t slot;
                            if (w_wait) {
                                slot = wit & 0xffu;
                                const uint32_t par = (wit >> 8) & 1u;
                                mbar_wait_cluster(bar0 + 8u * (SB_W_FULL + slot), par);           // both CTAs' halves
                                tc_fence_after();
                            } else {
                                slot = (uint32_t)(tt * stages_per_tile + s);     // resident: stage index == slot
                            }
                            const uint32_t b_lo = (((sm0 + p.ring_off + slot * p.stage_bytes) >> 4) & 0x3FFFu) | b_lbo_field;
                            uint32_t a_blk = a_base + (uint32_t)(kb * p.KB8) * a_rows + row, d_blk = d_tmem;
                            const uint32_t first = (i | kb) ? 1u : 0u;       // the first K step of a tile overwrites the accumulator
                            if (!(p.dbg & 2))
                            for (int blk = 0; blk < nb; ++blk, a_blk += 128u, d_blk += n_cols) {
                                uint32_t a = a_blk, b = b_lo;
                                mma_lo<2>(d_blk, a, b, idesc, first);
                                for (int ks = 1; ks < ksteps; ++ks) {
                                    a += a_step; b += b_step;
                                    mma_lo<2>(d_blk, a, b, idesc, 1u);
                                }
                            }
                            if (w_wait) {
                                if (!p.resident) mma_commit<2>(bar0 + 8u * (SB_W_EMPTY + slot));   // slot reusable once these retire
                                wit = ring_next(wit);
                            }
                        }
                    }
                    mma_commit<2>(bar0 + 8u * (SB_ACC_FULL + ab));        // the epilogue may drain this buffer
                }
                mma_commit<2>(bar0 + 8u * (SB_A_EMPTY + buf));            // input tile reusable
            }
        }
        __syncwarp();
    } else if (warp == 2) {
        // ================= peer relay: forward "bytes landed" to the leader's barriers =================
        if (lane == 0) {
            uint32_t wit = 0;
            int n = 0;
            for (int v = pair; v < p.n_units; v += n_pairs, ++n) {
                const int use = n / a_bufs, buf = n - use * a_bufs;
                mbar_wait(bar0 + 8u * (SB_A_FULL + buf), (uint32_t)(use & 1));
                mbar_arrive_remote(mapa_u32(bar0 + 8u * (SB_A_FULL + buf), 0));
                if (p.resident && n > 0) continue;
                for (int s = 0; s < p.tiles_per_group * stages_per_tile; ++s, wit = ring_next(wit)) {
                    const uint32_t slot = wit & 0xffu, par = (wit >> 8) & 1u;
                    mbar_wait(bar0 + 8u * (SB_W_FULL + slot), par);
                    mbar_arrive_remote(mapa_u32(bar0 + 8u * (SB_W_FULL + slot), 0));
                }
            }
        }
    } else if (warp >= SU_EPI_WARP0) {
        // ================= epilogue: TMEM -> bias -> leaky-relu -> bf16 stream =================
        // Two warps per TMEM lane quarter split the (row block, 32-column group) items; two TMEM loads in flight.
        const int q4 = warp & 3, half = (warp - SU_EPI_WARP0) >> 2;
        const int q = q4 * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
        const uint32_t acc_empty = mapa_u32(bar0 + 8u * SB_ACC_EMPTY, 0);
        const int G = p.n_cols >> 5, gpc = p.cb >> 5;                      // 32-column groups per tile / per phase of a tile
        const size_t o_step = (size_t)p.Lp_out * 8;                       // next 8-channel chunk of the stream
        const size_t f_step = (size_t)p.L_out * 4;                        // next 4-channel chunk of the fp32 probe output
        const bool paired = p.ph_tile >= 2;                               // the thread owns two consecutive output rows
        const int n_it = paired ? p.NB * (G >> 1) : p.NB * G;             // work items of one accumulator buffer
        uint32_t acc_it = 0;
        // bias + leaky-relu + bf16: 32 columns -> 4 x 16 bytes
        auto convert = [&](const uint32_t (&r)[32], const float* bias, uint4 (&w)[4], float (&v)[32]) {
#pragma unroll
            for (int c = 0; c < 32; c += 4) {     // packed fp32 (FADD2 / FMUL2): two IEEE operations per issue slot
                const float4 bv = __ldg((const float4*)(bias + c));
                add2(v[c], v[c + 1], __uint_as_float(r[c]), __uint_as_float(r[c + 1]), bv.x, bv.y);
                add2(v[c + 2], v[c + 3], __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]), bv.z, bv.w);
            }
            float a[32];                          // leaky_relu(v, slope) = max(v, v * slope) for 0 <= slope <= 1 (su_launch checks)
#pragma unroll
            for (int c = 0; c < 32; c += 2) {
                float n0, n1;
                mul2(n0, n1, v[c], v[c + 1], p.out_slope);
                a[c] = fmaxf(v[c], n0);
                a[c + 1] = fmaxf(v[c + 1], n1);
            }
#pragma unroll
            for (int h = 0; h < 4; ++h) {
                w[h].x = pack_bf16x2(a[8 * h + 0], a[8 * h + 1]);
                w[h].y = pack_bf16x2(a[8 * h + 2], a[8 * h + 3]);
                w[h].z = pack_bf16x2(a[8 * h + 4], a[8 * h + 5]);
                w[h].w = pack_bf16x2(a[8 * h + 6], a[8 * h + 7]);
            }
        };
        auto store_f32 = [&](const float (&v)[32], int b, int co, long long orow) {
            float* o = p.out_f32 + (((size_t)b * (p.Cout >> 2) + (co >> 2)) * p.L_out + orow) * 4;
#pragma unroll
            for (int c = 0; c < 8; ++c) *(float4*)(o + (size_t)c * f_step) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
        };
        for (int v = pair; v < p.n_units; v += n_pairs) {
            const SuTile t = su_tile(p, v, (int)rank);
            for (int tt = 0; tt < p.tiles_per_group; ++tt, ++acc_it) {
                const int nt = t.ng * p.tiles_per_group + tt;
                const int pg = nt / p.cblks, cblk = nt - pg * p.cblks;    // phase group and channel block of this N-tile
                const uint32_t ab = acc_it & 1u, au = acc_it >> 1;
                mbar_wait(bar0 + 8u * (SB_ACC_FULL + ab), au & 1u);
                tc_fence_after();
                const uint32_t t_acc = t_lane + ab * (uint32_t)(p.NB * p.n_cols);
                int idx = half;
                while (idx < n_it) {
                    // paired: item = (row block, phase pair jp, channel group cg): columns of phases 2 jp and 2 jp + 1
                    // single: item = (row block, 32-column group); the second load is the next item of this warp
                    int blk0, blk1, g0, g1;
                    bool two = true;
                    if (paired) {
                        const int per_blk = G >> 1, jc = idx % per_blk, jp = jc / gpc, cg = jc - jp * gpc;
                        blk0 = blk1 = idx / per_blk;
                        g0 = 2 * jp * gpc + cg;
                        g1 = g0 + gpc;
                    } else {
                        blk0 = idx / G; g0 = idx - blk0 * G;
                        two = idx + 2 < n_it;
                        blk1 = two ? (idx + 2) / G : blk0; g1 = two ? idx + 2 - blk1 * G : g0;
                    }
                    uint32_t ra[32], rb[32];
                    tmem_ld32(t_acc + (uint32_t)(blk0 * p.n_cols + g0 * 32), ra);
                    tmem_ld32(t_acc + (uint32_t)(blk1 * p.n_cols + g1 * 32), rb);
                    tmem_ld_wait();
                    if (t.live && !(p.dbg & 1)) {
                        float va[32], vb[32];
                        uint4 wa[4], wb[4];
                        if (paired) {
                            const int row = t.l0 + blk0 * 128 + q;
                            if (row < p.L_in) {
                                const int co = cblk * p.cb + (g0 % gpc) * 32;
                                const long long orow = (long long)p.up_s * row + pg * p.ph_tile + g0 / gpc;   // phase of group g0: even
                                convert(ra, p.bias + nt * p.n_cols + g0 * 32, wa, va);
                                convert(rb, p.bias + nt * p.n_cols + g1 * 32, wb, vb);
                                if (p.out_f32) { store_f32(va, t.b, co, orow); store_f32(vb, t.b, co, orow + 1); }
                                if (p.out) {
                                    // rows orow, orow + 1 of a chunk are 32 contiguous, 32-byte aligned bytes: one full sector
                                    __nv_bfloat16* o = p.out + (((size_t)t.b * (p.Cout >> 3) + (co >> 3)) * p.Lp_out + VO2W_PAD_ROWS + orow) * 8;
#pragma unroll
                                    for (int h = 0; h < 4; ++h) st_global_v8(o + (size_t)h * o_step, wa[h], wb[h]);
                                }
                            }
                        } else {
                            for (int e = 0; e < (two ? 2 : 1); ++e) {
                                const int blk = e ? blk1 : blk0, g = e ? g1 : g0;
                                const int row = t.l0 + blk * 128 + q;
                                if (row >= p.L_in) continue;
                                const int co = cblk * p.cb + g * 32;
                                const long long orow = (long long)p.up_s * row + pg * p.ph_tile;
                                convert(e ? rb : ra, p.bias + nt * p.n_cols + g * 32, wa, va);
                                if (p.out_f32) store_f32(va, t.b, co, orow);
                                if (p.out) {
                                    __nv_bfloat16* o = p.out + (((size_t)t.b * (p.Cout >> 3) + (co >> 3)) * p.Lp_out + VO2W_PAD_ROWS + orow) * 8;
#pragma unroll
                                    for (int h = 0; h < 4; ++h) *(uint4*)(o + (size_t)h * o_step) = wa[h];
                                }
                            }
                        }
                    }
                    idx += (!paired && two) ? 4 : 2;     // single: both items of this round trip are done
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_remote(acc_empty + 8u * ab);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 3) tmem_dealloc_cg<2>(tmem_base, 512u);
}

int round_up_i(int a, int b) { return (a + b - 1) / b * b; }
int floor_div(int a, int b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }
int ceil_div_i(int a, int b) { return -floor_div(-a, b); }

uint16_t bf16_bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

// geometry shared by both packers once Cin / N_total / halos are known
int su_geometry(SuLayer& L, int n_total) {
    L.KB = L.Cin <= 128 ? L.Cin : 64;                 // set again below once n_cols is known
    VO2W_REQUIRE(L.Cin % L.KB == 0, "stream conv: %d input channels do not tile by %d", L.Cin, L.KB);
    L.n_kb = L.Cin / L.KB;
    const int a_halo = L.halo_l + L.halo_r;
    auto fits = [&](int nb, int n_cols) {
        const size_t a = (size_t)(128 * nb + a_halo) * L.Cin * 2, ring = (size_t)4 * (n_cols / 2) * L.KB * 2;   // the ring has at least 4 stages
        return a + ring + 256 <= SU_SMEM_LIMIT;
    };
    if (n_total >= 256) {
        // weights are streamed per unit: prefer 2 row blocks x 128 columns (every weight stage then serves 512 rows of
        // the pair, half the L2 -> SM weight traffic per MMA) where the input tile fits, else 1 x 256
        VO2W_REQUIRE(n_total % 256 == 0, "stream conv: N = %d does not tile", n_total);
        if (fits(2, 128)) { L.NB = 2; L.n_cols = 128; }
        else { VO2W_REQUIRE(fits(1, 256), "stream conv: input tile does not fit shared memory (Cin %d)", L.Cin); L.NB = 1; L.n_cols = 256; }
    } else {
        VO2W_REQUIRE(n_total % 32 == 0, "stream conv: N = %d must be a multiple of 32", n_total);
        L.n_cols = n_total;
        L.NB = std::min(4, 512 / (2 * n_total));
        // a memory-bound layer wants the next tile in flight: leave room for two input buffers
        while (L.NB > 1 && (size_t)2 * (128 * L.NB + a_halo) * L.Cin * 2 + (size_t)4 * (n_total / 2) * L.KB * 2 + 256 > SU_SMEM_LIMIT) L.NB >>= 1;
        VO2W_REQUIRE(fits(L.NB, n_total), "stream conv: input tile does not fit shared memory (Cin %d)", L.Cin);
    }
    L.n_tiles = n_total / L.n_cols;
    VO2W_REQUIRE(L.n_tiles <= SU_MAX_TILES, "stream conv: %d N-tiles", L.n_tiles);
    // a weight stage costs ~570 cycles of bookkeeping in the issuer (wait, peer relay, commit): give it 128 input channels
    // where that keeps the stage at 16 KB, so that a stage carries >= 1000 cycles of MMAs (ups[1]: 8 -> 16 MMAs per stage)
    if (L.Cin > 128 && L.Cin % 128 == 0 && (size_t)(L.n_cols / 2) * 128 * 2 <= 16384) { L.KB = 128; L.n_kb = L.Cin / 128; }
    L.stage_bytes = (uint32_t)(L.n_cols / 2) * L.KB * 2;
    return VO2W_OK;
}

// get(n, ci, nt, i) -> weight of column n of tile nt, input channel ci, item i (0 outside the tensor)
template <typename F>
void su_fill(SuLayer& L, F get) {
    const int n_half = L.n_cols / 2, kb8 = L.KB / 8;
    const size_t stage = (size_t)n_half * L.KB;
    L.w_host.assign((size_t)2 * L.n_tiles * L.n_items * L.n_kb * stage, 0);
    size_t o = 0;
    for (int rank = 0; rank < 2; ++rank)
        for (int nt = 0; nt < L.n_tiles; ++nt)
            for (int i = 0; i < L.n_items; ++i)
                for (int kb = 0; kb < L.n_kb; ++kb)
                    for (int c = 0; c < kb8; ++c)
                        for (int n = 0; n < n_half; ++n)
                            for (int e = 0; e < 8; ++e)
                                L.w_host[o++] = bf16_bits(get(rank * n_half + n, kb * L.KB + c * 8 + e, nt, i));
}

}  // namespace

int su_pack_conv(SuLayer& L, const float* w, const float* bias, int Cout, int Cin, int k, int dil) {
    VO2W_REQUIRE(k >= 1 && k <= SU_MAX_ITEMS && (k & 1), "stream conv: kernel size %d unsupported", k);
    const int h = dil * (k - 1) / 2;
    VO2W_REQUIRE(h <= VO2W_PAD_ROWS, "stream conv: halo %d exceeds the %d pad rows", h, VO2W_PAD_ROWS);
    L = SuLayer();
    L.Cin = round_up_i(Cin, 16);
    L.Cout = round_up_i(Cout, 32);
    L.up_s = 1;
    L.halo_l = L.halo_r = h;
    L.n_items = k;
    int rc = su_geometry(L, L.Cout);
    if (rc != VO2W_OK) return rc;
    for (int j = 0; j < k; ++j) L.item_shift[j] = j * dil;        // staged row of x[q + (j - (k-1)/2) dil] for q = 0
    L.ph_tile = 1; L.cb = L.n_cols; L.cblks = L.n_tiles;          // tile nt = output channels [nt n_cols, (nt + 1) n_cols)
    su_fill(L, [&](int n, int ci, int nt, int i) {
        const int co = nt * L.n_cols + n;
        return (co < Cout && ci < Cin) ? w[((size_t)co * Cin + ci) * k + i] : 0.f;
    });
    L.bias_host.assign((size_t)L.n_tiles * L.n_cols, 0.f);
    for (int n = 0; n < Cout; ++n) L.bias_host[n] = bias[n];
    return VO2W_OK;
}

int su_pack_convT(SuLayer& L, const float* w, const float* bias, int Cin, int Cout, int k, int s) {
    VO2W_REQUIRE(s >= 1 && k >= s && (k - s) % 2 == 0, "stream convT: kernel %d / stride %d unsupported", k, s);
    VO2W_REQUIRE(Cin % 16 == 0 && Cout % 32 == 0, "stream convT: channels must be multiples of 16 (in) and 32 (out)");
    L = SuLayer();
    L.Cin = Cin;
    L.Cout = Cout;
    L.up_s = s;
    const int pad = (k - s) / 2, n_total = s * Cout;
    // out[s q + ph] = sum_d x[q + d] . w[:, :, ph + pad - s d]  for the d with 0 <= ph + pad - s d < k
    const int d_lo = ceil_div_i(0 + pad - k + 1, s), d_hi = floor_div(s - 1 + pad, s);
    L.halo_l = -d_lo;
    L.halo_r = d_hi;
    VO2W_REQUIRE(L.halo_l >= 0 && L.halo_r >= 0 && L.halo_l <= VO2W_PAD_ROWS && L.halo_r <= VO2W_PAD_ROWS, "stream convT: halo");
    int rc = su_geometry(L, n_total);
    if (rc != VO2W_OK) return rc;
    // Columns of a tile: (phase, channel) = ph_tile consecutive phases of cb channels, phase-major.  An even number of
    // phases per tile lets a thread store out[s q + 2j] and out[s q + 2j + 1] of a channel chunk as one 32-byte sector.
    if (n_total < 256) { L.ph_tile = s; L.cb = Cout; }
    else {
        L.ph_tile = s % 4 == 0 ? 4 : (s % 2 == 0 ? 2 : 1);
        while (L.ph_tile > 1 && (L.n_cols / L.ph_tile < 32 || Cout % (L.n_cols / L.ph_tile))) L.ph_tile >>= 1;
        L.cb = L.n_cols / L.ph_tile;
    }
    VO2W_REQUIRE(L.cb % 32 == 0 && Cout % L.cb == 0 && s % L.ph_tile == 0 && L.ph_tile * L.cb == L.n_cols,
                 "stream convT: cannot tile %d phases x %d channels by %d columns", s, Cout, L.n_cols);
    L.cblks = Cout / L.cb;
    int d_min[SU_MAX_TILES];
    L.n_items = 0;
    for (int nt = 0; nt < L.n_tiles; ++nt) {
        const int ph_lo = (nt / L.cblks) * L.ph_tile, ph_hi = ph_lo + L.ph_tile - 1;
        d_min[nt] = ceil_div_i(ph_lo + pad - k + 1, s);
        const int d_max = floor_div(ph_hi + pad, s);
        L.n_items = std::max(L.n_items, d_max - d_min[nt] + 1);
    }
    VO2W_REQUIRE(L.n_items >= 1 && L.n_items <= SU_MAX_ITEMS, "stream convT: %d items", L.n_items);
    for (int nt = 0; nt < L.n_tiles; ++nt) {
        L.tile_shift[nt] = d_min[nt] - d_lo;
        VO2W_REQUIRE(d_min[nt] + L.n_items - 1 <= d_hi, "stream convT: tile %d reads past the halo", nt);
    }
    for (int i = 0; i < L.n_items; ++i) L.item_shift[i] = i;       // halo_l + d_lo + i
    auto col = [&](int nt, int n, int& ph, int& co) {
        ph = (nt / L.cblks) * L.ph_tile + n / L.cb;
        co = (nt % L.cblks) * L.cb + n % L.cb;
    };
    su_fill(L, [&](int n, int ci, int nt, int i) {
        int ph, co;
        col(nt, n, ph, co);
        const int j = ph + pad - s * (d_min[nt] + i);
        return (j >= 0 && j < k) ? w[((size_t)ci * Cout + co) * k + j] : 0.f;
    });
    L.bias_host.assign((size_t)n_total, 0.f);
    for (int nt = 0; nt < L.n_tiles; ++nt)
        for (int n = 0; n < L.n_cols; ++n) {
            int ph, co;
            col(nt, n, ph, co);
            L.bias_host[(size_t)nt * L.n_cols + n] = bias[co];
        }
    return VO2W_OK;
}

int su_upload(SuLayer& L) {
    VO2W_CUDA_OK(cudaMalloc(&L.w_dev, L.w_host.size() * 2));
    VO2W_CUDA_OK(cudaMemcpy(L.w_dev, L.w_host.data(), L.w_host.size() * 2, cudaMemcpyHostToDevice));
    VO2W_CUDA_OK(cudaMalloc(&L.bias_dev, L.bias_host.size() * 4));
    VO2W_CUDA_OK(cudaMemcpy(L.bias_dev, L.bias_host.data(), L.bias_host.size() * 4, cudaMemcpyHostToDevice));
    std::vector<uint16_t>().swap(L.w_host);
    return VO2W_OK;
}

void su_free(SuLayer& L) {
    if (L.w_dev) cudaFree(L.w_dev);
    if (L.bias_dev) cudaFree(L.bias_dev);
    L.w_dev = nullptr;
    L.bias_dev = nullptr;
}

int su_plan(SuP& p, const SuLayer& L, int B, int L_in, int n_sms) {
    memset(&p, 0, sizeof(p));
    VO2W_REQUIRE(B > 0 && L_in > 0, "stream conv: empty batch");
    p.chunks = L.Cin / 8;
    p.KB8 = L.KB / 8; p.n_kb = L.n_kb; p.ksteps = L.KB / 16;
    p.halo_l = L.halo_l;
    p.ph_tile = L.ph_tile; p.cb = L.cb; p.cblks = L.cblks;
    // Row blocks per CTA tile and input buffers.  Weights streamed per unit (big layers): the packer's NB (every weight
    // stage serves NB x 256 rows).  Weights resident (the s = 2 upsamplers: HBM-bound streams): what bounds a stream is
    // the bytes each SM has in flight, so take the (NB, buffers) with the most bytes of input tiles in shared memory.
    const bool can_reside = L.n_tiles * L.n_items * L.n_kb <= 4;
    // streamed weights: a ring as deep as shared memory allows next to one input tile (4 or 8 stages): with 4 stages of
    // 8 KB in flight per SM the L2 -> SM weight stream of ups[1] alone took 119 us (load -> barrier -> relay -> MMA ->
    // commit is a ~1.3 us round trip per stage)
    const uint32_t a_max = ((uint32_t)(128 * L.NB + L.halo_l + L.halo_r) * L.Cin * 2 + 127u) & ~127u;
    p.n_slots = 4;
    if (!can_reside) {
        int max_slots = SU_SLOTS;
#ifdef VO2W_EXPERIMENTS
        if (getenv("VO2W_SU_SLOTS")) max_slots = std::max(4, std::min(SU_SLOTS, atoi(getenv("VO2W_SU_SLOTS"))));
#endif
        while (p.n_slots < max_slots && a_max + (uint32_t)(p.n_slots + 1) * L.stage_bytes <= SU_SMEM_LIMIT) ++p.n_slots;
    }
    p.slots_log2 = 0;                        // unused: the ring position carries slot and use count
    const uint32_t ring = (uint32_t)(can_reside ? L.n_tiles * L.n_items * L.n_kb : p.n_slots) * L.stage_bytes;
    int nb_sel = L.NB, bufs_sel = 1;
    {
        long long best_bytes = -1;
        for (int nb = L.NB; nb >= 1; nb >>= 1) {
            const uint32_t ab = ((uint32_t)(128 * nb + L.halo_l + L.halo_r) * L.Cin * 2 + 127u) & ~127u;
            if (ab + ring > SU_SMEM_LIMIT) continue;
            // measured on B200 (s = 2 upsamplers, B = 128): 2 buffers of the packer's NB beat both 3 buffers (171 vs 139 us at
            // 64 channels) and more, smaller tiles (NB = 1 x 5: 267 vs 147 us at 128 channels -- the per-unit barrier round
            // trip, ~2.4 us, is what small units cannot amortise)
            const int bufs = (int)std::min<uint32_t>(2, (SU_SMEM_LIMIT - ring) / ab);
            const long long bytes = (long long)bufs * ab;
            if (bytes > best_bytes) { best_bytes = bytes; nb_sel = nb; bufs_sel = bufs; }
            break;
        }
    }
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_SU_NB")) { const int f = atoi(getenv("VO2W_SU_NB")); if (f >= 1 && f <= L.NB) { nb_sel = f; bufs_sel = 0; } }
#endif
    p.NB = nb_sel;
    p.a_rows = 128 * p.NB + L.halo_l + L.halo_r;
    p.n_tiles = L.n_tiles; p.n_cols = L.n_cols; p.n_items = L.n_items;
    for (int i = 0; i < SU_MAX_ITEMS; ++i) p.item_shift[i] = L.item_shift[i];
    for (int i = 0; i < SU_MAX_TILES; ++i) p.tile_shift[i] = L.tile_shift[i];
    p.L_in = L_in; p.Lp_in = (int)vo2w_padded_len(L_in);
    p.L_out = L_in * L.up_s; p.Lp_out = (int)vo2w_padded_len(p.L_out);
    p.Cout = L.Cout; p.up_s = L.up_s;
    p.m_tiles = vo2w_ceil_div(L_in, 128 * p.NB);
    p.bm_tiles = B * p.m_tiles;
    p.pair_m = (p.bm_tiles + 1) / 2;
    const int n_pairs = std::max(1, n_sms / 2);
    const uint32_t a_bytes = ((uint32_t)p.a_rows * L.Cin * 2 + 127u) & ~127u;
    p.a_bufs = bufs_sel ? bufs_sel : (int)std::min<uint32_t>(2, (SU_SMEM_LIMIT - ring) / a_bytes);
    p.pf_extra = 1;                          // L2 prefetch runs this many units ahead of the last buffer being filled
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_SU_ABUFS")) p.a_bufs = std::max(1, std::min(p.a_bufs, atoi(getenv("VO2W_SU_ABUFS"))));
    if (getenv("VO2W_SU_PF")) p.pf_extra = atoi(getenv("VO2W_SU_PF"));
#endif
    // N-tile groups: split the N-tiles of an M tile over several units where that balances the persistent pairs.
    // Cost of a unit in tensor-pipe cycles: its N-tiles' MMAs (M = 256: n_cols / 2 cycles per K = 16 step and row block)
    // plus, where the input tile is single-buffered, its reload (nothing overlaps it: ~32 B/clk from L2 + latency).
    int best = 1;
    long long best_cost = -1;
    const long long tile_clk = (long long)L.n_items * L.n_kb * (L.KB / 16) * p.NB * (L.n_cols / 2);
    const long long reload_clk = p.a_bufs == 1 ? (long long)a_bytes / 32 + 1500 : 0;
    for (int ng = 1; ng <= (can_reside ? 1 : L.n_tiles); ng <<= 1) {
        if (L.n_tiles % ng) continue;
        const long long rounds = ((long long)p.pair_m * ng + n_pairs - 1) / n_pairs;
        const long long cost = rounds * ((L.n_tiles / ng) * tile_clk + reload_clk) + ng;        // + ng: ties go to fewer groups
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = ng; }
    }
#ifdef VO2W_EXPERIMENTS
    if (getenv("VO2W_SU_NG")) { const int f = atoi(getenv("VO2W_SU_NG")); if (f >= 1 && L.n_tiles % f == 0) best = f; }
#endif
    p.n_groups = best;
    p.tiles_per_group = L.n_tiles / best;
    p.n_units = p.pair_m * p.n_groups;
    p.stage_bytes = L.stage_bytes;
    p.tile_w_bytes = (uint32_t)(L.n_items * L.n_kb) * L.stage_bytes;
    p.rank_w_bytes = (uint32_t)L.n_tiles * p.tile_w_bytes;
    p.resident = can_reside ? 1 : 0;
    for (int i = 0; i < SU_A_BUFS; ++i) p.a_off[i] = i < p.a_bufs ? (uint32_t)i * a_bytes : 0u;
    p.ring_off = (uint32_t)p.a_bufs * a_bytes;
    p.smem_bytes = p.ring_off + ring;
    VO2W_REQUIRE(p.smem_bytes <= SU_SMEM_LIMIT, "stream conv: %u bytes of shared memory", p.smem_bytes);
    VO2W_REQUIRE(2 * p.NB * p.n_cols <= 512, "stream conv: accumulators exceed TMEM");
    return 2 * std::min(n_pairs, p.n_units);
}

int su_launch(const SuLayer& L, const __nv_bfloat16* in, __nv_bfloat16* out, float* out_f32, float out_slope, int B, int L_in,
              cudaStream_t st) {
    VO2W_REQUIRE(L.w_dev && L.bias_dev && in && (out || out_f32), "stream conv: layer not uploaded or null buffer");
    VO2W_REQUIRE(out_slope >= 0.f && out_slope <= 1.f, "stream conv: output slope %f outside [0, 1]", out_slope);
    static bool attr_set[VO2W_MAX_DEV];
    const int cur = vo2w_cur_dev();
    if (!attr_set[cur]) {
        VO2W_CUDA_OK(cudaFuncSetAttribute(stream_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SU_SMEM_LIMIT));
        attr_set[cur] = true;
    }
    SuP p;
    const int grid = su_plan(p, L, B, L_in, vo2w_sm_count());
    if (grid < 0) return grid;
    p.in = in; p.w = L.w_dev; p.bias = L.bias_dev; p.out = out; p.out_f32 = out_f32; p.out_slope = out_slope;
#ifdef VO2W_EXPERIMENTS
    // what-if timing switches (results are WRONG): 1 = no global stores, 2 = no MMAs
    static const int dbg = getenv("VO2W_SU_DBG") ? atoi(getenv("VO2W_SU_DBG")) : 0;
    p.dbg = dbg;
#endif
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(SU_THREADS);
    cfg.dynamicSmemBytes = p.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = vo2w_pdl_enabled() ? 2 : 1;
    VO2W_CUDA_OK(cudaLaunchKernelEx(&cfg, stream_umma_kernel, p));
    return VO2W_OK;
}

// Host-only probe of the packing (no device): the CPU tests replay the kernel's indexing in numpy from it.
// meta[0..15] = {Cin, Cout, up_s, KB, n_kb, NB, n_cols, n_tiles, n_items, halo_l, halo_r, stage_bytes, w_elems, bias_elems, 0, 0},
// then item_shift[SU_MAX_ITEMS], tile_shift[SU_MAX_TILES].
extern "C" int vo2w_debug_stream_pack(int transposed, const float* w, const float* bias, int Cin, int Cout, int k, int dil_or_stride,
                                      int32_t* meta, uint16_t* w_out, size_t w_capacity, float* bias_out, size_t bias_capacity) {
    VO2W_REQUIRE(w && bias && meta, "debug_stream_pack: null argument");
    SuLayer L;
    int rc = transposed ? su_pack_convT(L, w, bias, Cin, Cout, k, dil_or_stride) : su_pack_conv(L, w, bias, Cout, Cin, k, dil_or_stride);
    if (rc != VO2W_OK) return rc;
    const int32_t head[16] = {L.Cin, L.Cout, L.up_s, L.KB, L.n_kb, L.NB, L.n_cols, L.n_tiles, L.n_items, L.halo_l, L.halo_r,
                              (int32_t)L.stage_bytes, (int32_t)L.w_host.size(), (int32_t)L.bias_host.size(), L.ph_tile, L.cb};
    memcpy(meta, head, sizeof(head));
    for (int i = 0; i < SU_MAX_ITEMS; ++i) meta[16 + i] = L.item_shift[i];
    for (int i = 0; i < SU_MAX_TILES; ++i) meta[16 + SU_MAX_ITEMS + i] = L.tile_shift[i];
    if (w_out) {
        VO2W_REQUIRE(w_capacity >= L.w_host.size() && bias_capacity >= L.bias_host.size(), "debug_stream_pack: buffers too small");
        memcpy(w_out, L.w_host.data(), L.w_host.size() * 2);
        memcpy(bias_out, L.bias_host.data(), L.bias_host.size() * 4);
    }
    return VO2W_OK;
}
