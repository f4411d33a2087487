// fp32-accurate 1-d convolutions / linear layers of the acoustic tail on the tensor pipe.
//
// The reference computes these layers in fp32 (scripts/transformer/SubLayers.py, Layers.py).  tcgen05.mma kind::tf32
// keeps 10 mantissa bits per operand, so one MMA per product would miss the 1e-4 gate after a few layers.  Each
// operand is therefore split in two TF32-exact terms, a = hi + lo with hi = tf32(a), lo = tf32(a - hi), and the
// product is accumulated as  hi.hi + lo.hi + hi.lo  in the fp32 TMEM accumulator (the dropped lo.lo term is ~2^-22
// relative): three MMAs per k-step at the TF32 rate.  The tensor pipe adds into its fp32 accumulator with truncation,
// a bias that grows with the number of accumulations (measured 8.6e-5 max-abs at K = 2304 with one accumulator), so
// the hi.hi chain and the two small-term chains get SEPARATE TMEM accumulators (the small chain's truncation is
// 2^-11 of the main one's) that the epilogue adds in registers: a third of the accumulations at full magnitude.
// What remains is a shrink of the main accumulator that is linear in its number of accumulations n = Cin*k/8:
// measured -1.63e-8 * n (+- 2 %, n = 8..384) for zero-mean products, -5.3e-8 * n for same-sign products
// (profiles/r01_tf32_accumulation_bias.txt).  The epilogue multiplies the main accumulator by (1 + 1.63e-8 * n),
// which removes the mean of the error for activations-times-weights data: 3.5e-5 -> 1.3e-5 max-abs at n = 320.
//
// Geometry: activations are time-major [B,L,C]; one CTA owns NB blocks of 128 consecutive frames of one utterance x
// one tile of n_t output channels (accumulators: 2 x NB x n_t TMEM columns; n_t = 128, NB = 2 for the wide layers).
// Input channels are walked in chunks of 8 (one MMA k-step): per chunk the frames (with the k-1 halo rows, zero
// outside the utterance) are split and staged K-major in SWIZZLE_NONE core-matrix layout, so tap j of the convolution
// is the same tile at a row offset of j (16 B per row), exactly as in the bf16 vocoder kernels.  Weights are packed
// once (tf32_conv_pack) in consumption order and streamed by cp.async.bulk through an mbarrier ring; the NB row blocks
// share every weight stage.  A strided / per-head view of the same kernel (ldx, ldo, heads, w_zs) runs the two batched
// GEMMs of attention, with the keys / values of each (utterance, head) packed as that GEMM's weights.
//
// Warp roles (320 threads): warps 0-7 stage the activation chunks (two groups of four, alternating chunk pairs, so
// four chunk loads are in flight) and later run the epilogue; warp 8 streams the weights; warp 9 owns TMEM and issues
// the MMAs from one elected thread.
#include "conv_tf32.cuh"

namespace {

constexpr int T32_THREADS = 320;
constexpr int T32_AS = 4;        // activation stages
constexpr int T32_WS = 8;        // weight stages
constexpr int T32_KC = 8;        // input channels per chunk (one kind::tf32 k-step)
constexpr float T32_TRUNC_PER_ACC = 1.63e-8f;     // measured relative loss of the truncating accumulator per accumulation

struct T32K {
    Tf32Conv c;
    int n_t, n_tiles, chunks, R, R8;
    uint32_t a_stage_bytes, w_stage_bytes, tmem_cols;
};

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// instruction descriptor for kind::tf32: D fp32, A/B tf32 (format 2), both K-major, M x N
__device__ __forceinline__ uint32_t umma_idesc_tf32(uint32_t M, uint32_t N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// nearest TF32 (10 mantissa bits), ties away from zero; the result has its low 13 bits clear
__host__ __device__ __forceinline__ float tf32_round(float v) {
#ifdef __CUDA_ARCH__
    return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u);
#else
    union { float f; uint32_t u; } t;
    t.f = v;
    t.u = (t.u + 0x1000u) & 0xffffe000u;
    return t.f;
#endif
}
__device__ __forceinline__ void st_shared_f4(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// one instruction per element: nearest TF32, low 13 bits clear
__device__ __forceinline__ float tf32_rna(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}
// The staging loop is ALU-bound on the k = 1 layers, so the split is 2 instructions per element: hi = tf32(v), and
// lo = v - hi is stored as it is (exact in fp32; the tensor pipe reads its top 19 bits, an error of 2^-21 |v|).
__device__ __forceinline__ void split4(const float4& v, float4& hi, float4& lo) {
    hi.x = tf32_rna(v.x); hi.y = tf32_rna(v.y); hi.z = tf32_rna(v.z); hi.w = tf32_rna(v.w);
    lo.x = v.x - hi.x; lo.y = v.y - hi.y; lo.z = v.z - hi.z; lo.w = v.w - hi.w;
}

template <int NB>
__global__ void __launch_bounds__(T32_THREADS) tf32_conv_kernel(const __grid_constant__ T32K p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[2 * T32_AS + 2 * T32_WS + 1];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int l0 = blockIdx.x * 128 * NB, tile = blockIdx.y;
    const int b = (int)blockIdx.z / p.c.heads, zh = (int)blockIdx.z % p.c.heads;          // heads >= 1 (host)
    const int k = p.c.k, dil = p.c.dil, pad = dil * (k - 1) / 2, L = p.c.L, ldx = p.c.ldx, n_t = p.n_t, R8 = p.R8;

    const uint32_t a_ring = smem_u32(smem);
    const uint32_t w_ring = a_ring + T32_AS * p.a_stage_bytes;
    const uint32_t bar0 = smem_u32(bars);
    auto a_full = [&](int s) { return bar0 + 8u * s; };
    auto a_empty = [&](int s) { return bar0 + 8u * (T32_AS + s); };
    auto w_full = [&](int s) { return bar0 + 8u * (2 * T32_AS + s); };
    auto w_empty = [&](int s) { return bar0 + 8u * (2 * T32_AS + T32_WS + s); };
    const uint32_t acc_full = bar0 + 8u * (2 * T32_AS + 2 * T32_WS);

    if (threadIdx.x == 0) {
        for (int s = 0; s < T32_AS; ++s) { mbar_init(a_full(s), 128); mbar_init(a_empty(s), 1); }
        for (int s = 0; s < T32_WS; ++s) { mbar_init(w_full(s), 1); mbar_init(w_empty(s), 1); }
        mbar_init(acc_full, 1);
        fence_mbar_init();
    }
    if (threadIdx.x == 0) griddep_launch_dependents();   // programmatic dependent launch (common.cuh): the next kernel of
                                                         // the stream may set up on SMs this grid has left
    if (warp == 9) tmem_alloc(smem_u32(&tmem_base_s), p.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    griddep_wait();                                      // activations, residuals (and, in attention, the packed operand) are
                                                         // the previous kernels' outputs

    if (warp < 8) {
        // ===================== activation producers: global fp32 -> hi/lo core matrices =====================
        // Each group of four warps takes chunks in pairs (group 0: 0,1,4,5,..; group 1: 2,3,6,7,..).  A pair is 64
        // contiguous bytes per frame: lane = (16-byte segment, row) so that one warp load touches 8 rows x 64 B (8 cache
        // lines, not 32) and one quarter-warp store phase writes 128 contiguous bytes of one k-column.
        const int g = warp >> 2, wq = warp & 3, seg = lane >> 3, r_in = wq * 8 + (lane & 7);
        constexpr int NRI = NB * 4 + 2;                  // row iterations: (NB*128 + halo <= 50, rounded to 8) / 32 rows per pass
        const float slope = p.c.in_slope;
        const float* xb = p.c.x + (size_t)b * L * ldx + (size_t)zh * p.c.x_hs;
        for (int c0 = 2 * g; c0 < p.chunks; c0 += 4) {
            const bool pair = c0 + 1 < p.chunks;
            const bool mine = seg < 2 || pair;            // segments 2, 3 belong to the second chunk of the pair
            float4 v[NRI];
#pragma unroll
            for (int i = 0; i < NRI; ++i) {
                const int r = i * 32 + r_in, l = l0 - pad + r;
                // nothing but the loads in this loop: they must all be in flight before the first one is consumed
                if (mine && r < p.R && l >= 0 && l < L)
                    v[i] = __ldg(reinterpret_cast<const float4*>(xb + (size_t)l * ldx + c0 * T32_KC) + seg);
                else
                    v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            const int s0 = c0 % T32_AS, s1 = (c0 + 1) % T32_AS, use = c0 / T32_AS;      // c0 is even: same `use` for both
            if (use > 0) {
                mbar_wait(a_empty(s0), (uint32_t)((use - 1) & 1));
                if (pair) mbar_wait(a_empty(s1), (uint32_t)((use - 1) & 1));
            }
            if (mine) {
                const uint32_t base = a_ring + (uint32_t)(seg < 2 ? s0 : s1) * p.a_stage_bytes + (uint32_t)((seg & 1) * R8) * 16u;
#pragma unroll
                for (int i = 0; i < NRI; ++i) {
                    const int r = i * 32 + r_in;
                    if (r < R8) {
                        float4 t = v[i], hi, lo;
                        if (slope != 1.f) {          // warp-uniform
                            t.x = lrelu(t.x, slope); t.y = lrelu(t.y, slope); t.z = lrelu(t.z, slope); t.w = lrelu(t.w, slope);
                        }
                        split4(t, hi, lo);
                        st_shared_f4(base + (uint32_t)r * 16u, hi);
                        st_shared_f4(base + (uint32_t)(2 * R8 + r) * 16u, lo);
                    }
                }
            }
            fence_proxy_async();                          // generic-proxy stores -> visible to the tensor pipe
            mbar_arrive(a_full(s0));
            if (pair) mbar_arrive(a_full(s1));
        }

        // ===================== epilogue: TMEM -> bias / residual / activation -> global =====================
        mbar_wait(acc_full, 0);
        tc_fence_after();
        const int q = warp & 3, h = warp >> 2, m = q * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
        const int co0 = tile * n_t;
        const float comp = 1.0f + T32_TRUNC_PER_ACC * (float)(p.chunks * k);
#pragma unroll 1
        for (int blk = 0; blk < NB; ++blk) {
            const int l = l0 + blk * 128 + m;
            const bool live = l < L;
            const size_t row = ((size_t)b * L + (live ? l : 0)) * p.c.ldo + (size_t)zh * p.c.o_hs + co0;
#pragma unroll 1
            for (int cc = h; cc < n_t / 16; cc += 2) {
                uint32_t r0[8], r1[8], s0[8], s1[8];
                tmem_ld8(t_lane + (uint32_t)(blk * 2 * n_t + cc * 16), r0);
                tmem_ld8(t_lane + (uint32_t)(blk * 2 * n_t + cc * 16 + 8), r1);
                tmem_ld8(t_lane + (uint32_t)(blk * 2 * n_t + n_t + cc * 16), s0);
                tmem_ld8(t_lane + (uint32_t)(blk * 2 * n_t + n_t + cc * 16 + 8), s1);
                tmem_ld_wait();
                if (live) {
                float y[16];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    y[i] = fmaf(__uint_as_float(r0[i]), comp, __uint_as_float(s0[i]));
                    y[8 + i] = fmaf(__uint_as_float(r1[i]), comp, __uint_as_float(s1[i]));
                }
                if (p.c.bias) {
                    const float4* bp = reinterpret_cast<const float4*>(p.c.bias + co0 + cc * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const float4 t = __ldg(bp + i); y[4 * i] += t.x; y[4 * i + 1] += t.y; y[4 * i + 2] += t.z; y[4 * i + 3] += t.w; }
                }
                if (p.c.res) {
                    const float4* rp = reinterpret_cast<const float4*>(p.c.res + row + cc * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const float4 t = __ldg(rp + i); y[4 * i] += t.x; y[4 * i + 1] += t.y; y[4 * i + 2] += t.z; y[4 * i + 3] += t.w; }
                }
                if (p.c.acc) {
                    const float4* ap = reinterpret_cast<const float4*>(p.c.acc + row + cc * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const float4 t = ap[i]; y[4 * i] += t.x; y[4 * i + 1] += t.y; y[4 * i + 2] += t.z; y[4 * i + 3] += t.w; }
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) y[i] *= p.c.out_scale;
                if (p.c.act == 1) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) y[i] = fmaxf(y[i], 0.f);
                } else if (p.c.act == 2) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) y[i] = tanhf(y[i]);
                }
                float4* op = reinterpret_cast<float4*>(p.c.out + row + cc * 16);
#pragma unroll
                for (int i = 0; i < 4; ++i) op[i] = make_float4(y[4 * i], y[4 * i + 1], y[4 * i + 2], y[4 * i + 3]);
                }
                __syncwarp();
            }
        }
    } else if (warp == 8) {
        // ===================== weight producer: packed stages through the bulk-copy engine =====================
        if (lane == 0) {
            const int n_it = p.chunks * k;
            const float* src = p.c.wp + (size_t)blockIdx.z * p.c.w_zs + (size_t)tile * n_it * (p.w_stage_bytes / 4);
            for (int it = 0; it < n_it; ++it) {
                const int s = it % T32_WS;
                if (it >= T32_WS) mbar_wait(w_empty(s), (uint32_t)((it / T32_WS - 1) & 1));
                mbar_arrive_expect_tx(w_full(s), p.w_stage_bytes);
                bulk_g2s(w_ring + (uint32_t)s * p.w_stage_bytes, src + (size_t)it * (p.w_stage_bytes / 4), p.w_stage_bytes, w_full(s));
            }
        }
    } else {
        // ===================== MMA issuer: one elected thread =====================
        if (elect_one_sync()) {
            // w stage rows = [hi (n_t) ; lo (n_t)]: ONE MMA of width 2 n_t gives a_hi.w_hi | a_hi.w_lo side by side (the
            // activation tile is read once for both), a second of width n_t adds a_lo.w_hi onto the small half
            const uint32_t idesc2 = umma_idesc_tf32(128, (uint32_t)(2 * n_t)), idesc1 = umma_idesc_tf32(128, (uint32_t)n_t);
            // descriptors differ only in their 16-byte-unit start address (low 14 bits): build two templates, add offsets
            const uint64_t a_desc0 = umma_smem_desc(a_ring, (uint32_t)R8 * 16u, 128u);
            const uint64_t w_desc0 = umma_smem_desc(w_ring, (uint32_t)(2 * n_t) * 16u, 128u);
            const uint32_t a_stage_u = p.a_stage_bytes >> 4, w_stage_u = p.w_stage_bytes >> 4;
            const uint32_t a_lo_u = 2u * (uint32_t)R8;
            int wit = 0, ws = 0;
            uint32_t w_par = 0;
            for (int c = 0; c < p.chunks; ++c) {
                const int as = c % T32_AS;
                mbar_wait(a_full(as), (uint32_t)((c / T32_AS) & 1));
                tc_fence_after();
                const uint64_t a_stage = a_desc0 + (uint64_t)((uint32_t)as * a_stage_u);
                for (int j = 0; j < k; ++j, ++wit) {
                    mbar_wait(w_full(ws), w_par);
                    tc_fence_after();
                    const uint64_t b_desc = w_desc0 + (uint64_t)((uint32_t)ws * w_stage_u);
                    const uint32_t first = (c | j) ? 1u : 0u;
#pragma unroll
                    for (int blk = 0; blk < NB; ++blk) {
                        const uint64_t a_hi = a_stage + (uint64_t)(uint32_t)(blk * 128 + j * dil), a_lo = a_hi + a_lo_u;
                        const uint32_t d_main = tmem_base + (uint32_t)(blk * 2 * n_t);
                        umma_tf32(d_main, a_hi, b_desc, idesc2, first);                  // [main | small] (+)= a_hi . [w_hi ; w_lo]
                        umma_tf32(d_main + (uint32_t)n_t, a_lo, b_desc, idesc1, 1u);       // small += a_lo . w_hi
                    }
                    umma_commit(w_empty(ws));
                    if (++ws == T32_WS) { ws = 0; w_par ^= 1u; }
                }
                umma_commit(a_empty(as));
            }
            umma_commit(acc_full);
        }
        __syncwarp();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(tmem_base, p.tmem_cols);
}

__global__ void __launch_bounds__(256) tf32_pack_kernel(const float* __restrict__ w, int Cout, int Cin, int k, int n_t,
                                                        float* __restrict__ packed) {
    const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= (long long)Cout * Cin * k) return;
    const int j = (int)(e % k), ci = (int)((e / k) % Cin), co = (int)(e / ((long long)k * Cin));
    const float v = w[e], hi = tf32_round(v), lo = tf32_round(v - hi);
    const int tile = co / n_t, r = co % n_t, c = ci / T32_KC, kk = ci % T32_KC, chunks = Cin / T32_KC;
    const size_t stage = ((size_t)tile * chunks + c) * k + j;
    // stage = [2 k-columns][2 n_t rows: hi then lo][4 floats]
    const size_t idx = stage * (size_t)(16 * n_t) + (size_t)(kk >> 2) * (n_t * 8) + (size_t)r * 4 + (kk & 3);
    packed[idx] = hi;
    packed[idx + 4 * (size_t)n_t] = lo;
}

// strided gather of a [Cout, Cin] matrix per (b, h) into the packed k = 1 stage layout.  One thread per (output row co,
// chunk of T32_KC input channels), co fastest across threads: the four 16-byte stores of a thread (2 k-columns x hi / lo)
// are each contiguous over the lanes of a warp (r * 4 floats), and the 8 source values of a thread are one 32-byte
// sector when the input-channel axis is the contiguous one (keys: ld_ci = 1), or 8 loads each coalesced over co when
// the output axis is (values, transposed while packing: ld_co = 1).
__global__ void __launch_bounds__(256) tf32_pack_strided_kernel(const float* __restrict__ src, size_t bs, size_t hs, size_t ld_co,
                                                                size_t ld_ci, int heads, int Cout, int Cin, int co_valid, int ci_valid,
                                                                int n_t, float* __restrict__ packed) {
    const int z = blockIdx.y, chunks = Cin / T32_KC;
    const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= (long long)Cout * chunks) return;
    const int co = (int)(e % Cout), c = (int)(e / Cout);
    const float* s = src + (size_t)(z / heads) * bs + (size_t)(z % heads) * hs + (size_t)co * ld_co + (size_t)c * T32_KC * ld_ci;
    float v[T32_KC];
#pragma unroll
    for (int q = 0; q < T32_KC; ++q) {
        v[q] = 0.f;
        if (co < co_valid && c * T32_KC + q < ci_valid) v[q] = __ldg(s + (size_t)q * ld_ci);
    }
    const int tile = co / n_t, r = co % n_t;
    float* o = packed + (size_t)z * 2 * Cout * Cin + ((size_t)tile * chunks + c) * (size_t)(16 * n_t) + (size_t)r * 4;
#pragma unroll
    for (int h = 0; h < 2; ++h) {                     // k-columns 0..3 and 4..7 of the chunk
        float4 hi, lo;
        hi.x = tf32_round(v[4 * h + 0]); lo.x = tf32_round(v[4 * h + 0] - hi.x);
        hi.y = tf32_round(v[4 * h + 1]); lo.y = tf32_round(v[4 * h + 1] - hi.y);
        hi.z = tf32_round(v[4 * h + 2]); lo.z = tf32_round(v[4 * h + 2] - hi.z);
        hi.w = tf32_round(v[4 * h + 3]); lo.w = tf32_round(v[4 * h + 3] - hi.w);
        *reinterpret_cast<float4*>(o + (size_t)h * (n_t * 8)) = hi;
        *reinterpret_cast<float4*>(o + (size_t)h * (n_t * 8) + 4 * (size_t)n_t) = lo;
    }
}

// output channels per CTA: 128 wherever it divides (2 row blocks x [main | small] x 128 = all 512 TMEM columns; the
// stacked [w_hi ; w_lo] MMA is then 256 wide, which shared-memory operand bandwidth favours: measured 8 % faster on
// the k = 9 / k = 5 layers and 14 % on the k = 1 layers than 64-wide tiles with 4 row blocks), else 64, else the whole
// of a narrow layer (mel_linear / PostNet's 80)
int tile_width(int Cout, int k) {
    (void)k;
    if (Cout % 128 == 0) return 128;
    if (Cout % 64 == 0) return 64;
    if (Cout <= 128 && Cout % 16 == 0) return Cout;
    return 0;
}

template <int NB>
int launch(const T32K& kp, size_t smem, cudaStream_t st) {
    static bool attr[VO2W_MAX_DEV];
    const int cur = vo2w_cur_dev();
    if (!attr[cur]) {
        VO2W_CUDA_OK(cudaFuncSetAttribute(tf32_conv_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
        attr[cur] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)vo2w_ceil_div(kp.c.L, 128 * NB), (unsigned)kp.n_tiles, (unsigned)(kp.c.B * kp.c.heads));
    cfg.blockDim = dim3(T32_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = vo2w_pdl_enabled() ? 1 : 0;
    VO2W_CUDA_OK(cudaLaunchKernelEx(&cfg, tf32_conv_kernel<NB>, kp));
    return VO2W_OK;
}

}  // namespace

bool tf32_conv_supported(int Cout, int Cin, int k) {
    return tile_width(Cout, k) > 0 && Cin > 0 && Cin % T32_KC == 0 && (k & 1) && k >= 1 && k <= 11;
}

size_t tf32_conv_packed_floats(int Cout, int Cin, int k) { return (size_t)2 * Cout * Cin * k; }

int tf32_conv_pack(const float* w, int Cout, int Cin, int k, float* packed, cudaStream_t st) {
    VO2W_REQUIRE(tf32_conv_supported(Cout, Cin, k), "tf32 conv: shape Cout %d Cin %d k %d unsupported", Cout, Cin, k);
    const long long n = (long long)Cout * Cin * k;
    tf32_pack_kernel<<<(unsigned)vo2w_ceil_div(n, 256), 256, 0, st>>>(w, Cout, Cin, k, tile_width(Cout, k), packed);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int tf32_pack_strided(const float* src, size_t bs, size_t hs, size_t ld_co, size_t ld_ci, int B, int heads, int Cout, int Cin,
                      int co_valid, int ci_valid, float* packed, cudaStream_t st) {
    VO2W_REQUIRE(tf32_conv_supported(Cout, Cin, 1) && B > 0 && heads > 0 && (long long)B * heads <= 65535,
                 "tf32 pack: shape Cout %d Cin %d unsupported", Cout, Cin);
    VO2W_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0, "tf32 pack: the packed operand must be 16-byte aligned");
    dim3 grid((unsigned)vo2w_ceil_div((long long)Cout * (Cin / T32_KC), 256), (unsigned)(B * heads));
    tf32_pack_strided_kernel<<<grid, 256, 0, st>>>(src, bs, hs, ld_co, ld_ci, heads, Cout, Cin, co_valid, ci_valid, tile_width(Cout, 1),
                                                   packed);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int tf32_conv1d(const Tf32Conv& c, cudaStream_t st) {
    VO2W_REQUIRE(tf32_conv_supported(c.Cout, c.Cin, c.k), "tf32 conv: shape Cout %d Cin %d k %d unsupported", c.Cout, c.Cin, c.k);
    VO2W_REQUIRE(c.B > 0 && c.B <= 65535 && c.L > 0, "tf32 conv: bad batch (B=%d L=%d)", c.B, c.L);
    VO2W_REQUIRE((((uintptr_t)c.x | (uintptr_t)c.wp | (uintptr_t)c.bias | (uintptr_t)c.res | (uintptr_t)c.out) & 15) == 0,
                 "tf32 conv: pointers must be 16-byte aligned");
    T32K kp;
    kp.c = c;
    if (kp.c.ldx == 0) kp.c.ldx = c.Cin;
    if (kp.c.ldo == 0) kp.c.ldo = c.Cout;
    if (kp.c.heads < 1) kp.c.heads = 1;
    if (kp.c.dil < 1) kp.c.dil = 1;
    if (kp.c.in_slope == 0.f) kp.c.in_slope = 1.f;
    if (kp.c.out_scale == 0.f) kp.c.out_scale = 1.f;
    VO2W_REQUIRE((c.k - 1) * kp.c.dil <= 50 && (((uintptr_t)c.acc) & 15) == 0, "tf32 conv: halo (k-1)*dil = %d rows > 50", (c.k - 1) * kp.c.dil);
    VO2W_REQUIRE(kp.c.ldx % 4 == 0 && kp.c.ldo % 4 == 0 && kp.c.x_hs % 4 == 0 && kp.c.o_hs % 4 == 0 && kp.c.w_zs % 4 == 0 &&
                 (long long)c.B * kp.c.heads <= 65535, "tf32 conv: strides must be multiples of 4 floats, B * heads <= 65535");
    kp.n_t = tile_width(c.Cout, c.k);
    kp.n_tiles = c.Cout / kp.n_t;
    kp.chunks = c.Cin / T32_KC;
    // two row blocks at most: with 128-wide tiles that fills TMEM; narrower layers then fit two CTAs per SM, which hide
    // each other's prologue / pipeline fill / epilogue (4 blocks per CTA: 49.1 ms for the vocoder's fp32 mode, 2: 45.1)
    int NB = c.L > 128 ? 2 : 1;
    while (2 * NB * kp.n_t > 512) NB /= 2;
    // k = 1 layers are bound by the latency of the activation loads, not by the weight stream: half-size CTAs (two
    // per SM by registers, shared memory and TMEM columns) keep twice the loads in flight and hide each other's
    // prologue / epilogue
    if (c.k == 1) while (NB > 1 && 2 * NB * kp.n_t > 256) NB /= 2;      // decoder 8.8 -> 8.2 ms with 128-wide tiles
    kp.R = NB * 128 + (c.k - 1) * kp.c.dil;
    kp.R8 = (kp.R + 7) / 8 * 8;
    kp.a_stage_bytes = (uint32_t)(4 * kp.R8 * 16);
    kp.w_stage_bytes = (uint32_t)(64 * kp.n_t);
    uint32_t cols = 32;
    while (cols < (uint32_t)(2 * NB * kp.n_t)) cols *= 2;
    kp.tmem_cols = cols;
    const size_t smem = (size_t)T32_AS * kp.a_stage_bytes + (size_t)T32_WS * kp.w_stage_bytes;
    if (NB == 2) return launch<2>(kp, smem, st);
    return launch<1>(kp, smem, st);
}

// test / profiling entry: Conv1d weights [Cout,Cin,k] are packed on the fly into `scratch` (2*Cout*Cin*k floats)
extern "C" int vo2w_debug_tf32_conv_ex(const float* x, const float* w, const float* bias, const float* res, const float* acc,
                                       float* out, int B, int L, int Cin, int Cout, int k, int dil, float in_slope, float out_scale,
                                       int act, float* scratch, void* stream) {
    VO2W_REQUIRE(x && w && out && scratch, "debug_tf32_conv: null argument");
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if ((rc = tf32_conv_pack(w, Cout, Cin, k, scratch, st)) != VO2W_OK) return rc;
    Tf32Conv c{x, scratch, bias, res, out, B, L, Cin, Cout, k, act};
    c.dil = dil; c.in_slope = in_slope; c.acc = acc; c.out_scale = out_scale;
    return tf32_conv1d(c, st);
}

extern "C" int vo2w_debug_tf32_conv(const float* x, const float* w, const float* bias, const float* res, float* out, int B, int L,
                                    int Cin, int Cout, int k, int act, float* scratch, void* stream) {
    return vo2w_debug_tf32_conv_ex(x, w, bias, res, nullptr, out, B, L, Cin, Cout, k, 1, 1.f, 1.f, act, scratch, stream);
}
