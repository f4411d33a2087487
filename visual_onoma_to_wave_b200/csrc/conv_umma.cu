// tcgen05 / TMEM implicit-GEMM convolution kernel (the throughput path, VO2W_PREC_BF16).
//
// One CTA computes a [128 time samples] x [n_cols] tile of one utterance:
//   D[q, n] = sum over taps j, input channels ci of  X[l0 + q + shift_j, ci] * W_j[n, ci]
// - time is the UMMA M dimension (TMEM lane = sample), output channels (x phase for the
//   transposed convs) the N dimension, input channels the K dimension;
// - the input tile (128 + halo rows, all Cin) is staged ONCE in shared memory by 1-D bulk
//   async copies (one per 16-byte channel chunk; the operand buffer layout makes each a
//   contiguous span).  A tap is the same tile read at a different row offset: with the
//   SWIZZLE_NONE K-major core-matrix layout a row offset is a 16-byte multiple added to the
//   descriptor start address, so dilation halos cost no extra copies;
// - weights stream through a 4-stage mbarrier ring, pre-packed on the host in the same
//   core-matrix order (one bulk copy per stage);
// - accumulators live in TMEM; 4 epilogue warps read them back with tcgen05.ld and fuse
//   bias, residual add(s), the MRF average, leaky-relu for the next layer, bf16 packing,
//   or tanh (conv_post).
// Warp roles: warp 0 lane 0 = copy producer, warp 1 lane 0 = MMA issuer, warps 2-5 = epilogue.
//
// Reference semantics: scripts/hifigan/models.py:96-103 (ResBlock), :149-165 (Generator).
#include "conv_umma.cuh"

namespace {

__global__ void __launch_bounds__(UMMA_THREADS) umma_conv_kernel(const __grid_constant__ UConv p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[2 * UMMA_STAGES + 3];
    __shared__ uint32_t tmem_base_s;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int l0 = blockIdx.x * VO2W_TILE_M * p.n_blk, ty = blockIdx.y, b = blockIdx.z;   // n_blk row blocks of 128 per CTA

    const uint32_t a_smem = smem_u32(smem);
    const uint32_t b_smem = a_smem + p.a_bytes;
    const uint32_t bar0 = smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (UMMA_STAGES + s); };
    const uint32_t a_full = bar0 + 8u * (2 * UMMA_STAGES);
    const uint32_t a_empty = a_full + 8u;
    const uint32_t acc_full = a_full + 16u;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2 * UMMA_STAGES + 3; ++i) mbar_init(bar0 + 8u * i, 1);
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc(smem_u32(&tmem_base_s), (uint32_t)p.tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    const int KB = 16 * p.kb_steps;   // input channels per weight stage

    if (warp == 0) {
        // ===================== copy producer =====================
        if (lane == 0) {
            int it = 0;
            for (int s = 0; s < p.n_seg; ++s) {
                const USeg& sg = p.seg[s];
                if (s > 0) mbar_wait(a_empty, (uint32_t)((s - 1) & 1));
                const int start = VO2W_PAD_ROWS + l0 - sg.halo_l;         // first staged row inside the padded axis
                const uint32_t row_bytes = (uint32_t)min(sg.rows, p.Lp_in - start) * 16u;   // the last tile may be short
                const uint32_t chunk_stride = (uint32_t)sg.rows * 16u;
                const int chunks = p.Cin >> 3;
                mbar_arrive_expect_tx(a_full, row_bytes * (uint32_t)chunks);
                const __nv_bfloat16* src = sg.in + ((size_t)b * chunks * p.Lp_in + start) * 8;
                for (int c = 0; c < chunks; ++c)
                    bulk_g2s(a_smem + (uint32_t)c * chunk_stride, src + (size_t)c * p.Lp_in * 8, row_bytes, a_full);
                const uint8_t* wt = (const uint8_t*)sg.w + (size_t)ty * sg.w_tile_stride;
                for (int i = 0; i < sg.n_items; ++i) {
                    const uint32_t bytes = (uint32_t)sg.items[i].n_size * (uint32_t)KB * 2u;
                    const uint8_t* wi = wt + sg.items[i].w_off;
                    for (int kb = 0; kb < p.n_kblocks; ++kb, ++it) {
                        const int stage = it % UMMA_STAGES;
                        const uint32_t round = (uint32_t)(it / UMMA_STAGES);
                        mbar_wait(empty_bar(stage), (round & 1u) ^ 1u);
                        mbar_arrive_expect_tx(full_bar(stage), bytes);
                        bulk_g2s(b_smem + (uint32_t)stage * p.stage_bytes, wi + (size_t)kb * bytes, bytes,
                                 full_bar(stage));
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: the whole warp walks the loop, one elected lane issues ==========
        // (a lane-0 branch makes ptxas wrap every tcgen05.mma in a per-lane uniformisation loop, which
        //  roughly doubles the issue cost; profiles/r01_umma_issue_rates.txt)
        {
            int it = 0;
            // transposed conv: phases r >= s/2 read x[q+1], x[q] instead of x[q], x[q-1]
            const uint32_t extra_shift = p.up_s > 1 ? (uint32_t)((ty * p.phases_per_tile + p.up_pad) / p.up_s) : 0u;
            for (int s = 0; s < p.n_seg; ++s) {
                const USeg& sg = p.seg[s];
                mbar_wait(a_full, (uint32_t)(s & 1));
                tc_fence_after();
                const uint32_t a_lbo = (uint32_t)sg.rows * 16u;
                for (int i = 0; i < sg.n_items; ++i) {
                    const UItem item = sg.items[i];
                    const uint32_t idesc = umma_idesc_bf16(VO2W_TILE_M, (uint32_t)item.n_size);
                    const uint32_t b_lbo = (uint32_t)item.n_size * 16u;
                    const uint32_t d_tmem = tmem_base + (uint32_t)item.n_off;
                    for (int kb = 0; kb < p.n_kblocks; ++kb, ++it) {
                        const int stage = it % UMMA_STAGES;
                        const uint32_t round = (uint32_t)(it / UMMA_STAGES);
                        mbar_wait(full_bar(stage), round & 1u);
                        tc_fence_after();
                        if (elect_one_sync()) {
                            const uint32_t bs = b_smem + (uint32_t)stage * p.stage_bytes;
                            for (int blk = 0; blk < p.n_blk; ++blk)      // every weight stage serves all row blocks
                                for (int st = 0; st < p.kb_steps; ++st) {
                                    const uint32_t kstep = (uint32_t)(kb * p.kb_steps + st);
                                    const uint32_t a_addr = a_smem + (kstep * 2u * (uint32_t)sg.rows + (uint32_t)(blk * VO2W_TILE_M) +
                                                                      (uint32_t)item.row_shift + extra_shift) * 16u;
                                    const uint32_t b_addr = bs + (uint32_t)st * 2u * b_lbo;
                                    const uint32_t accumulate = (item.zero_init && kb == 0 && st == 0) ? 0u : 1u;
                                    umma_bf16(d_tmem + (uint32_t)(blk * p.n_cols), umma_smem_desc(a_addr, a_lbo, 128u),
                                              umma_smem_desc(b_addr, b_lbo, 128u), idesc, accumulate);
                                }
                            umma_commit(empty_bar(stage));   // stage reusable once these MMAs retire
                        }
                        __syncwarp();
                    }
                }
                if (elect_one_sync()) umma_commit(a_empty);  // input tile reusable
                __syncwarp();
            }
            if (elect_one_sync()) umma_commit(acc_full);
            __syncwarp();
        }
    } else {
        // ===================== epilogue: TMEM -> registers -> global =====================
        const int quarter = warp & 3;                    // TMEM lanes this warp may read
        const int q = quarter * 32 + lane;               // row inside a 128-row block
        mbar_wait(acc_full, 0);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const float* bias = p.bias + (size_t)ty * p.n_cols;
        // 8 accumulator columns [n, n + 8) of output row `row`: bias, residuals, activation, store
        auto emit8 = [&](const uint32_t (&r)[8], int n, int row) {
            int co, lo;
            if (p.up_s == 1) {
                co = ty * p.n_cols + n;
                lo = row;
            } else {
                const int ph = n / p.Cout;
                co = n - ph * p.Cout;
                lo = p.up_s * row + ty * p.phases_per_tile + ph;
            }
            float v[8];
            const float4 b0 = __ldg((const float4*)(bias + n)), b1 = __ldg((const float4*)(bias + n + 4));
            v[0] = __uint_as_float(r[0]) + b0.x; v[1] = __uint_as_float(r[1]) + b0.y;
            v[2] = __uint_as_float(r[2]) + b0.z; v[3] = __uint_as_float(r[3]) + b0.w;
            v[4] = __uint_as_float(r[4]) + b1.x; v[5] = __uint_as_float(r[5]) + b1.y;
            v[6] = __uint_as_float(r[6]) + b1.z; v[7] = __uint_as_float(r[7]) + b1.w;
            if (p.out_wave) {
                if (n == 0) p.out_wave[(size_t)b * p.L_out + lo] = tanhf(v[0]);
                return;
            }
            const size_t f_off = (((size_t)b * (p.Cout >> 2) + (co >> 2)) * p.L_out + lo) * 4;   // chunk co/4
            const size_t f_step = (size_t)p.L_out * 4;                                           // chunk co/4 + 1
            for (int k = 0; k < p.n_res; ++k) {
                const float4 x0 = __ldg((const float4*)(p.res[k] + f_off));
                const float4 x1 = __ldg((const float4*)(p.res[k] + f_off + f_step));
                v[0] += x0.x; v[1] += x0.y; v[2] += x0.z; v[3] += x0.w;
                v[4] += x1.x; v[5] += x1.y; v[6] += x1.z; v[7] += x1.w;
            }
            if (p.div != 1.0f) {
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = v[i] / p.div;
            }
            if (p.out_f32) {
                *(float4*)(p.out_f32 + f_off) = make_float4(v[0], v[1], v[2], v[3]);
                *(float4*)(p.out_f32 + f_off + f_step) = make_float4(v[4], v[5], v[6], v[7]);
            }
            if (p.out_bf16) {
                // HBM stream format of the bf16 mode: bf16(leaky_relu(value)) = the next layer's MMA operand; a
                // consumer that needs the value itself inverts the activation (pair_umma.cu, e2)
                __nv_bfloat162 h[4];
                const size_t o = (((size_t)b * (p.Cout >> 3) + (co >> 3)) * p.Lp_out + VO2W_PAD_ROWS + lo) * 8;
#pragma unroll
                for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(lrelu(v[2 * i], p.out_slope), lrelu(v[2 * i + 1], p.out_slope));
                *(uint4*)(p.out_bf16 + o) = *(const uint4*)h;
            }
        };
        for (int blk = 0; blk < p.n_blk; ++blk) {
            const int row = l0 + blk * VO2W_TILE_M + q;
            const bool valid = row < p.M_valid;
            const uint32_t t_blk = t_row + (uint32_t)(blk * p.n_cols);
            for (int g = 0; g < p.n_cols; g += 8) {
                uint32_t r[8];
                __syncwarp();                                // tcgen05.ld is warp-collective (.sync.aligned)
                tmem_ld8(t_blk + (uint32_t)g, r);
                tmem_ld_wait();
                if (valid) emit8(r, g, row);
                if (p.out_wave) break;                       // conv_post: only column 0 exists
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
}

// fp32 NCL -> bf16 operand buffer [B][C/8][Lp][8] (pad rows untouched), leaky-relu optional.
__global__ void pack_ncl_to_operand_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int C, int L,
                                           int Lp, float slope) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int chunk = blockIdx.y, b = blockIdx.z;
    if (l >= L) return;
    __nv_bfloat162 h[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = chunk * 8 + 2 * i;
        const float a = c < C ? x[((size_t)b * C + c) * L + l] : 0.f;
        const float d = c + 1 < C ? x[((size_t)b * C + c + 1) * L + l] : 0.f;
        h[i] = __floats2bfloat162_rn(lrelu(a, slope), lrelu(d, slope));
    }
    *(uint4*)(out + (((size_t)b * gridDim.y + chunk) * Lp + VO2W_PAD_ROWS + l) * 8) = *(const uint4*)h;
}

// fp32 chunked [B][Cp/4][L][4] -> NCL [B,C,L] (debug taps / layer probe)
__global__ void unchunk_f32_kernel(const float* __restrict__ x, float* __restrict__ out, int Cp, int C, int L) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y, b = blockIdx.z;
    if (l >= L) return;
    out[((size_t)b * C + c) * L + l] = x[(((size_t)b * (Cp >> 2) + (c >> 2)) * L + l) * 4 + (c & 3)];
}

// NCL [B,C,L] -> fp32 chunked [B][Cp/4][L][4] (layer probe residual)
__global__ void chunk_f32_kernel(const float* __restrict__ x, float* __restrict__ out, int C, int Cp, int L) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y, b = blockIdx.z;
    if (l >= L) return;
    out[(((size_t)b * (Cp >> 2) + (c >> 2)) * L + l) * 4 + (c & 3)] = c < C ? x[((size_t)b * C + c) * L + l] : 0.f;
}

}  // namespace

size_t umma_conv_smem(const UConv& p) { return (size_t)p.a_bytes + (size_t)UMMA_STAGES * p.stage_bytes; }

int umma_conv_launch(const UConv& p, int m_tiles, int n_tiles, int B, cudaStream_t st) {
    static bool attr_set[VO2W_MAX_DEV];
    const int cur = vo2w_cur_dev();
    if (!attr_set[cur]) {
        VO2W_CUDA_OK(cudaFuncSetAttribute(umma_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, UMMA_SMEM_LIMIT));
        attr_set[cur] = true;
    }
    const size_t smem = umma_conv_smem(p);
    VO2W_REQUIRE(smem <= UMMA_SMEM_LIMIT, "umma conv: %zu bytes of shared memory exceed the limit", smem);
    VO2W_REQUIRE(B <= 65535 && n_tiles <= 65535, "umma conv: grid too large");
    dim3 grid((unsigned)m_tiles, (unsigned)n_tiles, (unsigned)B);
    umma_conv_kernel<<<grid, UMMA_THREADS, smem, st>>>(p);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int pack_ncl_to_operand(const float* x, __nv_bfloat16* out, int B, int C, int L, int Lp, float slope, cudaStream_t st) {
    dim3 grid((unsigned)vo2w_ceil_div(L, 128), (unsigned)((C + 7) / 8), (unsigned)B);
    pack_ncl_to_operand_kernel<<<grid, 128, 0, st>>>(x, out, C, L, Lp, slope);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int unchunk_f32(const float* x, float* out, int B, int Cp, int C, int L, cudaStream_t st) {
    dim3 grid((unsigned)vo2w_ceil_div(L, 128), (unsigned)C, (unsigned)B);
    unchunk_f32_kernel<<<grid, 128, 0, st>>>(x, out, Cp, C, L);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int chunk_f32(const float* x, float* out, int B, int C, int Cp, int L, cudaStream_t st) {
    dim3 grid((unsigned)vo2w_ceil_div(L, 128), (unsigned)Cp, (unsigned)B);
    chunk_f32_kernel<<<grid, 128, 0, st>>>(x, out, C, Cp, L);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}
