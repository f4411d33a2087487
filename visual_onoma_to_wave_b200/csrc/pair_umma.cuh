// Fused ResBlock conv pair on tcgen05 (sm_100a):  r = x + conv2(lrelu(conv1(lrelu(x)) + b1)) + b2
// (scripts/hifigan/models.py:96-103, one iteration of the ResBlock loop).  See pair_umma.cu and DESIGN.md.
#pragma once
#include <vector>

#include "common.cuh"

#define PAIR_THREADS 256          // warp 0 input-tile producer, 1 MMA issuer / peer relay, 2 weight producer,
                                  // 3 TMEM owner, 4-7 epilogue (TMEM lane quarter = warp & 3)
#define PAIR_EPI_WARPS 4
#define PAIR_MAX_SLOTS 4
#define PAIR_SMEM_LIMIT (226 * 1024)      // dynamic part; 227 KB per CTA minus the static barriers

// epilogue of the second convolution
#define PAIR_OUT_STREAM 0         // out_raw = r (bf16), out_act = lrelu(r, out_slope) (bf16)
#define PAIR_OUT_SUM_FIRST 1      // S  = r                      (fp32, first ResBlock of the MRF sum)
#define PAIR_OUT_SUM_ADD 2        // S += r
#define PAIR_OUT_SUM_ONLY 4       // out_act = lrelu(r * inv_div, out_slope)         (num_kernels == 1)
#define PAIR_OUT_SUM_LAST 3       // out_act = lrelu((S + r) * inv_div, out_slope)   (x = xs / num_kernels, models.py:154-160)

struct PairP {
    const __nv_bfloat16* xa;      // conv1 operand = leaky_relu(x, 0.1) as bf16, [B][C/8][Lp][8], zero pad rows
    const __nv_bfloat16* x;       // residual stream x as raw bf16, same layout
    __nv_bfloat16* out_raw;
    __nv_bfloat16* out_act;
    float* S;                     // [B][C/4][L][4]
    const uint8_t* w1;            // packed conv1 weights: [cta rank][item = tap*kbs + kb][KB/8][C/CG][8] bf16
    const uint8_t* w2;
    const float* bias;            // [2][C]: conv1 bias, conv2 bias
    int32_t C, L, Lp;
    int32_t k, dil;               // conv1: k taps, dilation dil; conv2: k taps, dilation 1
    int32_t h1, h2;               // halos: dil*(k-1)/2 and (k-1)/2
    int32_t NB;                   // 128-row accumulator blocks per CTA tile
    int32_t V;                    // valid output rows per tile = 128*NB - 2*h2
    int32_t tiles_per_utt, n_tiles, iters;
    int32_t mode;
    float out_slope, inv_div;
    int32_t xa_rows, t_rows, xa_bufs;
    int32_t KB, ksteps, kbs_log2; // input channels per item, K=16 steps per item, log2(C / KB)
    int32_t n_items, items_per_stage, n_slots;
    uint32_t item_bytes, stage_bytes, conv_bytes_rank;
    uint32_t xa_off[2], t_off, ring_off, bias_off;   // byte offsets into dynamic shared memory
    uint32_t smem_bytes;
};

// Weights of one ResBlock conv pair packed for the kernel (device copies).
struct PackedPair {
    int C = 0, k = 0, dil = 1, cg = 1;
    uint8_t* w1_dev = nullptr;
    uint8_t* w2_dev = nullptr;
    float* bias_dev = nullptr;    // [2][C]
};

struct PairIO {
    const __nv_bfloat16* xa = nullptr;
    const __nv_bfloat16* x = nullptr;
    __nv_bfloat16* out_raw = nullptr;
    __nv_bfloat16* out_act = nullptr;
    float* S = nullptr;
    int mode = PAIR_OUT_STREAM;
    float out_slope = 0.1f;
    float inv_div = 1.0f;
};

uint16_t f32_to_bf16_rn(float f);   // pack.cu
// w1/w2: [C,C,k] fp32 (nn.Conv1d layout), b1/b2: [C].  cg = CTAs per cluster (1 or 2).
int pack_pair(PackedPair& P, const float* w1, const float* b1, const float* w2, const float* b2, int C, int k, int dil,
              int cg);
void free_pair(PackedPair& P);
// host-side layout of one conv ([rank][item][chunk][n_local][8] bf16 bits); exposed for the CPU tests
void pack_pair_conv_host(std::vector<uint16_t>& out, const float* w, int C, int k, int cg);
int pair_plan(PairP& p, int C, int k, int dil, int cg, int B, int L, int n_sms);
int pair_launch(const PackedPair& P, const PairIO& io, int B, int L, cudaStream_t st);
