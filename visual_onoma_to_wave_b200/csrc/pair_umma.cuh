// Fused ResBlock iterations on tcgen05 (sm_100a).  One launch computes, per 128*NB-row tile,
//     y = sum over segments s of [ x_s + conv2_s(lrelu(conv1_s(lrelu(x_s)) + b1_s)) + b2_s ],  out = bf16(lrelu(scale * y))
// (scripts/hifigan/models.py:96-103 for one segment; with the last iteration of all ResBlocks of a stage as
// segments it is also the MRF sum xs of models.py:154-160).  See pair_umma.cu and DESIGN.md.
#pragma once
#include <vector>

#include "common.cuh"

// warp roles:
//   0-7   epilogue 2 (conv2 accumulators + residuals -> HBM), two warps per TMEM lane quarter (= warp & 3)
//   8-11  epilogue 1 (conv1 accumulators -> conv2 operand in shared memory)
//   12 input-tile producer, 13 weight producer + TMEM owner, 14-15 MMA issuers (leader CTA; 14 = relay in the peer CTA)
#define PAIR_THREADS 512
#define PAIR_E2_WARPS 8
#define PAIR_E1_WARP0 8
#define PAIR_E1_WARPS 4
#define PAIR_WARP_XA 12
#define PAIR_WARP_W 13
#define PAIR_WARP_TMEM 13
#define PAIR_WARP_MMA0 14
#define PAIR_MAX_SLOTS 4
#define PAIR_W_BARS 8                     // weight barriers: ring slots, or one per convolution when the weights are resident
#define PAIR_RESIDENT_MAX (64 * 1024)     // per CTA: all convolutions of a launch, copied once (C = 32: every launch; C = 64: k = 3, 7)
#define PAIR_MAX_SEG 3
// Copies of the accumulators D1 / D2 (TMEM) and of the conv2 operand buffer (shared memory).  With 2 the MMA warp
// runs a skewed schedule -- conv1 of unit n+1 is issued before conv2 of unit n -- so that the conv1 epilogue (e1)
// of a unit overlaps tensor-core work instead of stalling it.  TMEM (512 columns) has the room for C <= 64.
#define PAIR_ND(C) ((C) <= 64 ? 2 : 1)
#define PAIR_SMEM_LIMIT (226 * 1024)      // dynamic part; 227 KB per CTA minus the static barriers
// conv_post (Conv1d C -> 1, k = 7, padding 3; models.py:145,161-163) riding the last MRF launch
#define PAIR_POST_K 7
#define PAIR_POST_HALO 3
#define PAIR_POST_ROWS 512                // rows of a tile (128 * NB at C = 32)

struct PairSegP {
    const __nv_bfloat16* xa;      // the stream: bf16(leaky_relu(x, 0.1)), [B][C/8][Lp][8], zero pad rows = conv1 operand;
                                  // the residual x is recovered from it (x = s >= 0 ? s : 10 s)
    const uint8_t* w1;            // packed conv1: [cta rank][item][KB/8][C/CG][8] bf16, item 0 = bias, item 1 + kb*k + tap = weights
    const uint8_t* w2;
    int32_t k, dil;               // conv1: k taps, dilation dil; conv2: k taps, dilation 1
    int32_t h1, h2;               // halos: dil*(k-1)/2 and (k-1)/2
    int32_t n_items;              // 1 + k * (C / KB)
    int32_t xa_rows;              // 128*NB + 2*h1
    uint32_t conv_bytes_rank;     // bytes of one packed conv per CTA rank
    int32_t n_stages;             // ring stages of one conv = ceil(n_items / items_per_stage)
};

struct PairP {
    PairSegP seg[PAIR_MAX_SEG];
    int32_t n_seg;
    __nv_bfloat16* out_act;       // bf16(leaky_relu(y * scale, out_slope)): the stream / operand of the next layer
    int32_t C, L, Lp;
    int32_t h2max;
    int32_t NB;                   // 128-row accumulator blocks per CTA tile
    int32_t V;                    // valid output rows per tile = 128*NB - 2*h2max
    int32_t Vs;                   // samples a tile is responsible for (= tile stride): V, or V - 6 with conv_post on board
    int32_t post;                 // 1: conv_post + tanh ride this launch (the 32-channel MRF launch); out_act is not written
    uint32_t part_off;            // post: [2][PAIR_POST_K][PAIR_POST_ROWS] fp32 partial sums in shared memory
    float post_w[PAIR_POST_K * 32];   // conv_post weights [tap][channel], fp32
    float post_b;
    float* wav;                   // post: fp32 waveform [B][L] (optional)
    int16_t* wav_i16;             // post: (int16)(wav * i16_scale) [B][L] (optional; utils/model.py:83-86)
    float i16_scale;
    int32_t tiles_per_utt, n_tiles, iters;
    float inv_tiles_per_utt;      // 1 / tiles_per_utt (tile_of's division-free split of the tile index)
    float out_slope, scale;
    int32_t t_rows, xa_bufs;      // input tile buffers: 1 .. 4
    int32_t resident;             // 1: every convolution's weights stay in shared memory for the whole kernel (no ring)
    int32_t split;                // 1: conv1 and conv2 are issued by two threads on their own schedules (needs resident)
    int32_t rinit;                // 1: the e2 warps initialise D2 with the residual (from the staged input tile) before conv2
    uint32_t wres_off[2 * PAIR_MAX_SEG];   // resident: byte offset of (segment, conv1 / conv2) inside the weight area
    int32_t t_bufs;               // = PAIR_ND(C): copies of the conv2 operand buffer (and of D1 / D2 in TMEM)
    uint32_t t_bytes;             // bytes of one conv2 operand buffer
    int32_t KB, ksteps, kbs_log2; // input channels per item, K=16 steps per item, log2(C / KB)
    int32_t items_per_stage, n_slots, slots_log2;   // n_slots is 2 or 4
    uint32_t item_bytes, stage_bytes;
    uint32_t xa_off[4], t_off, ring_off, ones_off;   // byte offsets into dynamic shared memory
    uint32_t smem_bytes;
    int32_t pf;                   // units (tile x segment) the input rows are prefetched into L2 ahead of their bulk copy
    int32_t dbg;                  // VO2W_PAIR_DBG bit mask: what-if timing experiments (1: e2 without global traffic,
                                  // 2: e1 without conversion, 4: no MMAs); results are wrong when set
    unsigned long long* prof;     // optional [grid][16] cycle counters (VO2W_PAIR_PROF=1), else null
};

// Weights of one ResBlock iteration (convs1[m], convs2[m]) packed for the kernel.
struct PackedPair {
    int C = 0, k = 0, dil = 1, cg = 1;
    uint8_t* w1_dev = nullptr;
    uint8_t* w2_dev = nullptr;
};

// What one launch runs: 1..3 iterations whose results are summed.
struct PackedGroup {
    int n_seg = 0;
    const PackedPair* seg[PAIR_MAX_SEG] = {nullptr, nullptr, nullptr};
};

struct PairIO {
    const __nv_bfloat16* xa[PAIR_MAX_SEG] = {nullptr, nullptr, nullptr};
    __nv_bfloat16* out_act = nullptr;
    float out_slope = 0.1f;
    float scale = 1.0f;
    // conv_post + tanh on board (last stage, 32 channels, CTA pairs): weights [PAIR_POST_K][32] fp32 on the host
    const float* post_w = nullptr;
    float post_b = 0.f;
    float* wav = nullptr;
    int16_t* wav_i16 = nullptr;
    float i16_scale = 0.f;
};

uint16_t f32_to_bf16_rn(float f);   // pack.cu
// w1/w2: [C,C,k] fp32 (nn.Conv1d layout), b1/b2: [C].  cg = CTAs per tcgen05 group (1 or 2).
int pack_pair(PackedPair& P, const float* w1, const float* b1, const float* w2, const float* b2, int C, int k, int dil,
              int cg);
void free_pair(PackedPair& P);
int make_group(PackedGroup& G, const PackedPair* const* pairs, int n_seg);
void free_group(PackedGroup& G);
// host-side layout of one conv ([rank][item][chunk][n_local][8] bf16 bits, item 0 = bias as bf16 hi/lo in K
// columns 0/1); exposed for the CPU tests
void pack_pair_conv_host(std::vector<uint16_t>& out, const float* w, const float* bias, int C, int k, int cg);
// fills the geometry of a launch (no pointers); returns the grid size or a negative error
int pair_plan(PairP& p, int C, int n_seg, const int* ks, const int* dils, int cg, int B, int L, int n_sms, bool post = false);
int pair_launch(const PackedGroup& G, const PairIO& io, int B, int L, cudaStream_t st);
