// Persistent single-layer implicit-GEMM convolution on tcgen05 CTA pairs (sm_100a): conv_pre and the four
// ConvTranspose1d upsamplers of the HiFi-GAN Generator (scripts/hifigan/models.py:118-135,150-153).
// See stream_umma.cu and DESIGN.md section 4.3.
#pragma once
#include <vector>

#include "common.cuh"

// warps: 0 input-tile producer, 1 weight producer, 2 MMA issuer (leader CTA) / barrier relay (peer CTA), 3 TMEM owner,
//        4-11 epilogue (two warps per TMEM lane quarter)
#define SU_THREADS 384
#define SU_EPI_WARP0 4
#define SU_EPI_WARPS 8
#define SU_SLOTS 8                       // weight ring stages (at most; 4 or 8 are used)
#define SU_A_BUFS 6                      // input tile buffers (at most; as many as shared memory holds)
#define SU_MAX_ITEMS 12
#define SU_MAX_TILES 16
#define SU_SMEM_LIMIT (226 * 1024)

// One layer packed for the kernel.  The GEMM: M = input time (128-row blocks), K = input channels x taps,
// N = output channels (Conv1d) or phase x output channels (ConvTranspose1d, polyphase: out[s q + ph] is column
// (ph, co) of row q).  N is cut into tiles of n_cols accumulator columns; every "item" of a tile is one input time
// shift d (a tap of the convolution, or one of the 2-3 input samples a phase group depends on) over all n_cols.
struct SuLayer {
    int Cin = 0;               // padded to a multiple of 16 (= channels of the operand buffer)
    int Cout = 0;              // channels of the output tensor
    int up_s = 1;              // 1 = Conv1d; s = ConvTranspose1d with stride s
    int KB = 0, n_kb = 0;      // input channels per weight stage, stages per item
    int NB = 1;                // 128-row blocks per CTA tile
    int n_cols = 0, n_tiles = 0, n_items = 0;
    int ph_tile = 1, cb = 0, cblks = 1;  // a tile's columns = ph_tile phases x cb channels (phase-major); cblks = Cout / cb;
                                         // tile nt = phase group nt / cblks, channel block nt % cblks
    int halo_l = 0, halo_r = 0;
    int item_shift[SU_MAX_ITEMS] = {};   // staged-tile row of item i for output row 0 (before the tile's own shift)
    int tile_shift[SU_MAX_TILES] = {};
    uint32_t stage_bytes = 0;            // one (item, channel block) of one CTA rank: (n_cols / 2) x KB bf16
    std::vector<uint16_t> w_host;        // [rank][tile][item][kb][KB/8][n_cols/2][8] bf16 bits
    std::vector<float> bias_host;        // [n_tiles * n_cols]
    uint8_t* w_dev = nullptr;
    float* bias_dev = nullptr;
};

struct SuP {
    const __nv_bfloat16* in;   // operand buffer [B][Cin/8][Lp_in][8] (zero pad rows = the layer's padding)
    const uint8_t* w;
    const float* bias;
    __nv_bfloat16* out;        // optional: the HBM stream bf16(leaky_relu(y, out_slope)), [B][Cout/8][Lp_out][8]
    float* out_f32;            // optional (probe): y as fp32 [B][Cout/4][L_out][4]
    float out_slope;
    int32_t chunks;            // Cin / 8
    int32_t KB8, n_kb, ksteps; // KB / 8, stages per item, K=16 steps per stage
    int32_t NB, a_rows, halo_l;
    int32_t n_tiles, n_cols, n_groups, tiles_per_group, n_items;
    int32_t ph_tile, cb, cblks;
    int32_t item_shift[SU_MAX_ITEMS];
    int32_t tile_shift[SU_MAX_TILES];
    int32_t L_in, Lp_in, L_out, Lp_out, Cout, up_s;
    int32_t m_tiles, bm_tiles, pair_m, n_units;
    uint32_t stage_bytes, tile_w_bytes, rank_w_bytes;
    int32_t a_bufs, pf_extra;
    int32_t n_slots, slots_log2;   // weight ring stages in use (4 .. 8; slots_log2 unused)
    uint32_t a_off[SU_A_BUFS], ring_off, smem_bytes;
    int32_t resident;          // the weights of a unit fit the ring and every unit uses the same ones: loaded once
    int32_t dbg;               // what-if timing switches, only settable in -DVO2W_EXPERIMENTS builds (results are wrong)
};

// w [Cout,Cin,k] (nn.Conv1d, dilation dil, "same" padding dil*(k-1)/2)
int su_pack_conv(SuLayer& L, const float* w, const float* bias, int Cout, int Cin, int k, int dil);
// w [Cin,Cout,k] (nn.ConvTranspose1d, stride s, padding (k-s)/2, output_padding 0: L_out = s * L_in)
int su_pack_convT(SuLayer& L, const float* w, const float* bias, int Cin, int Cout, int k, int s);
int su_upload(SuLayer& L);
void su_free(SuLayer& L);
// geometry of one launch (no pointers); returns the grid size or a negative error
int su_plan(SuP& p, const SuLayer& L, int B, int L_in, int n_sms);
// out = bf16(leaky_relu(layer(in), out_slope)) as an operand buffer; out_f32 (optional) = layer(in) in fp32 chunks
int su_launch(const SuLayer& L, const __nv_bfloat16* in, __nv_bfloat16* out, float* out_f32, float out_slope, int B, int L_in,
              cudaStream_t st);
