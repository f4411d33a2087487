// tcgen05 / TMEM implicit-GEMM convolution for sm_100a.  See conv_umma.cu and DESIGN.md.
#pragma once
#include "common.cuh"

#define UMMA_MAX_ITEMS 12
#define UMMA_MAX_SEGS 3
#define UMMA_STAGES 4
#define UMMA_THREADS 192          // warp 0 copy producer, warp 1 MMA issuer, warps 2-5 epilogue
#define UMMA_SMEM_LIMIT (224 * 1024)

// One product  D[:, n_off : n_off+n_size] (+)= A[row_shift : row_shift+128, :] . W_item^T
// i.e. one convolution tap (or one polyphase tap of a transposed convolution).
struct UItem {
    int32_t row_shift;   // first row of the staged input tile this tap reads
    int32_t n_off;       // first accumulator column
    int32_t n_size;      // accumulator columns (multiple of 16, <= 256)
    int32_t zero_init;   // 1: first product into this column range (overwrite, don't accumulate)
    uint32_t w_off;      // byte offset of the item's packed weights inside one N-tile
};

// One input tensor with its taps.  Several segments accumulate into the same TMEM tile:
// the last conv of the three MRF ResBlocks is ONE launch (xs = rb0 + rb1 + rb2, models.py:154-160).
struct USeg {
    const __nv_bfloat16* in;   // operand buffer [B][Cin/8][Lp_in][8]
    const __nv_bfloat16* w;    // packed weights [n_tiles][item][kblock][KB/8][n_size][8]
    uint32_t w_tile_stride;    // bytes between N-tiles
    int32_t n_items;
    int32_t halo_l;            // staged tile starts at sample l0 - halo_l
    int32_t rows;              // staged rows = 128 + halo_l + halo_r
    UItem items[UMMA_MAX_ITEMS];
};

struct UConv {
    USeg seg[UMMA_MAX_SEGS];
    int32_t n_seg;
    int32_t Cin;               // multiple of 16
    int32_t kb_steps;          // K=16 MMA steps per weight stage
    int32_t n_kblocks;         // Cin / (16 * kb_steps)
    int32_t Lp_in;             // padded length of the operand buffers
    int32_t M_valid;           // accumulator rows that exist (input time axis length)
    int32_t n_blk;             // 128-row blocks per CTA (they share every weight stage); block b accumulates at column b * n_cols
    int32_t n_cols;            // accumulator columns of one N-tile
    int32_t tmem_cols;         // power of two >= max(32, n_blk * n_cols)
    uint32_t a_bytes;          // shared memory reserved for the staged input tile
    uint32_t stage_bytes;      // shared memory per weight stage
    // ---- epilogue ----
    const float* bias;         // [n_tiles][n_cols]
    const float* res[3];       // fp32 residual streams [B][Cout/4][L_out][4], summed
    int32_t n_res;
    float div;                 // 1 or num_kernels (x = xs / 3)
    float* out_f32;            // optional [B][Cout/4][L_out][4]
    __nv_bfloat16* out_bf16;   // optional operand buffer of the next layer: bf16(leaky_relu(value, out_slope)), the HBM stream
    float out_slope;
    float* out_wave;           // optional [B][L_out]: tanh(column 0)   (conv_post)
    int32_t Cout;              // channels of the output tensor
    int32_t up_s;              // 1 = convolution; s = transposed convolution (out row = s*q + phase)
    int32_t phases_per_tile;   // phases covered by one N-tile (up_s > 1)
    int32_t up_pad;            // transposed-conv padding; tile ty adds (ty*phases_per_tile + up_pad) / up_s rows
    int32_t L_out;
    int32_t Lp_out;
};

int umma_conv_launch(const UConv& p, int m_tiles, int n_tiles, int B, cudaStream_t st);
size_t umma_conv_smem(const UConv& p);
int pack_ncl_to_operand(const float* x, __nv_bfloat16* out, int B, int C, int L, int Lp, float slope, cudaStream_t st);
// inv_slope != 1: undo the stream's leaky-relu (negative values are multiplied by inv_slope)
int unchunk_bf16(const __nv_bfloat16* x, float* out, int B, int C, int L, int Lp, cudaStream_t st, float inv_slope = 1.0f);
int unchunk_f32(const float* x, float* out, int B, int Cp, int C, int L, cudaStream_t st);
int chunk_f32(const float* x, float* out, int B, int C, int Cp, int L, cudaStream_t st);
