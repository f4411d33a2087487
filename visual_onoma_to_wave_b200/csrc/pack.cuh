// Host-side weight packing and launch assembly for the tcgen05 convolution kernel.
#pragma once
#include <vector>

#include "conv_umma.cuh"

// One reference layer (nn.Conv1d or nn.ConvTranspose1d) packed for umma_conv_kernel.
// Packed weights per N-tile: [item][Cin/8][n_size][8] bf16, i.e. for every tap the B operand
// W_tap[n][ci] in the K-major SWIZZLE_NONE core-matrix order (8 input channels = 16 bytes
// contiguous, then the n index, then the next channel chunk).  A weight stage of the kernel's
// ring is any run of 2*kb_steps consecutive channel chunks of one item.
struct PackedLayer {
    int Cin = 0;              // padded to a multiple of 16
    int Cout = 0;             // true output channels
    int k = 0, dil = 1;
    int up_s = 1;             // > 1: ConvTranspose1d with stride up_s, kernel 2*up_s, padding up_s/2
    int n_cols = 0;           // accumulator columns per N-tile
    int n_tiles = 0;
    int phases_per_tile = 1;
    int halo_l = 0, halo_r = 0;
    int n_items = 0;
    UItem items[UMMA_MAX_ITEMS];
    uint32_t w_tile_stride = 0;
    std::vector<uint16_t> w_host;      // bf16 bits, all tiles
    std::vector<float> bias_host;      // [n_tiles][n_cols]
    __nv_bfloat16* w_dev = nullptr;
    float* bias_dev = nullptr;
};

uint16_t f32_to_bf16_rn(float f);

// w [Cout,Cin,k] (nn.Conv1d), dilation dil, "same" padding dil*(k-1)/2
int pack_conv(PackedLayer& L, const float* w, const float* bias, int Cout, int Cin, int k, int dil);
// w [Cin,Cout,k] (nn.ConvTranspose1d), stride s, k == 2s, padding s/2  (polyphase, 2 taps per phase)
int pack_convT(PackedLayer& L, const float* w, const float* bias, int Cin, int Cout, int k, int s);
int upload_layer(PackedLayer& L);
void free_layer(PackedLayer& L);

struct EpilogueSpec {
    const float* bias = nullptr;       // override (fused MRF launch); default = layer bias
    const float* res[3] = {nullptr, nullptr, nullptr};
    int n_res = 0;
    float div = 1.0f;
    float* out_f32 = nullptr;
    __nv_bfloat16* out_bf16 = nullptr;
    float out_slope = 1.0f;
    float* out_wave = nullptr;
};

// Assemble and launch: `layers[i]` reads operand buffer `ins[i]`; all accumulate into one tile.
// L_in = length of the input time axis (accumulator rows).
int launch_layers(const PackedLayer* const* layers, const __nv_bfloat16* const* ins, int n_seg, int B, int L_in,
                  const EpilogueSpec& ep, cudaStream_t st);
