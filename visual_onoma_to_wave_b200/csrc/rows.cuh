// fp32 helpers on time-major rows [B*L][D] shared by the acoustic-model tail (postnet.cu) and the FFT blocks
// (decoder.cu).  See rows.cu / decoder.cu.
#pragma once
#include "common.cuh"

// [B,R,C] -> [B,C,R]; optional `add` [B,C,R] is added to the result
int rows_transpose_add(const float* in, const float* add, float* out, int B, int R, int C, cudaStream_t st);
// nn.Linear on the last axis: y[r,:] = W x[r,:] + b   (x [rows,K], w [N,K], bias [N] or null, y [rows,N])
int rows_linear(const float* x, const float* w, const float* bias, long long rows, int K, int N, float* y, cudaStream_t st);
// y[b,l,:] = x[b,l,:] + table[l,:]
int rows_add_table(const float* x, const float* table, float* y, int B, int L, int D, cudaStream_t st);
// time-major pieces of the FFT blocks (decoder.cu)
// y[r,:] = mask[r] ? 0 : LayerNorm(x[r,:]) * gamma + beta
int rows_layernorm_mask(const float* x, const float* gamma, const float* beta, const uint8_t* mask, float* y, long long rows, int D,
                        cudaStream_t st);
// qkv [B*L][3D] -> out [B*L][D]: softmax(q k^T / sqrt(128), key padding mask) v per head of 128 columns
int rows_attention(const float* qkv, const uint8_t* mask, float* out, int B, int L, int D, int n_head, cudaStream_t st);
