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
// channel-major [B,D,L] pieces of the FFT blocks (decoder.cu)
int ncl_layernorm_mask(const float* x, const float* gamma, const float* beta, const uint8_t* mask, float* y, int B, int D, int L,
                       cudaStream_t st);
int ncl_attention(const float* qkv, const uint8_t* mask, float* out, int B, int L, int D, int n_head, cudaStream_t st);
