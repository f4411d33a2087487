// fp32 CUDA-core convolution launches (exact mode).  See conv_simt.cu.
#pragma once
#include "common.cuh"

struct SimtConv {
    const float* x;      // [B,Cin,L]
    const float* w;      // Conv1d [Cout,Cin,k]; ConvTranspose1d [Cin,Cout,k]
    const float* bias;   // [Cout]
    const float* res;    // optional residual [B,Cout,L]
    float* out;          // [B,Cout,L] (conv) or [B,Cout,L*stride] (transposed)
    int B, Cin, Cout, L, k, dil, pad, stride;
    float in_slope;      // leaky-relu slope applied to x while staging (1 = identity)
    int accum;           // 0 store, 1 out += v, 2 out = (out + v) / div
    float div;
    int tanh_out;
};

int simt_conv1d(const SimtConv& p, cudaStream_t st);
int simt_convt1d(const SimtConv& p, cudaStream_t st);
