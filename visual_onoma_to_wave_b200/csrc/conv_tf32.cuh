// fp32-accurate convolutions on the tensor pipe: tcgen05.mma kind::tf32 with a 3-term operand split
// (a = a_hi + a_lo, both exactly representable in TF32; a.w ~= a_hi.w_hi + a_lo.w_hi + a_hi.w_lo, fp32 accumulation
// in TMEM).  Time-major activations [B,L,C].  See conv_tf32.cu.
#pragma once
#include "common.cuh"

struct Tf32Conv {
    const float* x;      // [B,L,Cin]
    const float* wp;     // packed by tf32_conv_pack
    const float* bias;   // [Cout] or null
    const float* res;    // [B,L,Cout] added before the activation, or null
    float* out;          // [B,L,Cout]
    int B, L, Cin, Cout, k;     // "same" padding (k - 1) / 2, dilation 1, stride 1
    int act;             // 0 none, 1 relu, 2 tanh
};

bool tf32_conv_supported(int Cout, int Cin, int k);
size_t tf32_conv_packed_floats(int Cout, int Cin, int k);
// w [Cout,Cin,k] (Conv1d) on the device -> packed hi/lo stages in consumption order
int tf32_conv_pack(const float* w, int Cout, int Cin, int k, float* packed, cudaStream_t st);
int tf32_conv1d(const Tf32Conv& p, cudaStream_t st);
