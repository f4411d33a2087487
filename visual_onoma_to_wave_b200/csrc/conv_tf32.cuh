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
    // optional batched-GEMM view (attention): zero-initialised fields mean the plain layout above
    int ldx = 0;         // row stride of x in floats (0: Cin)
    int ldo = 0;         // row stride of out / res in floats (0: Cout)
    int heads = 0;       // > 1: grid z = B * heads; head h reads x columns h * x_hs.., writes out columns h * o_hs..
    int x_hs = 0, o_hs = 0;
    size_t w_zs = 0;     // packed-weight stride per (b, h) in floats (0: one weight set for every z)
    // optional vocoder fusions (fp32 mode of the Generator): zero-initialised fields mean "off"
    int dil = 0;                 // dilation (0: 1); padding dil * (k - 1) / 2
    float in_slope = 0.f;        // leaky-relu slope applied to x while staging (0: identity)
    const float* acc = nullptr;  // second addend [B,L,Cout] (MRF sum); may alias out
    float out_scale = 0.f;       // result scaled before the activation (0: 1), e.g. 1 / num_kernels
};

bool tf32_conv_supported(int Cout, int Cin, int k);
size_t tf32_conv_packed_floats(int Cout, int Cin, int k);
// w [Cout,Cin,k] (Conv1d) on the device -> packed hi/lo stages in consumption order
int tf32_conv_pack(const float* w, int Cout, int Cin, int k, float* packed, cudaStream_t st);
int tf32_conv1d(const Tf32Conv& p, cudaStream_t st);
// per (b, h): w(co, ci) = src[b * bs + h * hs + co * ld_co + ci * ld_ci] for co < co_valid, ci < ci_valid, else 0;
// packed as a k = 1 layer [Cout, Cin] at packed + (b * heads + h) * 2 * Cout * Cin
int tf32_pack_strided(const float* src, size_t bs, size_t hs, size_t ld_co, size_t ld_ci, int B, int heads, int Cout, int Cin,
                      int co_valid, int ci_valid, float* packed, cudaStream_t st);
