// fp32 CUDA-core kernels for the exact precision mode (VO2W_PREC_FP32) and for the
// per-layer cross-checks of the tcgen05 path.  Plain NCL fp32 tensors, fp32 FMA
// accumulation: results agree with the reference's fp32 CPU convolutions to ~1e-6,
// far inside the 1e-4 gate.  The throughput mode is conv_umma.cu.
//
// Fusions mirror Generator.forward / ResBlock.forward (scripts/hifigan/models.py:96-103,
// 149-165): leaky-relu on the input while staging, bias, residual add, MRF accumulate and
// the final /num_kernels, tanh.
#include "conv_simt.cuh"

namespace {

constexpr int TCO = 32;      // output channels per CTA
constexpr int TL = 128;      // output samples per CTA
constexpr int CIB = 8;       // input channels per shared-memory stage
constexpr int KMAX = 11;
constexpr int XW = TL + 64;  // staged row: TL + (k-1)*dil <= 128 + 50

__global__ void __launch_bounds__(256) conv1d_f32_kernel(SimtConv p) {
    __shared__ float xs[CIB][XW];
    __shared__ float ws[TCO][CIB][KMAX];
    const int tid = threadIdx.x, lg = tid & 31, cg = tid >> 5;
    const int b = blockIdx.z, co0 = blockIdx.y * TCO, l0 = blockIdx.x * TL;
    const int span = TL + (p.k - 1) * p.dil;
    const float* xb = p.x + (size_t)b * p.Cin * p.L;

    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[a][i] = 0.f;

    for (int ci0 = 0; ci0 < p.Cin; ci0 += CIB) {
        for (int e = tid; e < CIB * span; e += 256) {
            const int c = e / span, o = e - c * span;
            const int l = l0 - p.pad + o, ci = ci0 + c;
            float v = 0.f;
            if (ci < p.Cin && l >= 0 && l < p.L) v = lrelu(xb[(size_t)ci * p.L + l], p.in_slope);
            xs[c][o] = v;
        }
        for (int e = tid; e < TCO * CIB * p.k; e += 256) {
            const int j = e % p.k, c = (e / p.k) % CIB, o = e / (p.k * CIB);
            const int co = co0 + o, ci = ci0 + c;
            ws[o][c][j] = (co < p.Cout && ci < p.Cin) ? p.w[((size_t)co * p.Cin + ci) * p.k + j] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < CIB; ++c) {
            for (int j = 0; j < p.k; ++j) {
                float xv[4], wv[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) xv[i] = xs[c][lg + 32 * i + j * p.dil];
#pragma unroll
                for (int a = 0; a < 4; ++a) wv[a] = ws[cg * 4 + a][c][j];
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[a][i] = fmaf(wv[a], xv[i], acc[a][i]);
            }
        }
        __syncthreads();
    }

#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int co = co0 + cg * 4 + a;
        if (co >= p.Cout) continue;
        const float bias = p.bias[co];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int l = l0 + lg + 32 * i;
            if (l >= p.L) continue;
            const size_t o = ((size_t)b * p.Cout + co) * p.L + l;
            float v = acc[a][i] + bias;
            if (p.res) v += p.res[o];
            if (p.accum == 1) v = p.out[o] + v;                       // xs += resblock(x)
            else if (p.accum == 2) v = (p.out[o] + v) / p.div;         // ... ; x = xs / num_kernels
            if (p.tanh_out) v = tanhf(v);
            p.out[o] = v;
        }
    }
}

// ConvTranspose1d, out[l] = bias + sum_{j : (l+p-j) % s == 0} x[(l+p-j)/s] . w[:, :, j]
__global__ void __launch_bounds__(256) convt1d_f32_kernel(SimtConv p) {
    const int Lo = p.L * p.stride;
    const int l = blockIdx.x * 256 + threadIdx.x;
    const int co = blockIdx.y, b = blockIdx.z;
    if (l >= Lo) return;
    const float* xb = p.x + (size_t)b * p.Cin * p.L;
    float acc = p.bias[co];
    for (int j = (l + p.pad) % p.stride; j < p.k; j += p.stride) {
        const int i = (l + p.pad - j) / p.stride;
        if (i < 0 || i >= p.L) continue;
        const float* wj = p.w + (size_t)co * p.k + j;          // w[ci][co][j]
        float a = 0.f;
        for (int ci = 0; ci < p.Cin; ++ci)
            a = fmaf(lrelu(xb[(size_t)ci * p.L + i], p.in_slope), wj[(size_t)ci * p.Cout * p.k], a);
        acc += a;
    }
    p.out[((size_t)b * p.Cout + co) * Lo + l] = acc;
}

}  // namespace

int simt_conv1d(const SimtConv& p, cudaStream_t st) {
    VO2W_REQUIRE(p.k <= KMAX && (p.k - 1) * p.dil + TL <= XW, "simt conv: kernel %d dil %d unsupported", p.k, p.dil);
    dim3 grid(vo2w_ceil_div(p.L, TL), vo2w_ceil_div(p.Cout, TCO), p.B);
    conv1d_f32_kernel<<<grid, 256, 0, st>>>(p);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int simt_convt1d(const SimtConv& p, cudaStream_t st) {
    dim3 grid(vo2w_ceil_div((long long)p.L * p.stride, 256), p.Cout, p.B);
    convt1d_f32_kernel<<<grid, 256, 0, st>>>(p);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}
