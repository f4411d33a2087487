// VarianceAdaptor pieces in front of LengthRegulator.
//
//   VariancePredictor.forward (scripts/model/modules.py:161-213, eval):  x [B,T,d_in] ->
//       LayerNorm(relu(Conv1d k3(x))) -> LayerNorm(relu(Conv1d k3(.))) -> Linear(F,1) -> squeeze -> masked_fill(mask, 0)
//   get_energy_embedding / get_kurtosis_embedding (modules.py:53-78):  prediction de-normalised, scaled by the control,
//       re-normalised; x + embedding[bucketize(value, bins)]
//   duration rounding (modules.py:110-113):  clamp(round(exp(log_d) - 1) * d_control, min=0)
//
// The predictor's two convolutions run on the tensor pipe with fp32-level accuracy (conv_tf32.cu), time-major, with the
// utterances laid end to end (frame_rows_kernel); shapes that kernel does not tile take the fp32 CUDA-core path
// (channel-major, conv_simt.cu, relu folded into the LayerNorm kernel's load).  The affine steps around the
// bucket search use explicit round-to-nearest multiplies / adds / divides (no FMA contraction) so that a value sits on
// the same side of a bin edge as in the reference's fp32 arithmetic.
#include <mutex>
#include <vector>

#include "conv_simt.cuh"
#include "conv_tf32.cuh"
#include "rows.cuh"

struct vo2w_varpred {
    int d_in = 0, d_filter = 0, k = 0;
    float* w_dev = nullptr;
    float* p_dev = nullptr;      // both convolutions packed for tf32_conv_kernel (null: shapes it does not tile -> CUDA cores)
    size_t p2 = 0;               // offset of the second convolution inside p_dev
    size_t w1, b1, g1, be1, w2, b2, g2, be2, wl, bl;
};

namespace {

size_t var_align(size_t v) { return (v + 255) / 256 * 256; }

// out[b,l] = mask[b,l] ? 0 : bias + sum_c w[c] x[b,c,l]          (Linear(F,1).squeeze(-1).masked_fill(mask, 0))
__global__ void __launch_bounds__(256) project_mask_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                           const float* __restrict__ bias, const uint8_t* __restrict__ mask,
                                                           float* __restrict__ out, int C, int L) {
    __shared__ float red[8][33];
    const int b = blockIdx.y, tl = threadIdx.x & 31, g = threadIdx.x >> 5, l = blockIdx.x * 32 + tl;
    const float* xb = x + (size_t)b * C * L;
    float acc = 0.f;
    if (l < L) for (int c = g; c < C; c += 8) acc = fmaf(w[c], xb[(size_t)c * L + l], acc);
    red[g][tl] = acc;
    __syncthreads();
    if (g == 0 && l < L) {
        float v = bias[0];
#pragma unroll
        for (int i = 0; i < 8; ++i) v += red[i][tl];
        out[(size_t)b * L + l] = (mask && mask[(size_t)b * L + l]) ? 0.f : v;
    }
}

// Tensor-pipe path: the B utterances are laid end to end as ONE sequence of B * (T + 2) rows, every utterance framed by a
// zero row on either side (= the convolutions' zero padding), so that the 128-row tiles of tf32_conv_kernel are full
// whatever T is (the reference's T_src is ~32).  xp[b * (T + 2) + 1 + t] = x[b, t]; pad[r] = 1 on the framing rows.
__global__ void __launch_bounds__(256) frame_rows_kernel(const float* __restrict__ x, float* __restrict__ xp,
                                                         uint8_t* __restrict__ pad, int T, int C4, long long rows_p) {
    const long long r = (long long)blockIdx.x * 4 + (threadIdx.x >> 6);       // 64 threads per row
    if (r >= rows_p) return;
    const int t = (int)(r % (T + 2)) - 1;
    const long long b = r / (T + 2);
    const bool framing = t < 0 || t >= T;
    float4* o = reinterpret_cast<float4*>(xp) + (size_t)r * C4;
    const float4* i = reinterpret_cast<const float4*>(x) + (size_t)(b * T + (framing ? 0 : t)) * C4;
    for (int c = threadIdx.x & 63; c < C4; c += 64) o[c] = framing ? make_float4(0.f, 0.f, 0.f, 0.f) : i[c];
    if ((threadIdx.x & 63) == 0) pad[r] = framing ? 1 : 0;
}

// one warp per row (b,t); x is the framed sequence [B * (T + 2), C]
__global__ void __launch_bounds__(256) project_mask_rows_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                                const float* __restrict__ bias, const uint8_t* __restrict__ mask,
                                                                float* __restrict__ out, long long rows, int T, int C) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    const float* xr = x + (size_t)(r + 2 * (r / T) + 1) * C;
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(w[c], xr[c], acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[r] = (mask && mask[r]) ? 0.f : acc + bias[0];
}

// one warp per (b,t) row: bucket search on lane 0's value, then the embedding row is added to x
__global__ void __launch_bounds__(256) variance_embed_kernel(const float* __restrict__ x, const float* __restrict__ value,
                                                             const float* __restrict__ bins, int n_edges,
                                                             const float* __restrict__ table, float mean, float stdv, float control,
                                                             int denorm, float* __restrict__ value_out, float* __restrict__ out,
                                                             long long rows, int D) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    float v = value[r];
    if (denorm) {
        v = __fadd_rn(__fmul_rn(v, stdv), mean);          // prediction * std + mean
        v = __fmul_rn(v, control);                        // * control
        v = __fdiv_rn(__fsub_rn(v, mean), stdv);          // (. - mean) / std
    }
    // torch.bucketize(v, bins) (right=False): the first edge that is >= v
    int lo = 0, hi = n_edges;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (bins[mid] >= v) hi = mid; else lo = mid + 1;
    }
    if (value_out && lane == 0) value_out[r] = v;
    const float* e = table + (size_t)lo * D;
    const float* xr = x + (size_t)r * D;
    float* o = out + (size_t)r * D;
    for (int c = lane; c < D; c += 32) o[c] = xr[c] + e[c];
}

__global__ void __launch_bounds__(256) duration_round_kernel(const float* __restrict__ log_d, float d_control,
                                                             float* __restrict__ out, long long n) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float r = __fmul_rn(rintf(__fsub_rn(expf(log_d[i]), 1.0f)), d_control);    // torch.round = half to even
    out[i] = r < 0.f ? 0.f : r;                                                       // clamp(min=0); NaN stays NaN
}

// y = LayerNorm_over_channels(relu(x)) (no mask): relu_1 -> layer_norm_1 of the predictor
__global__ void __launch_bounds__(256) relu_layernorm_ncl_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta, float* __restrict__ y, int D, int L) {
    __shared__ float red[8][33];
    const int b = blockIdx.y, tl = threadIdx.x & 31, g = threadIdx.x >> 5, l = blockIdx.x * 32 + tl;
    const bool live = l < L;
    const float* xb = x + (size_t)b * D * L;
    float* yb = y + (size_t)b * D * L;
    float sum = 0.f;
    if (live) for (int c = g; c < D; c += 8) sum += fmaxf(xb[(size_t)c * L + l], 0.f);
    red[g][tl] = sum;
    __syncthreads();
    float mean = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) mean += red[i][tl];
    mean *= 1.0f / D;
    __syncthreads();
    float sq = 0.f;
    if (live) for (int c = g; c < D; c += 8) { const float d = fmaxf(xb[(size_t)c * L + l], 0.f) - mean; sq = fmaf(d, d, sq); }
    red[g][tl] = sq;
    __syncthreads();
    float var = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) var += red[i][tl];
    const float rstd = rsqrtf(var * (1.0f / D) + 1e-5f);
    if (!live) return;
    for (int c = g; c < D; c += 8) yb[(size_t)c * L + l] = (fmaxf(xb[(size_t)c * L + l], 0.f) - mean) * rstd * gamma[c] + beta[c];
}

int conv_same(const float* x, const float* w, const float* bias, float* out, int B, int Cin, int Cout, int L, int k, cudaStream_t st) {
    SimtConv q;
    memset(&q, 0, sizeof(q));
    q.x = x; q.w = w; q.bias = bias; q.out = out;
    q.B = B; q.Cin = Cin; q.Cout = Cout; q.L = L; q.k = k; q.dil = 1; q.pad = (k - 1) / 2; q.stride = 1;
    q.in_slope = 1.f; q.div = 1.f;
    return simt_conv1d(q, st);
}

}  // namespace

// blob: conv1d_1 weight [F,d_in,k], bias [F], layer_norm_1 gamma, beta [F], conv1d_2 weight [F,F,k], bias, layer_norm_2
// gamma, beta, linear_layer weight [F], bias [1]
extern "C" int vo2w_varpred_create(int d_in, int d_filter, int k, const float* blob_host, size_t n_floats, vo2w_varpred** out) {
    VO2W_REQUIRE(blob_host && out, "varpred_create: null argument");
    VO2W_REQUIRE(d_in > 0 && d_filter > 0, "varpred_create: bad shape (d_in %d, filter %d)", d_in, d_filter);
    // conv1d_2 pads by 1 whatever the kernel size (modules.py:189): only k = 3 keeps the length the mask needs
    VO2W_REQUIRE(k == 3, "varpred_create: kernel_size %d: the reference's second convolution pads by 1, only 3 keeps the length", k);
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    vo2w_varpred* p = new vo2w_varpred();
    p->d_in = d_in; p->d_filter = d_filter; p->k = k;
    const size_t F = (size_t)d_filter;
    size_t off = 0;
    auto take = [&](size_t n) { const size_t o = off; off += n; return o; };
    p->w1 = take(F * d_in * k); p->b1 = take(F); p->g1 = take(F); p->be1 = take(F);
    p->w2 = take(F * F * k); p->b2 = take(F); p->g2 = take(F); p->be2 = take(F);
    p->wl = take(F); p->bl = take(1);
    if (off != n_floats) {
        vo2w_set_error("varpred_create: weight blob has %zu floats, shape needs %zu", n_floats, off);
        delete p;
        return VO2W_E_INVALID;
    }
    if (cudaMalloc(&p->w_dev, n_floats * 4) != cudaSuccess ||
        cudaMemcpy(p->w_dev, blob_host, n_floats * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
        vo2w_set_error("varpred_create: uploading weights failed");
        if (p->w_dev) cudaFree(p->w_dev);
        delete p;
        return VO2W_E_CUDA;
    }
    // Tensor pipe with fp32-level accuracy (conv_tf32.cu, as the decoder and PostNet) where the kernel tiles the shapes
    // (the reference's 256 -> 256 -> 256, k = 3, does): time-major, so nothing is transposed.
    if (tf32_conv_supported(d_filter, d_in, k) && tf32_conv_supported(d_filter, d_filter, k)) {
        p->p2 = tf32_conv_packed_floats(d_filter, d_in, k);
        const size_t total = p->p2 + tf32_conv_packed_floats(d_filter, d_filter, k);
        if (cudaMalloc(&p->p_dev, total * 4) != cudaSuccess) {
            vo2w_set_error("varpred_create: allocating packed weights failed");
            vo2w_varpred_destroy(p);
            return VO2W_E_CUDA;
        }
        if ((rc = tf32_conv_pack(p->w_dev + p->w1, d_filter, d_in, k, p->p_dev, 0)) != VO2W_OK ||
            (rc = tf32_conv_pack(p->w_dev + p->w2, d_filter, d_filter, k, p->p_dev + p->p2, 0)) != VO2W_OK) {
            vo2w_varpred_destroy(p);
            return rc;
        }
        if (cudaDeviceSynchronize() != cudaSuccess) {
            vo2w_set_error("varpred_create: packing weights failed");
            vo2w_varpred_destroy(p);
            return VO2W_E_CUDA;
        }
    }
    *out = p;
    return VO2W_OK;
}

extern "C" int vo2w_varpred_destroy(vo2w_varpred* p) {
    if (!p) return VO2W_OK;
    if (p->w_dev) cudaFree(p->w_dev);
    if (p->p_dev) cudaFree(p->p_dev);
    delete p;
    return VO2W_OK;
}

extern "C" size_t vo2w_varpred_workspace_bytes(const vo2w_varpred* p, int B, int T) {
    if (!p || B <= 0 || T <= 0) return 0;
    const size_t rows = (size_t)B * (T + 2);        // every utterance framed by two zero rows (tensor-pipe path)
    return var_align(rows * 4 * p->d_in) + 2 * var_align(rows * 4 * p->d_filter) + var_align(rows);
}

extern "C" int vo2w_varpred_forward(vo2w_varpred* p, const float* x, const uint8_t* mask, int B, int T, float* out,
                                    void* workspace, size_t workspace_bytes, void* stream) {
    VO2W_REQUIRE(p && x && out, "varpred_forward: null argument");
    VO2W_REQUIRE(B > 0 && T > 0 && B <= 65535, "varpred_forward: bad batch (B=%d T=%d)", B, T);
    const size_t need = vo2w_varpred_workspace_bytes(p, B, T);
    if (!workspace || workspace_bytes < need) {
        vo2w_set_error("varpred_forward: workspace %zu bytes, need %zu", workspace_bytes, need);
        return VO2W_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int F = p->d_filter;
    const size_t bt = (size_t)B * (T + 2) * 4;
    uint8_t* ws = (uint8_t*)workspace;
    float* xn = (float*)ws;
    float* a = (float*)(ws + var_align(bt * p->d_in));
    float* b = (float*)(ws + var_align(bt * p->d_in) + var_align(bt * F));
    uint8_t* pad = ws + var_align(bt * p->d_in) + 2 * var_align(bt * F);
    const float* W = p->w_dev;
    const long long rows_p = (long long)B * (T + 2);
    if (p->p_dev && rows_p < (1ll << 31) && (((uintptr_t)x) & 15) == 0) {
        // relu is the convolution kernel's activation; LayerNorm runs on its time-major rows and zeroes the framing
        // rows again (the reference itself masks only the final projection, modules.py:205-211)
        frame_rows_kernel<<<(unsigned)vo2w_ceil_div(rows_p, 4), 256, 0, st>>>(x, xn, pad, T, p->d_in / 4, rows_p);
        VO2W_CUDA_OK(cudaGetLastError());
        Tf32Conv c1{xn, p->p_dev, W + p->b1, nullptr, a, 1, (int)rows_p, p->d_in, F, p->k, 1};
        int rc = tf32_conv1d(c1, st);
        if (rc != VO2W_OK) return rc;
        if ((rc = rows_layernorm_mask(a, W + p->g1, W + p->be1, pad, b, rows_p, F, st)) != VO2W_OK) return rc;
        Tf32Conv c2{b, p->p_dev + p->p2, W + p->b2, nullptr, a, 1, (int)rows_p, F, F, p->k, 1};
        if ((rc = tf32_conv1d(c2, st)) != VO2W_OK) return rc;
        if ((rc = rows_layernorm_mask(a, W + p->g2, W + p->be2, pad, b, rows_p, F, st)) != VO2W_OK) return rc;
        const long long rows = (long long)B * T;
        project_mask_rows_kernel<<<(unsigned)vo2w_ceil_div(rows, 8), 256, 0, st>>>(b, W + p->wl, W + p->bl, mask, out, rows, T, F);
        VO2W_CUDA_OK(cudaGetLastError());
        return VO2W_OK;
    }
    dim3 grid((unsigned)vo2w_ceil_div(T, 32), (unsigned)B);
    int rc = rows_transpose_add(x, nullptr, xn, B, T, p->d_in, st);
    if (rc != VO2W_OK) return rc;
    if ((rc = conv_same(xn, W + p->w1, W + p->b1, a, B, p->d_in, F, T, p->k, st)) != VO2W_OK) return rc;
    relu_layernorm_ncl_kernel<<<grid, 256, 0, st>>>(a, W + p->g1, W + p->be1, b, F, T);
    VO2W_CUDA_OK(cudaGetLastError());
    if ((rc = conv_same(b, W + p->w2, W + p->b2, a, B, F, F, T, p->k, st)) != VO2W_OK) return rc;
    relu_layernorm_ncl_kernel<<<grid, 256, 0, st>>>(a, W + p->g2, W + p->be2, b, F, T);
    VO2W_CUDA_OK(cudaGetLastError());
    project_mask_kernel<<<grid, 256, 0, st>>>(b, W + p->wl, W + p->bl, mask, out, F, T);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

// ---- the same forward for callers that may not allocate (no counterpart in the reference) ---------------------------
// The predictor's three intermediates live in a file-scope device array instead of a caller workspace.  Every GPU that
// loads the library holds VO2W_VARPRED_STATIC_BYTES of it until the process exits.  In host code `varpred_static_scratch`
// is only the symbol's host-side stand-in: its device address is looked up once per device, with that device current.
static __device__ __align__(256) unsigned char varpred_static_scratch[VO2W_VARPRED_STATIC_BYTES];

namespace {
std::mutex static_scratch_mu;
void* static_scratch_ptr[VO2W_MAX_DEV] = {};

int static_scratch_address(void** out) {
    int dev = -1;
    VO2W_CUDA_OK(cudaGetDevice(&dev));
    VO2W_REQUIRE(dev >= 0 && dev < VO2W_MAX_DEV, "varpred_forward_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(static_scratch_mu);
    if (!static_scratch_ptr[dev]) {
        void* ptr = nullptr;
        VO2W_CUDA_OK(cudaGetSymbolAddress(&ptr, varpred_static_scratch));
        VO2W_REQUIRE(ptr != nullptr && (uintptr_t)ptr % 256 == 0, "varpred_forward_static: bad scratch address %p", ptr);
        static_scratch_ptr[dev] = ptr;
    }
    *out = static_scratch_ptr[dev];
    return VO2W_OK;
}
}  // namespace

extern "C" int vo2w_varpred_forward_static(vo2w_varpred* p, const float* x, const uint8_t* mask, int B, int T, float* out,
                                           void* stream) {
    VO2W_REQUIRE(p && x && out, "varpred_forward_static: null argument");
    VO2W_REQUIRE(B > 0 && T > 0 && B <= 65535, "varpred_forward_static: bad batch (B=%d T=%d)", B, T);
    const size_t need = vo2w_varpred_workspace_bytes(p, B, T);
    if (need > (size_t)VO2W_VARPRED_STATIC_BYTES) {
        vo2w_set_error("varpred_forward_static: B=%d x T=%d needs %zu bytes of scratch, the built-in array holds %zu "
                       "(use vo2w_varpred_forward with a caller workspace)", B, T, need, (size_t)VO2W_VARPRED_STATIC_BYTES);
        return VO2W_E_WORKSPACE;
    }
    void* scratch = nullptr;
    const int rc = static_scratch_address(&scratch);
    if (rc != VO2W_OK) return rc;
    // Nothing in the scratch has to start at zero: the transposed input, both convolution outputs and both LayerNorm
    // outputs are written in full before they are read, so the previous call's contents are simply overwritten.
    return vo2w_varpred_forward(p, x, mask, B, T, out, scratch, (size_t)VO2W_VARPRED_STATIC_BYTES, stream);
}

extern "C" int vo2w_variance_embed(const float* x, const float* value, const float* bins, int n_edges, const float* table,
                                   float mean, float stdv, float control, int denorm, float* value_out, float* out,
                                   long long rows, int D, void* stream) {
    VO2W_REQUIRE(x && value && bins && table && out, "variance_embed: null argument");
    VO2W_REQUIRE(rows >= 0 && D > 0 && n_edges >= 0, "variance_embed: bad shape");
    if (rows == 0) return VO2W_OK;
    variance_embed_kernel<<<(unsigned)vo2w_ceil_div(rows, 8), 256, 0, (cudaStream_t)stream>>>(
        x, value, bins, n_edges, table, mean, stdv, control, denorm, value_out, out, rows, D);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

extern "C" int vo2w_duration_round(const float* log_duration, float d_control, float* out, long long n, void* stream) {
    VO2W_REQUIRE(log_duration && out && n >= 0, "duration_round: bad argument");
    if (n == 0) return VO2W_OK;
    duration_round_kernel<<<(unsigned)vo2w_ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(log_duration, d_control, out, n);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}
