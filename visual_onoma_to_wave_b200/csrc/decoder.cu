// FFT-block stack of the mel decoder: the step LengthRegulator feeds.
//
//   Decoder.forward (scripts/transformer/Models.py:168-197, eval):  x + sinusoid position table, then per layer
//   FFTBlock.forward (scripts/transformer/Layers.py:21-30):
//       MultiHeadAttention (SubLayers.py:31-62): q,k,v = Linear(x); softmax(q k^T / sqrt(d_k), key mask) v;
//                                                LayerNorm(fc(.) + x);            masked_fill(mask, 0)
//       PositionwiseFeedForward (SubLayers.py:95-104): LayerNorm(w_2(relu(w_1(x^T)))^T + x);  masked_fill(mask, 0)
//   (dropout is the identity in eval mode).
//
// Time-major [B,L,D] rows throughout (the reference's own layout: no transposes).  Every Linear / Conv1d runs on the
// tensor pipe with fp32-level accuracy (conv_tf32.cu: kind::tf32 with a hi/lo operand split; q|k|v fused into one
// [3D,D] layer, relu and the residual adds fused into the epilogues), attention is a tiled online-softmax kernel
// (64 queries x 64 keys per step, d_k = 128, fp32 CUDA cores), LayerNorm + padding mask one warp-per-row kernel.
#include <vector>

#include "conv_tf32.cuh"
#include "rows.cuh"

namespace {

constexpr int AQ = 64, AK = 64, AD = 128;          // queries / keys per tile, head size
constexpr int AQS = AD + 1, APS = AK + 1;           // padded shared-memory strides
constexpr size_t ATTN_SMEM = ((size_t)AQ * AQS + (size_t)AK * AQS + (size_t)AQ * APS + 3 * AQ) * 4;

// out[b, q, h*128 + :] = softmax_k(q_h . k_h / sqrt(128) masked) v_h   for one (query tile, head, utterance)
// qkv [B*L][3*D] rows (q | k | v, head h at columns h*128), mask [B][L] (1 = padding key) or null, out [B*L][D]
__global__ void __launch_bounds__(256) attention_kernel(const float* __restrict__ qkv, const uint8_t* __restrict__ mask,
                                                        float* __restrict__ out, int L, int D, float scale) {
    extern __shared__ float sm[];
    float* Qs = sm;                       // [AQ][AQS]
    float* KVs = Qs + AQ * AQS;           // [AK][AQS]   keys, then values
    float* Ps = KVs + AK * AQS;           // [AQ][APS]   scores, then probabilities
    float* rowM = Ps + AQ * APS;          // running max, running sum, rescale factor per query row
    float* rowL = rowM + AQ;
    float* rowA = rowL + AQ;
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * AQ;
    const int tid = threadIdx.x, tq = tid >> 4, tk = tid & 15, warp = tid >> 5, lane = tid & 31;
    const size_t row_stride = (size_t)3 * D;
    const float* base = qkv + (size_t)b * L * row_stride + (size_t)h * AD;
    const uint8_t* mb = mask ? mask + (size_t)b * L : nullptr;

    for (int e = tid; e < AQ * AD; e += 256) {
        const int r = e >> 7, d = e & 127;
        Qs[r * AQS + d] = (q0 + r < L) ? base[(size_t)(q0 + r) * row_stride + d] * scale : 0.f;
    }
    if (tid < AQ) { rowM[tid] = -1e30f; rowL[tid] = 0.f; rowA[tid] = 0.f; }
    float O[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int e = 0; e < 8; ++e) O[i][e] = 0.f;
    __syncthreads();

    for (int k0 = 0; k0 < L; k0 += AK) {
        for (int e = tid; e < AK * AD; e += 256) {
            const int r = e >> 7, d = e & 127;
            KVs[r * AQS + d] = (k0 + r < L) ? base[(size_t)(k0 + r) * row_stride + D + d] : 0.f;
        }
        __syncthreads();
        float S[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) S[i][j] = 0.f;
        for (int d = 0; d < AD; ++d) {
            float qv[4], kv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) qv[i] = Qs[(tq * 4 + i) * AQS + d];
#pragma unroll
            for (int j = 0; j < 4; ++j) kv[j] = KVs[(tk * 4 + j) * AQS + d];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) S[i][j] = fmaf(qv[i], kv[j], S[i][j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int kk = k0 + tk * 4 + j;
            const bool dead = kk >= L || (mb && mb[kk]);          // masked_fill(mask, -inf) before the softmax
#pragma unroll
            for (int i = 0; i < 4; ++i) Ps[(tq * 4 + i) * APS + tk * 4 + j] = dead ? -1e30f : S[i][j];
        }
        __syncthreads();
        // online softmax: warp w owns rows 8w .. 8w+7, lanes walk the 64 keys
        for (int r = warp * 8; r < warp * 8 + 8; ++r) {
            const float s0 = Ps[r * APS + lane], s1 = Ps[r * APS + lane + 32];
            float mx = fmaxf(s0, s1);
#pragma unroll
            for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            const float m_old = rowM[r], m_new = fmaxf(m_old, mx);
            const float p0 = s0 <= -1e30f ? 0.f : expf(s0 - m_new), p1 = s1 <= -1e30f ? 0.f : expf(s1 - m_new);
            float sum = p0 + p1;
#pragma unroll
            for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            Ps[r * APS + lane] = p0;
            Ps[r * APS + lane + 32] = p1;
            if (lane == 0) {
                const float a = expf(m_old - m_new);
                rowA[r] = a;
                rowL[r] = rowL[r] * a + sum;
                rowM[r] = m_new;
            }
        }
        __syncthreads();
        for (int e = tid; e < AK * AD; e += 256) {                  // values over the keys
            const int r = e >> 7, d = e & 127;
            KVs[r * AQS + d] = (k0 + r < L) ? base[(size_t)(k0 + r) * row_stride + 2 * D + d] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float a = rowA[tq * 4 + i];
#pragma unroll
            for (int e = 0; e < 8; ++e) O[i][e] *= a;
        }
        for (int k = 0; k < AK; ++k) {
            float pv[4], vv[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) pv[i] = Ps[(tq * 4 + i) * APS + k];
#pragma unroll
            for (int e = 0; e < 8; ++e) vv[e] = KVs[k * AQS + tk + 16 * e];      // column tk + 16 e: conflict-free
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int e = 0; e < 8; ++e) O[i][e] = fmaf(pv[i], vv[e], O[i][e]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = q0 + tq * 4 + i;
        if (q >= L) continue;
        const float inv = 1.f / rowL[tq * 4 + i];
        float* o = out + ((size_t)b * L + q) * D + (size_t)h * AD;
#pragma unroll
        for (int e = 0; e < 8; ++e) o[tk + 16 * e] = O[i][e] * inv;
    }
}

// y[r,:] = mask[r] ? 0 : LayerNorm(x[r,:]) * gamma + beta   (eps 1e-5, biased variance; nn.LayerNorm(d_model), then
// masked_fill(mask.unsqueeze(-1), 0): SubLayers.py:60,103, Layers.py:24-28).  One warp per row, D = 32 * PER.
template <int PER>
__global__ void __launch_bounds__(256) layernorm_rows_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, const uint8_t* __restrict__ mask,
                                                             float* __restrict__ y, long long rows) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    constexpr int D = 32 * PER;
    const float* xr = x + (size_t)r * D;
    float v[PER], sum = 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) { v[i] = xr[lane + 32 * i]; sum += v[i]; }
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum * (1.0f / D);
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) { const float d = v[i] - mean; sq = fmaf(d, d, sq); }
#pragma unroll
    for (int o = 16; o; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = rsqrtf(sq * (1.0f / D) + 1e-5f);
    const bool dead = mask && mask[r];
    float* yr = y + (size_t)r * D;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int c = lane + 32 * i;
        yr[c] = dead ? 0.f : (v[i] - mean) * rstd * gamma[c] + beta[c];
    }
}

// in place: S[r, 0..Lp) -> softmax(S[r, :] * scale) over the valid keys (j < L and not padding), zeros elsewhere
// (attn / temperature, masked_fill(mask, -inf), Softmax(dim=2): Modules.py:15-22).  rows r = (b, l, h); one warp per row,
// the row in registers: MAXP values per lane (Lp <= 32 * MAXP), instantiated for 8 / 16 / 32 so that a 512-key row does
// not pay for 1 024 (the results do not depend on MAXP: the extra slots hold -inf and are never summed).
template <int MAXP>
__global__ void __launch_bounds__(256) softmax_rows_kernel(float* __restrict__ S, const uint8_t* __restrict__ mask, long long rows,
                                                           int L, int Lp, int rows_per_b, float scale) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    float* row = S + (size_t)r * Lp;
    const uint8_t* mb = mask ? mask + (size_t)(r / rows_per_b) * L : nullptr;
    float v[MAXP];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
        const int j = lane + 32 * i;
        v[i] = -INFINITY;
        if (j < L && !(mb && mb[j])) v[i] = row[j] * scale;
        mx = fmaxf(mx, v[i]);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
        v[i] = expf(v[i] - mx);                       // -inf - mx -> 0; an all-masked row gives NaN like the reference
        if (lane + 32 * i < Lp) sum += v[i];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.f / sum;
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
        const int j = lane + 32 * i;
        if (j < Lp) row[j] = v[i] * inv;
    }
}

// Same row softmax for rows of any length (Lp > 1024: long-form inputs, which the reference accepts -- it builds its
// sinusoid table on the fly above max_seq_len, Models.py:171-188).  Three passes over the row (max, sum, write); the
// row (<= 16 KB at 4096 keys) stays in L1 between them.  One warp per row.
__global__ void __launch_bounds__(256) softmax_rows_long_kernel(float* __restrict__ S, const uint8_t* __restrict__ mask, long long rows,
                                                                int L, int Lp, int rows_per_b, float scale) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    float* row = S + (size_t)r * Lp;
    const uint8_t* mb = mask ? mask + (size_t)(r / rows_per_b) * L : nullptr;
    float mx = -INFINITY;
    for (int j = lane; j < L; j += 32)
        if (!(mb && mb[j])) mx = fmaxf(mx, row[j] * scale);
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int j = lane; j < L; j += 32)
        if (!(mb && mb[j])) sum += expf(row[j] * scale - mx);
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.f / sum;
    const bool none = mx == -INFINITY;                // an all-masked row gives NaN like the reference (and the short kernel)
    for (int j = lane; j < Lp; j += 32) {
        const bool dead = j >= L || (mb && mb[j]);
        row[j] = none ? __int_as_float(0x7fc00000) : (dead ? 0.f : expf(row[j] * scale - mx) * inv);
    }
}

}  // namespace

// Attention as two batched GEMMs on the tensor pipe (conv_tf32.cu, fp32-level accuracy) around a row softmax:
//   S[b,l,h,:] = q_h[b,l,:] . k_h[b,:,:]^T      (k = 1 "convolution" whose weights are head h's keys of utterance b)
//   P = softmax(S / sqrt(d_k)) over the unmasked keys
//   out[b,l,h*128:] = P[b,l,h,:] . v_h[b,:,:]   (weights = the values, transposed while packing)
// S / P live in the workspace as [B, L, H, Lp], Lp = L rounded up to 64 (zero probability on the padding columns).
int rows_attention_gemm(const float* qkv, const uint8_t* mask, float* out, int B, int L, int D, int n_head, float* S, float* pk,
                        float* pv, cudaStream_t st) {
    VO2W_REQUIRE(D == n_head * AD, "attention: head size must be %d (d_model %d, heads %d)", AD, D, n_head);
    const int Lp = (L + 63) / 64 * 64;
    VO2W_REQUIRE(Lp <= 16384, "attention: %d frames", L);
    const size_t row3 = (size_t)3 * D;
    int rc = tf32_pack_strided(qkv + D, (size_t)L * row3, AD, row3, 1, B, n_head, Lp, AD, L, AD, pk, st);
    if (rc != VO2W_OK) return rc;
    Tf32Conv qk{qkv, pk, nullptr, nullptr, S, B, L, AD, Lp, 1, 0};
    qk.ldx = (int)row3; qk.x_hs = AD; qk.ldo = n_head * Lp; qk.o_hs = Lp; qk.heads = n_head; qk.w_zs = (size_t)2 * Lp * AD;
    if ((rc = tf32_conv1d(qk, st)) != VO2W_OK) return rc;
    const long long rows = (long long)B * L * n_head;
    const unsigned sgrid = (unsigned)vo2w_ceil_div(rows, 8);
    const float sscale = 1.0f / sqrtf((float)AD);
    if (Lp <= 256)       // the row fits one warp's registers: one pass
        softmax_rows_kernel<8><<<sgrid, 256, 0, st>>>(S, mask, rows, L, Lp, L * n_head, sscale);
    else if (Lp <= 512)
        softmax_rows_kernel<16><<<sgrid, 256, 0, st>>>(S, mask, rows, L, Lp, L * n_head, sscale);
    else if (Lp <= 1024)
        softmax_rows_kernel<32><<<sgrid, 256, 0, st>>>(S, mask, rows, L, Lp, L * n_head, sscale);
    else
        softmax_rows_long_kernel<<<(unsigned)vo2w_ceil_div(rows, 8), 256, 0, st>>>(S, mask, rows, L, Lp, L * n_head, 1.0f / sqrtf((float)AD));
    VO2W_CUDA_OK(cudaGetLastError());
    if ((rc = tf32_pack_strided(qkv + 2 * D, (size_t)L * row3, AD, 1, row3, B, n_head, AD, Lp, AD, L, pv, st)) != VO2W_OK) return rc;
    Tf32Conv pvc{S, pv, nullptr, nullptr, out, B, L, Lp, AD, 1, 0};
    pvc.ldx = n_head * Lp; pvc.x_hs = Lp; pvc.ldo = D; pvc.o_hs = AD; pvc.heads = n_head; pvc.w_zs = (size_t)2 * Lp * AD;
    return tf32_conv1d(pvc, st);
}

int rows_layernorm_mask(const float* x, const float* gamma, const float* beta, const uint8_t* mask, float* y, long long rows, int D,
                        cudaStream_t st) {
    const unsigned grid = (unsigned)vo2w_ceil_div(rows, 8);
    if (D == 256) layernorm_rows_kernel<8><<<grid, 256, 0, st>>>(x, gamma, beta, mask, y, rows);
    else if (D == 128) layernorm_rows_kernel<4><<<grid, 256, 0, st>>>(x, gamma, beta, mask, y, rows);
    else if (D == 384) layernorm_rows_kernel<12><<<grid, 256, 0, st>>>(x, gamma, beta, mask, y, rows);
    else if (D == 512) layernorm_rows_kernel<16><<<grid, 256, 0, st>>>(x, gamma, beta, mask, y, rows);
    else {
        vo2w_set_error("layernorm: width %d unsupported (128, 256, 384, 512)", D);
        return VO2W_E_INVALID;
    }
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int rows_attention(const float* qkv, const uint8_t* mask, float* out, int B, int L, int D, int n_head, cudaStream_t st) {
    VO2W_REQUIRE(D == n_head * AD, "attention: head size must be %d (d_model %d, heads %d)", AD, D, n_head);
    static bool attr_set[VO2W_MAX_DEV];
    const int cur = vo2w_cur_dev();
    if (!attr_set[cur]) {
        VO2W_CUDA_OK(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ATTN_SMEM));
        attr_set[cur] = true;
    }
    dim3 grid((unsigned)vo2w_ceil_div(L, AQ), (unsigned)n_head, (unsigned)B);
    attention_kernel<<<grid, 256, ATTN_SMEM, st>>>(qkv, mask, out, L, D, 1.0f / sqrtf((float)AD));
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// handle: the layer stack's parameters in one device blob
// ---------------------------------------------------------------------------------------------------------------
struct vo2w_decoder {
    int d_model = 0, n_head = 0, d_inner = 0, k1 = 0, k2 = 0, n_layers = 0;
    float* w_dev = nullptr;       // the caller's blob (biases, LayerNorm parameters are used from here)
    float* p_dev = nullptr;       // the four weight matrices of every layer, packed for conv_tf32.cu
    struct Layer { size_t wqkv, bqkv, wfc, bfc, ln1g, ln1b, w1, b1, w2, b2, ln2g, ln2b; size_t pqkv, pfc, p1, p2; };
    std::vector<Layer> layers;
};

namespace {

size_t dec_align(size_t v) { return (v + 255) / 256 * 256; }

int conv(const float* x, const float* wp, const float* bias, const float* res, float* out, int B, int Cin, int Cout, int L, int k,
         int act, cudaStream_t st) {
    Tf32Conv c{x, wp, bias, res, out, B, L, Cin, Cout, k, act};
    return tf32_conv1d(c, st);
}

}  // namespace

// blob, per layer: [w_qs | w_ks | w_vs] [3D,D], their biases [3D], fc [D,D] + bias, attention LayerNorm gamma, beta,
// w_1 [d_inner,D,k1] + bias, w_2 [D,d_inner,k2] + bias, feed-forward LayerNorm gamma, beta
extern "C" int vo2w_decoder_create(int d_model, int n_head, int d_inner, int k1, int k2, int n_layers, const float* blob_host,
                                   size_t n_floats, vo2w_decoder** out) {
    VO2W_REQUIRE(blob_host && out, "decoder_create: null argument");
    VO2W_REQUIRE(d_model > 0 && n_head > 0 && d_model == n_head * AD, "decoder_create: d_model %d / heads %d: head size must be %d",
                 d_model, n_head, AD);
    VO2W_REQUIRE(d_inner > 0 && n_layers > 0 && tf32_conv_supported(d_inner, d_model, k1) && tf32_conv_supported(d_model, d_inner, k2),
                 "decoder_create: unsupported shape (d_inner %d must be a multiple of 64, odd kernels <= 11: %d/%d, layers %d)",
                 d_inner, k1, k2, n_layers);
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    vo2w_decoder* d = new vo2w_decoder();
    d->d_model = d_model; d->n_head = n_head; d->d_inner = d_inner; d->k1 = k1; d->k2 = k2; d->n_layers = n_layers;
    const size_t D = (size_t)d_model, H = (size_t)d_inner;
    size_t off = 0, poff = 0;
    auto take = [&](size_t n) { const size_t o = off; off += n; return o; };
    auto ptake = [&](size_t n) { const size_t o = poff; poff += 2 * n; return o; };
    for (int i = 0; i < n_layers; ++i) {
        vo2w_decoder::Layer l;
        l.pqkv = ptake(3 * D * D); l.pfc = ptake(D * D); l.p1 = ptake(H * D * k1); l.p2 = ptake(D * H * k2);
        l.wqkv = take(3 * D * D); l.bqkv = take(3 * D);
        l.wfc = take(D * D); l.bfc = take(D);
        l.ln1g = take(D); l.ln1b = take(D);
        l.w1 = take(H * D * k1); l.b1 = take(H);
        l.w2 = take(D * H * k2); l.b2 = take(D);
        l.ln2g = take(D); l.ln2b = take(D);
        d->layers.push_back(l);
    }
    if (off != n_floats) {
        vo2w_set_error("decoder_create: weight blob has %zu floats, shape needs %zu", n_floats, off);
        delete d;
        return VO2W_E_INVALID;
    }
    if (cudaMalloc(&d->w_dev, n_floats * 4) != cudaSuccess || cudaMalloc(&d->p_dev, poff * 4) != cudaSuccess ||
        cudaMemcpy(d->w_dev, blob_host, n_floats * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
        vo2w_set_error("decoder_create: uploading weights failed");
        vo2w_decoder_destroy(d);
        return VO2W_E_CUDA;
    }
    for (const auto& l : d->layers) {
        if ((rc = tf32_conv_pack(d->w_dev + l.wqkv, 3 * d_model, d_model, 1, d->p_dev + l.pqkv, 0)) != VO2W_OK ||
            (rc = tf32_conv_pack(d->w_dev + l.wfc, d_model, d_model, 1, d->p_dev + l.pfc, 0)) != VO2W_OK ||
            (rc = tf32_conv_pack(d->w_dev + l.w1, d_inner, d_model, k1, d->p_dev + l.p1, 0)) != VO2W_OK ||
            (rc = tf32_conv_pack(d->w_dev + l.w2, d_model, d_inner, k2, d->p_dev + l.p2, 0)) != VO2W_OK) {
            vo2w_decoder_destroy(d);
            return rc;
        }
    }
    if (cudaDeviceSynchronize() != cudaSuccess) {
        vo2w_set_error("decoder_create: packing weights failed");
        vo2w_decoder_destroy(d);
        return VO2W_E_CUDA;
    }
    *out = d;
    return VO2W_OK;
}

extern "C" int vo2w_decoder_destroy(vo2w_decoder* d) {
    if (!d) return VO2W_OK;
    if (d->w_dev) cudaFree(d->w_dev);
    if (d->p_dev) cudaFree(d->p_dev);
    delete d;
    return VO2W_OK;
}

extern "C" size_t vo2w_decoder_workspace_bytes(const vo2w_decoder* d, int B, int L) {
    if (!d || B <= 0 || L <= 0) return 0;
    const size_t bl = (size_t)B * L * 4;
    const size_t Lp = ((size_t)L + 63) / 64 * 64;
    return 2 * dec_align(bl * d->d_model) + dec_align(bl * 3 * d->d_model) + dec_align(bl * d->d_inner) +
           dec_align(bl * d->n_head * Lp) + 2 * dec_align((size_t)B * d->n_head * 2 * Lp * AD * 4);
}

// x [B,L,D] time-major, pos [L,D] (the sinusoid table's first L rows), mask [B,L] bytes (1 = padding) or null,
// out [B,L,D]
extern "C" int vo2w_decoder_forward(vo2w_decoder* d, const float* x, const float* pos, const uint8_t* mask, int B, int L,
                                    float* out, void* workspace, size_t workspace_bytes, void* stream) {
    VO2W_REQUIRE(d && x && pos && out, "decoder_forward: null argument");
    VO2W_REQUIRE(B > 0 && L > 0 && B <= 65535, "decoder_forward: bad batch (B=%d L=%d)", B, L);
    const size_t need = vo2w_decoder_workspace_bytes(d, B, L);
    if (!workspace || workspace_bytes < need) {
        vo2w_set_error("decoder_forward: workspace %zu bytes, need %zu", workspace_bytes, need);
        return VO2W_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int D = d->d_model, H = d->d_inner;
    const size_t bl = (size_t)B * L * 4;
    uint8_t* ws = (uint8_t*)workspace;
    float* cur = (float*)ws;                                   // [B,L,D] layer input / output
    float* tmp = (float*)(ws + dec_align(bl * D));             // [B,L,D] pre-LayerNorm sums
    float* qkv = (float*)(ws + 2 * dec_align(bl * D));         // [B,L,3D]
    float* hid = (float*)(ws + 2 * dec_align(bl * D) + dec_align(bl * 3 * D));   // [B,L,d_inner]; first holds the attention output
    const long long rows = (long long)B * L;
    const size_t Lp = ((size_t)L + 63) / 64 * 64;
    uint8_t* wa = ws + 2 * dec_align(bl * D) + dec_align(bl * 3 * D) + dec_align(bl * H);
    float* att_s = (float*)wa;                                  // [B,L,heads,Lp] scores / probabilities
    float* att_k = (float*)(wa + dec_align(bl * d->n_head * Lp));            // packed keys / values per (b, head)
    float* att_v = (float*)(wa + dec_align(bl * d->n_head * Lp) + dec_align((size_t)B * d->n_head * 2 * Lp * AD * 4));

    // enc_seq + position_enc[:, :L]  (Models.py:184-186)
    int rc = rows_add_table(x, pos, cur, B, L, D, st);
    if (rc != VO2W_OK) return rc;
    const float* W = d->w_dev;
    const float* P = d->p_dev;
    for (size_t i = 0; i < d->layers.size(); ++i) {
        const auto& l = d->layers[i];
        float* dst = i + 1 == d->layers.size() ? out : cur;
        // MultiHeadAttention (SubLayers.py:31-62)
        if ((rc = conv(cur, P + l.pqkv, W + l.bqkv, nullptr, qkv, B, D, 3 * D, L, 1, 0, st)) != VO2W_OK) return rc;
        // tensor-pipe attention for every length the score tensor is sized for; the CUDA-core online-softmax kernel
        // (no score tensor) stays for sequences beyond that
        rc = Lp > 16384 ? rows_attention(qkv, mask, hid, B, L, D, d->n_head, st)
                        : rows_attention_gemm(qkv, mask, hid, B, L, D, d->n_head, att_s, att_k, att_v, st);
        if (rc != VO2W_OK) return rc;
        if ((rc = conv(hid, P + l.pfc, W + l.bfc, cur, tmp, B, D, D, L, 1, 0, st)) != VO2W_OK) return rc;
        if ((rc = rows_layernorm_mask(tmp, W + l.ln1g, W + l.ln1b, mask, cur, rows, D, st)) != VO2W_OK) return rc;
        // PositionwiseFeedForward (SubLayers.py:95-104); relu rides the first convolution's epilogue
        if ((rc = conv(cur, P + l.p1, W + l.b1, nullptr, hid, B, D, H, L, d->k1, 1, st)) != VO2W_OK) return rc;
        if ((rc = conv(hid, P + l.p2, W + l.b2, cur, tmp, B, H, D, L, d->k2, 0, st)) != VO2W_OK) return rc;
        if ((rc = rows_layernorm_mask(tmp, W + l.ln2g, W + l.ln2b, mask, dst, rows, D, st)) != VO2W_OK) return rc;
    }
    return VO2W_OK;
}
