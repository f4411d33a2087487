// mel_linear + PostNet of the acoustic model's tail: the step between the mel decoder and
// the vocoder.
//
//   PostNet.forward (scripts/transformer/Layers.py:126-137, eval mode): x [B,L,n_mel] -> transpose ->
//       4 x tanh(BatchNorm1d(Conv1d k5)) -> BatchNorm1d(Conv1d k5) -> transpose      (dropout is the identity)
//   vTTS.forward (scripts/model/vtts.py:106-107):  output = mel_linear(output); postnet_output = postnet(output) + output
//
// BatchNorm (eval) is an affine map per channel, folded into the convolution's weight and bias by the caller
// (postnet.py), as weight_norm is for the vocoder.  The convolutions run time-major on the tensor pipe with fp32-level
// accuracy (conv_tf32.cu: kind::tf32 with a hi/lo operand split): no transposes, tanh and the `+ output` of
// vtts.py:107 ride the epilogues.  Shapes the tensor-pipe kernel does not tile (channel counts that are not multiples
// of 8 / 16) fall back to the fp32 CUDA-core kernel (conv_simt.cu) between two tiled transposes.
#include <vector>

#include "conv_simt.cuh"
#include "conv_tf32.cuh"
#include "rows.cuh"

struct vo2w_postnet {
    int n_mel = 0, dim = 0, k = 0, n_conv = 0;
    float* w_dev = nullptr;
    float* p_dev = nullptr;       // packed for conv_tf32.cu (null: CUDA-core fallback)
    std::vector<size_t> w_off, b_off, p_off;
};

namespace {

size_t align256(size_t v) { return (v + 255) / 256 * 256; }

}  // namespace

extern "C" int vo2w_postnet_create(int n_mel, int dim, int k, int n_conv, const float* folded_host, size_t n_floats,
                                   vo2w_postnet** out) {
    VO2W_REQUIRE(folded_host && out, "postnet_create: null argument");
    VO2W_REQUIRE(n_mel > 0 && dim > 0 && n_conv >= 2 && (k & 1) && k >= 1 && k <= 11, "postnet_create: bad shape (n_mel %d dim %d k %d n_conv %d)",
                 n_mel, dim, k, n_conv);
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    vo2w_postnet* p = new vo2w_postnet();
    p->n_mel = n_mel; p->dim = dim; p->k = k; p->n_conv = n_conv;
    size_t off = 0;
    for (int i = 0; i < n_conv; ++i) {
        const int cin = i == 0 ? n_mel : dim, cout = i == n_conv - 1 ? n_mel : dim;
        p->w_off.push_back(off);
        off += (size_t)cout * cin * k;
        p->b_off.push_back(off);
        off += (size_t)cout;
    }
    if (off != n_floats) {
        vo2w_set_error("postnet_create: weight blob has %zu floats, shape needs %zu", n_floats, off);
        delete p;
        return VO2W_E_INVALID;
    }
    if (cudaMalloc(&p->w_dev, n_floats * 4) != cudaSuccess ||
        cudaMemcpy(p->w_dev, folded_host, n_floats * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
        vo2w_set_error("postnet_create: uploading weights failed");
        if (p->w_dev) cudaFree(p->w_dev);
        delete p;
        return VO2W_E_CUDA;
    }
    bool tensor = true;
    size_t poff = 0;
    for (int i = 0; i < n_conv; ++i) {
        const int cin = i == 0 ? n_mel : dim, cout = i == n_conv - 1 ? n_mel : dim;
        tensor = tensor && tf32_conv_supported(cout, cin, k);
        p->p_off.push_back(poff);
        poff += tf32_conv_packed_floats(cout, cin, k);
    }
    if (tensor) {
        if (cudaMalloc(&p->p_dev, poff * 4) != cudaSuccess) {
            vo2w_set_error("postnet_create: allocating packed weights failed");
            vo2w_postnet_destroy(p);
            return VO2W_E_CUDA;
        }
        for (int i = 0; i < n_conv; ++i) {
            const int cin = i == 0 ? n_mel : dim, cout = i == n_conv - 1 ? n_mel : dim;
            if ((rc = tf32_conv_pack(p->w_dev + p->w_off[i], cout, cin, k, p->p_dev + p->p_off[i], 0)) != VO2W_OK) {
                vo2w_postnet_destroy(p);
                return rc;
            }
        }
        if (cudaDeviceSynchronize() != cudaSuccess) {
            vo2w_set_error("postnet_create: packing weights failed");
            vo2w_postnet_destroy(p);
            return VO2W_E_CUDA;
        }
    }
    *out = p;
    return VO2W_OK;
}

extern "C" int vo2w_postnet_destroy(vo2w_postnet* p) {
    if (!p) return VO2W_OK;
    if (p->w_dev) cudaFree(p->w_dev);
    if (p->p_dev) cudaFree(p->p_dev);
    delete p;
    return VO2W_OK;
}

extern "C" size_t vo2w_postnet_workspace_bytes(const vo2w_postnet* p, int B, int L) {
    if (!p || B <= 0 || L <= 0) return 0;
    return align256((size_t)B * p->n_mel * L * 4) + 2 * align256((size_t)B * p->dim * L * 4);
}

extern "C" int vo2w_postnet_forward(vo2w_postnet* p, const float* x, int B, int L, float* out, int add_input,
                                    void* workspace, size_t workspace_bytes, void* stream) {
    VO2W_REQUIRE(p && x && out, "postnet_forward: null argument");
    VO2W_REQUIRE(B > 0 && L > 0 && B <= 65535, "postnet_forward: bad batch (B=%d L=%d)", B, L);
    const size_t need = vo2w_postnet_workspace_bytes(p, B, L);
    if (!workspace || workspace_bytes < need) {
        vo2w_set_error("postnet_forward: workspace %zu bytes, need %zu", workspace_bytes, need);
        return VO2W_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* ws = (uint8_t*)workspace;
    if (p->p_dev) {
        // time-major all the way: [B,L,n_mel] -> tanh(conv) x (n_conv - 1) -> conv (+ x)
        float* buf[2] = {(float*)(ws + align256((size_t)B * p->n_mel * L * 4)),
                         (float*)(ws + align256((size_t)B * p->n_mel * L * 4) + align256((size_t)B * p->dim * L * 4))};
        const float* cur = x;
        for (int i = 0; i < p->n_conv; ++i) {
            const bool last = i == p->n_conv - 1;
            Tf32Conv c{cur, p->p_dev + p->p_off[i], p->w_dev + p->b_off[i], (last && add_input) ? x : nullptr,
                       last ? out : buf[i & 1], B, L, i == 0 ? p->n_mel : p->dim, last ? p->n_mel : p->dim, p->k,
                       last ? 0 : 2};           // torch.tanh on all but the last convolution (Layers.py:129-131)
            const int rc = tf32_conv1d(c, st);
            if (rc != VO2W_OK) return rc;
            cur = c.out;
        }
        return VO2W_OK;
    }
    float* x_ncl = (float*)ws;
    float* buf[2] = {(float*)(ws + align256((size_t)B * p->n_mel * L * 4)),
                     (float*)(ws + align256((size_t)B * p->n_mel * L * 4) + align256((size_t)B * p->dim * L * 4))};
    // x.contiguous().transpose(1, 2): [B,L,n_mel] -> [B,n_mel,L]
    int rc = rows_transpose_add(x, nullptr, x_ncl, B, L, p->n_mel, st);
    if (rc != VO2W_OK) return rc;
    const float* cur = x_ncl;
    for (int i = 0; i < p->n_conv; ++i) {
        const bool last = i == p->n_conv - 1;
        SimtConv q;
        memset(&q, 0, sizeof(q));
        q.x = cur; q.w = p->w_dev + p->w_off[i]; q.bias = p->w_dev + p->b_off[i];
        q.out = last ? buf[i & 1] : buf[i & 1];
        q.B = B; q.Cin = i == 0 ? p->n_mel : p->dim; q.Cout = last ? p->n_mel : p->dim;
        q.L = L; q.k = p->k; q.dil = 1; q.pad = (p->k - 1) / 2; q.stride = 1;
        q.in_slope = 1.f; q.div = 1.f;
        q.tanh_out = last ? 0 : 1;          // torch.tanh on all but the last convolution (Layers.py:129-131)
        rc = simt_conv1d(q, st);
        if (rc != VO2W_OK) return rc;
        cur = q.out;
    }
    // transpose back; `+ output` of vtts.py:107 fused on request (the addend is the time-major input itself)
    return rows_transpose_add(cur, add_input ? x : nullptr, out, B, p->n_mel, L, st);
}

extern "C" int vo2w_linear_forward(const float* x, const float* w, const float* bias, long long rows, int K, int N, float* y,
                                   float* scratch, void* stream) {
    VO2W_REQUIRE(x && w && y, "linear_forward: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    if (scratch && rows > 0 && rows < (1ll << 31) && tf32_conv_supported(N, K, 1) && ((uintptr_t)bias & 15) == 0) {
        // a k=1 convolution over one "utterance" of `rows` frames on the tensor pipe (fp32-level accuracy)
        const int rc = tf32_conv_pack(w, N, K, 1, scratch, st);
        if (rc != VO2W_OK) return rc;
        Tf32Conv c{x, scratch, bias, nullptr, y, 1, (int)rows, K, N, 1, 0};
        return tf32_conv1d(c, st);
    }
    return rows_linear(x, w, bias, rows, K, N, y, st);
}
