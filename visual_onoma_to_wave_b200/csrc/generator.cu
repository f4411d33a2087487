// HiFi-GAN V1 Generator on sm_100a: handle, workspace plan and the forward schedule.
//
// Replaces hifigan.Generator (scripts/hifigan/models.py:112-174) behind the C ABI of
// include/vo2w.h.  Two precision modes share this schedule:
//   VO2W_PREC_BF16  tcgen05 kernels on chunked bf16 operand buffers: every ResBlock iteration
//                   (conv1 -> lrelu -> conv2 -> +x) is ONE fused launch (pair_umma.cu), conv_pre /
//                   upsamplers / conv_post are single-conv launches (conv_umma.cu); the last
//                   iteration of all ResBlocks of a stage plus the MRF average is one launch too;
//                   activations travel through HBM as bf16;
//   VO2W_PREC_FP32  tcgen05 kind::tf32 kernels with a hi/lo operand split (conv_tf32.cu) on time-major fp32 tensors.
#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "conv_simt.cuh"
#include "conv_tf32.cuh"
#include "rows.cuh"
#include "pack.cuh"
#include "pair_umma.cuh"
#include "stream_umma.cuh"

namespace {

struct LayerW {            // one reference layer inside the fp32 blob
    size_t w_off, b_off;   // float offsets
    int d0, d1, k;         // weight shape
};

struct StageBufs {         // per upsampling stage, offsets in bytes into the workspace
    int C = 0;
    long long L = 0, Lp = 0;
    // bf16 mode: operand-layout buffers [B][C/8][Lp][8]
    // every tensor is a "stream": bf16(leaky_relu(value, 0.1)) = the next conv's MMA operand; the value itself
    // (the residual of ResBlock.forward) is recovered from it by the consumer
    size_t xa = 0;                 // ups output
    size_t ra[3][2] = {};          // ResBlock j after iteration 0 / 1
    size_t mrf = 0;                // leaky-relu(xs / num_kernels) = operand of the next layer
    // fp32 mode (time-major [B,L,C])
    size_t x_f32 = 0, rr = 0, tt = 0, xs = 0;
};

struct Plan {
    int B = 0, T = 0;
    size_t melb = 0, a0 = 0, c0_f32 = 0;
    size_t i16 = 0;                // int16 PCM staging of the host path (its own region: the operand buffers' zero pad
                                   // rows are an invariant between forwards, so nothing may be staged on top of them)
    StageBufs st[4];
    size_t total = 0;
};

}  // namespace

struct vo2w_generator {
    vo2w_gen_config cfg;
    int precision = 0;
    bool debug_taps = false;
    int n_stage = 0, nk = 0;
    std::vector<LayerW> lw;        // conv_pre, ups[i], resblocks[n].convs1[m], .convs2[m], conv_post (state-dict order)
    size_t n_floats = 0;
    float* w_dev = nullptr;        // fp32 blob on the device (fp32 mode + biases)
    // fp32 mode: every layer but conv_post packed for the tensor-pipe kernel (conv_tf32.cu); a ConvTranspose1d becomes a
    // k = 3 convolution onto stride * Cout phase-major channels (see ups_as_conv_kernel)
    float* p32 = nullptr;
    std::vector<size_t> p32_off;   // per layer of `lw`
    float* ups_bias = nullptr;     // bias replicated per phase, ups_bias_off[i]
    size_t ups_bias_off[4] = {};
    // bf16 mode
    PackedLayer conv_pre, conv_post, ups[4];   // umma_conv_kernel packing (conv_post when it is not fused; A/B fallback)
    SuLayer s_pre, s_ups[4];       // stream_umma_kernel packing: conv_pre and the upsamplers on persistent CTA pairs
    bool use_stream = true;        // VO2W_SINGLE_CONV=old selects umma_conv_kernel for conv_pre / ups (A/B runs)
    bool fuse_post = false;        // conv_post + tanh ride the last stage's MRF launch (32 channels, k = 7, CTA pairs)
    float post_w[PAIR_POST_K * 32] = {};   // conv_post weights [tap][channel] fp32
    float post_b = 0.f;
    PackedPair pairs[12][3];       // resblocks[n] iteration m: convs1[m] + convs2[m]
    PackedGroup iter_group[12][2]; // one launch = iteration m < 2 of resblocks[n]
    PackedGroup mrf_group[4];      // one launch = last iteration of all ResBlocks of a stage + the MRF average
    int pair_cg = 2;               // CTAs per tcgen05 group of the fused kernel
    // last forward (debug taps, launch count)
    Plan last;
    void* last_ws = nullptr;
    int last_launches = 0;
    bool last_fused_post = false;  // the last forward ran conv_post inside the last MRF launch ("mrf_act.<last>" was never written)
    // optional per-launch device timing (vo2w_generator_set_timing): events around every launch of a forward
    bool timing = false;
    std::vector<cudaEvent_t> ev;
    std::vector<int> ev_kind;     // kind of the launch that ENDS at event i + 1
};

namespace {

int idx_conv_pre() { return 0; }
int idx_ups(int i) { return 1 + i; }
int idx_rb(const vo2w_generator* g, int n, int grp, int m) { return 1 + g->n_stage + n * 6 + grp * 3 + m; }
int idx_conv_post(const vo2w_generator* g) { return 1 + g->n_stage + g->n_stage * g->nk * 6; }

int stage_channels(const vo2w_generator* g, int i) { return g->cfg.upsample_initial_channel >> (i + 1); }

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

Plan make_plan(const vo2w_generator* g, int B, int T) {
    Plan p;
    p.B = B;
    p.T = T;
    size_t off = 0;
    auto take = [&](size_t bytes) {
        size_t o = off;
        off = align_up(off + bytes, 256);
        return o;
    };
    const int C0 = g->cfg.upsample_initial_channel;
    long long L = T;
    if (g->precision == VO2W_PREC_BF16) {
        const long long Lp0 = vo2w_padded_len(T);
        p.melb = take((size_t)B * 80 * Lp0 * 2);
        p.a0 = take((size_t)B * C0 * Lp0 * 2);
        for (int i = 0; i < g->n_stage; ++i) {
            StageBufs& s = p.st[i];
            L *= g->cfg.upsample_rates[i];
            s.C = stage_channels(g, i);
            s.L = L;
            s.Lp = vo2w_padded_len(L);
            const size_t bfb = (size_t)B * s.C * s.Lp * 2;
            s.xa = take(bfb);
            for (int j = 0; j < g->nk; ++j)
                for (int m = 0; m < 2; ++m) {
                    s.ra[j][m] = take(bfb);
                }
            s.mrf = take(bfb);
        }
    } else {
        p.c0_f32 = take((size_t)B * C0 * T * 4);
        for (int i = 0; i < g->n_stage; ++i) {
            StageBufs& s = p.st[i];
            L *= g->cfg.upsample_rates[i];
            s.C = stage_channels(g, i);
            s.L = L;
            const size_t f32b = (size_t)B * s.C * L * 4;
            s.x_f32 = take(f32b);
            s.rr = take(f32b);
            s.tt = take(f32b);
            s.xs = take(f32b);
        }
    }
    long long hop = 1;
    for (int i = 0; i < g->n_stage; ++i) hop *= g->cfg.upsample_rates[i];
    p.i16 = take((size_t)B * T * hop * 2);
    p.total = off;
    return p;
}

int build_layer_table(vo2w_generator* g, size_t n_floats) {
    const vo2w_gen_config& c = g->cfg;
    size_t off = 0;
    auto add = [&](int d0, int d1, int k, int nbias) {
        LayerW l;
        l.w_off = off;
        l.d0 = d0;
        l.d1 = d1;
        l.k = k;
        off += (size_t)d0 * d1 * k;
        l.b_off = off;
        off += nbias;
        g->lw.push_back(l);
    };
    const int C0 = c.upsample_initial_channel;
    add(C0, 80, 7, C0);                                                     // conv_pre (models.py:118-120)
    for (int i = 0; i < g->n_stage; ++i)                                    // ups (models.py:123-135)
        add(C0 >> i, C0 >> (i + 1), c.upsample_kernel_sizes[i], C0 >> (i + 1));
    for (int i = 0; i < g->n_stage; ++i)                                    // resblocks (models.py:137-143)
        for (int j = 0; j < g->nk; ++j)
            for (int grp = 0; grp < 2; ++grp)
                for (int m = 0; m < 3; ++m) {
                    const int ch = stage_channels(g, i);
                    add(ch, ch, c.resblock_kernel_sizes[j], ch);
                }
    add(1, stage_channels(g, g->n_stage - 1), 7, 1);                        // conv_post (models.py:145)
    VO2W_REQUIRE(off == n_floats, "generator_create: weight blob has %zu floats, config needs %zu", n_floats, off);
    return VO2W_OK;
}

}  // namespace

namespace {

// ConvTranspose1d(stride s, kernel ks = 2 s, padding (ks - s) / 2) as a Conv1d with 3 taps onto s * Cout channels:
//   out[l * s + ph] = sum_t x[l + floor((ph + p) / s) - t] . w[:, :, (ph + p) % s + t * s],  t = 0 .. ks / s - 1,
// so with channel ph * Cout + co the time-major result [L, s * Cout] IS the upsampled [L * s, Cout] tensor.
// w [Cin, Cout, ks] (ConvTranspose1d layout) -> weff [s * Cout, Cin, 3] (Conv1d layout), zero where a tap is unused.
__global__ void __launch_bounds__(256) ups_as_conv_kernel(const float* __restrict__ w, int Cin, int Cout, int ks, int s,
                                                          float* __restrict__ weff) {
    const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
    if (e >= (long long)s * Cout * Cin * 3) return;
    const int tau = (int)(e % 3), ci = (int)((e / 3) % Cin), ce = (int)(e / (3LL * Cin));
    const int ph = ce / Cout, co = ce % Cout, p = (ks - s) / 2;
    const int t = (ph + p) / s - (tau - 1), j = (ph + p) % s + t * s;
    weff[e] = (t >= 0 && j < ks) ? w[((size_t)ci * Cout + co) * ks + j] : 0.f;
}

// conv_post of the fp32 mode on time-major input: wav[b,l] = tanh(bias + sum_{j,c} w[c,j] leaky_relu(x[b,l+j-pad,c], slope)).
// One output channel, so no tensor-pipe tile: a CTA stages 256 + k - 1 frames (activated, rows padded to Cin + 1 floats
// against bank conflicts) and every thread reduces its own frame window.
__global__ void __launch_bounds__(256) post_conv_rows_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                             const float* __restrict__ bias, float* __restrict__ wav, int L, int Cin,
                                                             int k, float slope) {
    extern __shared__ float psm[];
    const int pad = (k - 1) / 2, rows = 256 + k - 1, ld = Cin + 1;
    float* xs = psm;                    // [rows][ld]
    float* ws = psm + rows * ld;        // [k][Cin]
    const int b = blockIdx.y, l0 = blockIdx.x * 256;
    const float* xb = x + (size_t)b * L * Cin;
    for (int e = threadIdx.x; e < rows * Cin; e += 256) {
        const int r = e / Cin, c = e - r * Cin, l = l0 - pad + r;
        xs[r * ld + c] = (l >= 0 && l < L) ? lrelu(xb[(size_t)l * Cin + c], slope) : 0.f;
    }
    for (int e = threadIdx.x; e < k * Cin; e += 256) { const int j = e / Cin, c = e - j * Cin; ws[e] = w[c * k + j]; }   // w [1,Cin,k]
    __syncthreads();
    const int l = l0 + threadIdx.x;
    if (l >= L) return;
    float acc = bias[0];
    for (int j = 0; j < k; ++j) {
        const float* xr = xs + (threadIdx.x + j) * ld;
        const float* wr = ws + j * Cin;
        for (int c = 0; c < Cin; ++c) acc = fmaf(wr[c], xr[c], acc);
    }
    wav[(size_t)b * L + l] = tanhf(acc);
}

int pack_fp32_mode(vo2w_generator* g) {
    const vo2w_gen_config& cfg = g->cfg;
    size_t total = 0, bias_total = 0, scratch = 0;
    g->p32_off.assign(g->lw.size(), 0);
    for (size_t i = 0; i < g->lw.size(); ++i) {
        const LayerW& l = g->lw[i];
        const bool up = (int)i >= idx_ups(0) && (int)i < idx_ups(g->n_stage);
        g->p32_off[i] = total;
        if ((int)i == idx_conv_post(g)) continue;                 // Cout = 1: stays on the CUDA-core kernel
        if (up) {
            const int s = cfg.upsample_rates[i - idx_ups(0)];
            VO2W_REQUIRE(l.k == 2 * s && tf32_conv_supported(s * l.d1, l.d0, 3), "fp32 mode: upsampler %zu (k %d, stride %d, %d -> %d) unsupported",
                         i - idx_ups(0), l.k, s, l.d0, l.d1);
            total += tf32_conv_packed_floats(s * l.d1, l.d0, 3);
            scratch = std::max(scratch, (size_t)s * l.d1 * l.d0 * 3);
            g->ups_bias_off[i - idx_ups(0)] = bias_total;
            bias_total += (size_t)s * l.d1;
        } else {
            VO2W_REQUIRE(tf32_conv_supported(l.d0, l.d1, l.k), "fp32 mode: layer %zu (%d -> %d, k %d) unsupported", i, l.d1, l.d0, l.k);
            total += tf32_conv_packed_floats(l.d0, l.d1, l.k);
        }
    }
    float* tmp = nullptr;
    if (cudaMalloc(&g->p32, total * 4) != cudaSuccess || cudaMalloc(&g->ups_bias, bias_total * 4) != cudaSuccess ||
        cudaMalloc(&tmp, scratch * 4) != cudaSuccess) {
        if (tmp) cudaFree(tmp);
        vo2w_set_error("generator_create: allocating packed fp32-mode weights failed");
        return VO2W_E_CUDA;
    }
    int rc = VO2W_OK;
    for (size_t i = 0; i < g->lw.size() && rc == VO2W_OK; ++i) {
        const LayerW& l = g->lw[i];
        const bool up = (int)i >= idx_ups(0) && (int)i < idx_ups(g->n_stage);
        if ((int)i == idx_conv_post(g)) continue;
        if (up) {
            const int st = (int)i - idx_ups(0), s = cfg.upsample_rates[st];
            const long long n = (long long)s * l.d1 * l.d0 * 3;
            ups_as_conv_kernel<<<(unsigned)vo2w_ceil_div(n, 256), 256>>>(g->w_dev + l.w_off, l.d0, l.d1, l.k, s, tmp);
            rc = tf32_conv_pack(tmp, s * l.d1, l.d0, 3, g->p32 + g->p32_off[i], 0);
            for (int ph = 0; ph < s && rc == VO2W_OK; ++ph)
                if (cudaMemcpy(g->ups_bias + g->ups_bias_off[st] + (size_t)ph * l.d1, g->w_dev + l.b_off, (size_t)l.d1 * 4,
                               cudaMemcpyDeviceToDevice) != cudaSuccess) rc = VO2W_E_CUDA;
            if (cudaDeviceSynchronize() != cudaSuccess) rc = VO2W_E_CUDA;       // tmp is reused by the next upsampler
        } else {
            rc = tf32_conv_pack(g->w_dev + l.w_off, l.d0, l.d1, l.k, g->p32 + g->p32_off[i], 0);
        }
    }
    if (cudaDeviceSynchronize() != cudaSuccess) rc = VO2W_E_CUDA;
    cudaFree(tmp);
    if (rc != VO2W_OK) vo2w_set_error("generator_create: packing fp32-mode weights failed");
    return rc;
}

}  // namespace

extern "C" int vo2w_generator_create(const vo2w_gen_config* cfg, const float* weights_host, size_t n_floats,
                                     int precision, vo2w_generator** out) {
    VO2W_REQUIRE(cfg && weights_host && out, "generator_create: null argument");
    VO2W_REQUIRE(precision == VO2W_PREC_FP32 || precision == VO2W_PREC_BF16, "generator_create: bad precision %d", precision);
    VO2W_REQUIRE(cfg->num_upsamples >= 1 && cfg->num_upsamples <= 4, "generator_create: num_upsamples %d", cfg->num_upsamples);
    VO2W_REQUIRE(cfg->num_kernels >= 1 && cfg->num_kernels <= 3, "generator_create: num_kernels %d", cfg->num_kernels);
    VO2W_REQUIRE(cfg->num_mels == 80, "generator_create: conv_pre takes 80 mel bins (models.py:119)");
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    vo2w_generator* g = new vo2w_generator();
    g->cfg = *cfg;
    g->precision = precision;
    g->n_stage = cfg->num_upsamples;
    g->nk = cfg->num_kernels;
    const char* dbg = getenv("VO2W_DEBUG_TAPS");
    g->debug_taps = dbg && dbg[0] == '1';
    g->n_floats = n_floats;
    rc = build_layer_table(g, n_floats);
    if (rc != VO2W_OK) { delete g; return rc; }
#define VO2W_TRY(expr) do { int _rc = (expr); if (_rc != VO2W_OK) { vo2w_generator_destroy(g); return _rc; } } while (0)
    if (cudaMalloc(&g->w_dev, n_floats * 4) != cudaSuccess ||
        cudaMemcpy(g->w_dev, weights_host, n_floats * 4, cudaMemcpyHostToDevice) != cudaSuccess) {
        vo2w_set_error("generator_create: uploading weights failed");
        vo2w_generator_destroy(g);
        return VO2W_E_CUDA;
    }
    if (precision == VO2W_PREC_BF16) {
        const float* W = weights_host;
        auto pc = [&](PackedLayer& L, int idx, int dil) {
            const LayerW& l = g->lw[idx];
            int r = pack_conv(L, W + l.w_off, W + l.b_off, l.d0, l.d1, l.k, dil);
            return r != VO2W_OK ? r : upload_layer(L);
        };
        VO2W_TRY(pc(g->conv_pre, idx_conv_pre(), 1));
        VO2W_TRY(pc(g->conv_post, idx_conv_post(g), 1));
        const char* cge = getenv("VO2W_PAIR_CG");
        g->pair_cg = (cge && cge[0] == '1') ? 1 : 2;
        const char* sce = getenv("VO2W_SINGLE_CONV");
        g->use_stream = !(sce && sce[0] == 'o');
        {
            const LayerW& l = g->lw[idx_conv_pre()];
            VO2W_TRY(su_pack_conv(g->s_pre, W + l.w_off, W + l.b_off, l.d0, l.d1, l.k, 1));
            VO2W_TRY(su_upload(g->s_pre));
            // conv_post can ride the last MRF launch when that launch is the 32-channel CTA-pair kernel and the layer
            // is the reference's Conv1d(32, 1, 7, padding=3).  VO2W_FUSE_POST=0 or debug taps keep it a launch of its own.
            const LayerW& lp = g->lw[idx_conv_post(g)];
            const char* fpe = getenv("VO2W_FUSE_POST");
            g->fuse_post = g->pair_cg == 2 && lp.d0 == 1 && lp.d1 == 32 && lp.k == PAIR_POST_K && stage_channels(g, g->n_stage - 1) == 32 &&
                           !(fpe && fpe[0] == '0') && !g->debug_taps;
            if (lp.d0 == 1 && lp.d1 == 32 && lp.k == PAIR_POST_K) {
                for (int c = 0; c < 32; ++c)
                    for (int j = 0; j < PAIR_POST_K; ++j) g->post_w[j * 32 + c] = W[lp.w_off + (size_t)c * PAIR_POST_K + j];
                g->post_b = W[lp.b_off];
            }
        }
        for (int i = 0; i < g->n_stage; ++i) {
            const LayerW& l = g->lw[idx_ups(i)];
            VO2W_TRY(pack_convT(g->ups[i], W + l.w_off, W + l.b_off, l.d0, l.d1, l.k, cfg->upsample_rates[i]));
            VO2W_TRY(upload_layer(g->ups[i]));
            VO2W_TRY(su_pack_convT(g->s_ups[i], W + l.w_off, W + l.b_off, l.d0, l.d1, l.k, cfg->upsample_rates[i]));
            VO2W_TRY(su_upload(g->s_ups[i]));
            const int ch = stage_channels(g, i);
            for (int j = 0; j < g->nk; ++j) {
                const int n = i * g->nk + j;
                for (int m = 0; m < 3; ++m) {
                    const LayerW& l1 = g->lw[idx_rb(g, n, 0, m)];
                    const LayerW& l2 = g->lw[idx_rb(g, n, 1, m)];
                    VO2W_TRY(pack_pair(g->pairs[n][m], W + l1.w_off, W + l1.b_off, W + l2.w_off, W + l2.b_off, ch,
                                       cfg->resblock_kernel_sizes[j], cfg->resblock_dilation_sizes[j][m], g->pair_cg));
                }
                for (int m = 0; m < 2; ++m) {
                    const PackedPair* one[1] = {&g->pairs[n][m]};
                    VO2W_TRY(make_group(g->iter_group[n][m], one, 1));
                }
            }
            const PackedPair* last[3] = {nullptr, nullptr, nullptr};
            for (int j = 0; j < g->nk; ++j) last[j] = &g->pairs[i * g->nk + j][2];
            VO2W_TRY(make_group(g->mrf_group[i], last, g->nk));
        }
    }
    if (precision == VO2W_PREC_FP32) VO2W_TRY(pack_fp32_mode(g));
#undef VO2W_TRY
    *out = g;
    return VO2W_OK;
}

extern "C" int vo2w_generator_destroy(vo2w_generator* g) {
    if (!g) return VO2W_OK;
    if (g->w_dev) cudaFree(g->w_dev);
    if (g->p32) cudaFree(g->p32);
    if (g->ups_bias) cudaFree(g->ups_bias);
    for (cudaEvent_t e : g->ev) cudaEventDestroy(e);
    free_layer(g->conv_pre);
    free_layer(g->conv_post);
    su_free(g->s_pre);
    for (int i = 0; i < 4; ++i) { free_layer(g->ups[i]); su_free(g->s_ups[i]); }
    for (int i = 0; i < 4; ++i) free_group(g->mrf_group[i]);
    for (int n = 0; n < 12; ++n) {
        for (int m = 0; m < 2; ++m) free_group(g->iter_group[n][m]);
        for (int m = 0; m < 3; ++m) free_pair(g->pairs[n][m]);
    }
    delete g;
    return VO2W_OK;
}

extern "C" size_t vo2w_generator_workspace_bytes(const vo2w_generator* g, int B, int T) {
    if (!g || B <= 0 || T <= 0) return 0;
    return make_plan(g, B, T).total;
}

extern "C" int vo2w_generator_last_launches(const vo2w_generator* g) { return g ? g->last_launches : 0; }

extern "C" int vo2w_generator_set_timing(vo2w_generator* g, int on) {
    VO2W_REQUIRE(g, "generator_set_timing: null handle");
    g->timing = on != 0;
    return VO2W_OK;
}

extern "C" int vo2w_generator_get_timing(vo2w_generator* g, float* ms, int32_t* kind, int capacity) {
    VO2W_REQUIRE(g && ms && kind, "generator_get_timing: null argument");
    VO2W_REQUIRE(g->timing && (int)g->ev.size() > g->last_launches, "generator_get_timing: no timed forward has run");
    const int n = g->last_launches;
    VO2W_REQUIRE(n <= capacity, "generator_get_timing: %d launches, capacity %d", n, capacity);
    VO2W_CUDA_OK(cudaEventSynchronize(g->ev[n]));
    for (int i = 0; i < n; ++i) {
        VO2W_CUDA_OK(cudaEventElapsedTime(&ms[i], g->ev[i], g->ev[i + 1]));
        kind[i] = g->ev_kind[i];
    }
    return n;
}

// ---------------------------------------------------------------------------------------
// forward schedules
// ---------------------------------------------------------------------------------------
namespace {

// The bf16 kernels write rows [0, L) of an operand buffer [B][C/8][Lp][8] and rely on every other row -- the 32 pad rows
// in front, everything behind sample L -- reading as zero (= the convolutions' zero padding).  On a workspace that has not
// run this (B, T) before, those rows are cleared by ONE launch over all 34 buffers (~0.3 % of the workspace; clearing the
// whole workspace cost 0.6 ms of HBM time at B = 128 whenever a serving loop changed its batch shape).
#define ZP_MAX_BUFS 40
struct ZeroPads {
    int n;
    unsigned long long off[ZP_MAX_BUFS];     // byte offset of the buffer inside the workspace
    int segs[ZP_MAX_BUFS];                   // B * C / 8 segments of Lp rows
    int L[ZP_MAX_BUFS], Lp[ZP_MAX_BUFS];
};

__global__ void __launch_bounds__(256) zero_pad_rows_kernel(uint8_t* ws, const __grid_constant__ ZeroPads z) {
    const int buf = blockIdx.y;
    if (buf >= z.n) return;
    const int L = z.L[buf], Lp = z.Lp[buf], tail0 = VO2W_PAD_ROWS + L, n_pad = VO2W_PAD_ROWS + (Lp - tail0);
    uint4* base = reinterpret_cast<uint4*>(ws + z.off[buf]);
    const long long total = (long long)z.segs[buf] * n_pad;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const long long seg = i / n_pad;
        const int r = (int)(i - seg * n_pad);
        const int row = r < VO2W_PAD_ROWS ? r : tail0 + (r - VO2W_PAD_ROWS);
        base[seg * Lp + row] = make_uint4(0u, 0u, 0u, 0u);
    }
}

int zero_pad_rows(const vo2w_generator* g, const Plan& p, void* workspace, cudaStream_t st) {
    ZeroPads z;
    z.n = 0;
    auto add = [&](size_t off, int C, long long L, long long Lp) {
        z.off[z.n] = off; z.segs[z.n] = p.B * (C / 8); z.L[z.n] = (int)L; z.Lp[z.n] = (int)Lp;
        ++z.n;
    };
    add(p.melb, 80, p.T, vo2w_padded_len(p.T));
    add(p.a0, g->cfg.upsample_initial_channel, p.T, vo2w_padded_len(p.T));
    for (int i = 0; i < g->n_stage; ++i) {
        const StageBufs& s = p.st[i];
        add(s.xa, s.C, s.L, s.Lp);
        for (int j = 0; j < g->nk; ++j)
            for (int m = 0; m < 2; ++m) add(s.ra[j][m], s.C, s.L, s.Lp);
        add(s.mrf, s.C, s.L, s.Lp);
    }
    VO2W_REQUIRE(z.n <= ZP_MAX_BUFS, "zero_pad_rows: %d buffers", z.n);
    zero_pad_rows_kernel<<<dim3(64, (unsigned)z.n), 256, 0, st>>>((uint8_t*)workspace, z);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

// kind: 0 = layout / conversion helper, 1 = single convolution (umma_conv_kernel / fp32 kernels), 2 = fused ResBlock iterations
int mark_launch(vo2w_generator* g, int kind, cudaStream_t st) {
    ++g->last_launches;
    if (!g->timing) return VO2W_OK;
    while ((int)g->ev.size() <= g->last_launches) {
        cudaEvent_t e;
        VO2W_CUDA_OK(cudaEventCreate(&e));
        g->ev.push_back(e);
    }
    g->ev_kind.resize(g->ev.size(), 0);
    g->ev_kind[g->last_launches - 1] = kind;
    VO2W_CUDA_OK(cudaEventRecord(g->ev[g->last_launches], st));
    return VO2W_OK;
}
#define VO2W_RUNK(kind, expr) do { int _rc = (expr); if (_rc != VO2W_OK) return _rc; _rc = mark_launch(g, kind, st); if (_rc != VO2W_OK) return _rc; } while (0)
#define VO2W_RUN(expr) VO2W_RUNK(1, expr)

// NVTX ranges (header-only nvtx3: a no-op costing a function-pointer check unless a profiler is attached) around a forward
// and its stages, so that `ncu --nvtx --nvtx-include "vo2w.stage2/"` or a timeline tool can pick a stage out of the 34 launches.
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
static const char* const kStageRange[4] = {"vo2w.stage0", "vo2w.stage1", "vo2w.stage2", "vo2w.stage3"};

// Generator.forward (models.py:149-165) on the tcgen05 path.
int forward_bf16(vo2w_generator* g, const Plan& p, const float* mel, float* wav, short* wav_i16, float i16_scale, uint8_t* ws,
                 cudaStream_t st) {
    const int B = p.B, T = p.T;
    auto bf = [&](size_t off) { return (__nv_bfloat16*)(ws + off); };
    const bool fused_post = g->fuse_post;
    g->last_fused_post = fused_post;

    // mel [B,80,T] fp32 -> operand buffer
    VO2W_RUNK(0, pack_ncl_to_operand(mel, bf(p.melb), B, 80, T, (int)vo2w_padded_len(T), 1.0f, st));
    {   // x = conv_pre(mel); the only consumer is leaky_relu(0.1) -> ups[0]
        NvtxRange r("vo2w.conv_pre");
        if (g->use_stream) {
            VO2W_RUN(su_launch(g->s_pre, bf(p.melb), bf(p.a0), nullptr, 0.1f, B, T, st));
        } else {
            const PackedLayer* L[1] = {&g->conv_pre};
            const __nv_bfloat16* in[1] = {bf(p.melb)};
            EpilogueSpec ep;
            ep.out_bf16 = bf(p.a0);
            ep.out_slope = 0.1f;
            VO2W_RUN(launch_layers(L, in, 1, B, T, ep, st));
        }
    }
    const __nv_bfloat16* stage_in = bf(p.a0);
    long long L_in = T;
    for (int i = 0; i < g->n_stage; ++i) {
        const StageBufs& s = p.st[i];
        NvtxRange r(kStageRange[i & 3]);                 // ups[i] + the stage's three ResBlocks + MRF (+ conv_post)
        if (g->use_stream) {   // x = ups[i](leaky_relu(x)): polyphase implicit GEMM on persistent CTA pairs
            VO2W_RUN(su_launch(g->s_ups[i], stage_in, bf(s.xa), nullptr, 0.1f, B, (int)L_in, st));
        } else {
            const PackedLayer* L[1] = {&g->ups[i]};
            const __nv_bfloat16* in[1] = {stage_in};
            EpilogueSpec ep;
            ep.out_bf16 = bf(s.xa);
            ep.out_slope = 0.1f;
            VO2W_RUN(launch_layers(L, in, 1, B, (int)L_in, ep, st));
        }
        // F.leaky_relu(x) before conv_post uses the default slope 0.01 (models.py:161)
        const float next_slope = i + 1 < g->n_stage ? 0.1f : 0.01f;
        for (int j = 0; j < g->nk; ++j) {
            const int n = i * g->nk + j;
            for (int m = 0; m < 2; ++m) {   // one ResBlock iteration per launch (models.py:97-103)
                PairIO io;
                io.xa[0] = m == 0 ? bf(s.xa) : bf(s.ra[j][0]);
                io.out_act = bf(s.ra[j][m]);
                io.out_slope = 0.1f;
                VO2W_RUNK(2, pair_launch(g->iter_group[n][m], io, B, (int)s.L, st));
            }
        }
        {   // last iteration of every ResBlock + xs = sum_j resblocks[j](x); x = xs / num_kernels (models.py:154-160)
            // in ONE launch: the three conv2 accumulate in TMEM, the epilogue adds the three residual streams
            PairIO io;
            for (int j = 0; j < g->nk; ++j) {
                io.xa[j] = bf(s.ra[j][1]);
            }
            io.out_act = bf(s.mrf);
            io.out_slope = next_slope;
            io.scale = 1.0f / (float)g->nk;
            if (fused_post && i + 1 == g->n_stage) {
                // ... and wav = tanh(conv_post(leaky_relu(x, 0.01))) (models.py:161-163) in the same launch: the
                // 32-channel tensor x is never written to HBM
                io.out_act = nullptr;
                io.post_w = g->post_w;
                io.post_b = g->post_b;
                io.wav = wav;
                io.wav_i16 = wav_i16;
                io.i16_scale = i16_scale;
            }
            VO2W_RUNK(2, pair_launch(g->mrf_group[i], io, B, (int)s.L, st));
        }
        stage_in = bf(s.mrf);
        L_in = s.L;
    }
    if (!fused_post) {   // wav = tanh(conv_post(leaky_relu(x)))
        const PackedLayer* L[1] = {&g->conv_post};
        const __nv_bfloat16* in[1] = {stage_in};
        EpilogueSpec ep;
        ep.out_wave = wav;
        VO2W_RUN(launch_layers(L, in, 1, B, (int)L_in, ep, st));
    }
    return VO2W_OK;
}

// fp32 mode: the same schedule on the tensor pipe with fp32-level accuracy (conv_tf32.cu: kind::tf32, hi/lo operand
// split), time-major [B,L,C] tensors.  leaky-relu rides the activation staging, bias / residual / MRF sum / average
// the epilogue; a ConvTranspose1d is a 3-tap convolution onto stride x Cout phase-major channels, whose time-major
// result is the upsampled tensor as it stands.  conv_post (one output channel) runs on the CUDA-core kernel.
int forward_fp32(vo2w_generator* g, const Plan& p, const float* mel, float* wav, uint8_t* ws, cudaStream_t st) {
    const int B = p.B, T = p.T;
    auto f32 = [&](size_t off) { return (float*)(ws + off); };
    auto Bi = [&](int idx) { return g->w_dev + g->lw[idx].b_off; };
    auto P = [&](int idx) { return g->p32 + g->p32_off[idx]; };
    auto conv = [&](const float* x, int idx, const float* bias, const float* res, const float* acc, float* out, int L, int Cin,
                    int Cout, int k, int dil, float in_slope, float out_scale) {
        Tf32Conv c{x, P(idx), bias, res, out, B, L, Cin, Cout, k, 0};
        c.dil = dil; c.in_slope = in_slope; c.acc = acc; c.out_scale = out_scale;
        return tf32_conv1d(c, st);
    };
    {
        // mel [B,80,T] -> [B,T,80]; conv_pre
        const LayerW& l = g->lw[idx_conv_pre()];
        float* mel_t = f32(p.st[0].tt);
        VO2W_RUNK(0, rows_transpose_add(mel, nullptr, mel_t, B, l.d1, T, st));
        VO2W_RUN(conv(mel_t, idx_conv_pre(), Bi(0), nullptr, nullptr, f32(p.c0_f32), T, l.d1, l.d0, l.k, 1, 1.f, 1.f));
    }
    const float* x_in = f32(p.c0_f32);
    long long L_in = T;
    for (int i = 0; i < g->n_stage; ++i) {
        const StageBufs& s = p.st[i];
        {
            const int idx = idx_ups(i), up = g->cfg.upsample_rates[i];
            const LayerW& l = g->lw[idx];
            VO2W_RUN(conv(x_in, idx, g->ups_bias + g->ups_bias_off[i], nullptr, nullptr, f32(s.x_f32), (int)L_in, l.d0, up * l.d1, 3, 1,
                          0.1f, 1.f));
        }
        for (int j = 0; j < g->nk; ++j) {
            const int n = i * g->nk + j;
            const int k = g->cfg.resblock_kernel_sizes[j];
            for (int m = 0; m < 3; ++m) {
                const int d = g->cfg.resblock_dilation_sizes[j][m];
                const float* xin = m == 0 ? f32(s.x_f32) : f32(s.rr);
                int idx = idx_rb(g, n, 0, m);
                VO2W_RUN(conv(xin, idx, Bi(idx), nullptr, nullptr, f32(s.tt), (int)s.L, s.C, s.C, k, d, 0.1f, 1.f));
                idx = idx_rb(g, n, 1, m);
                if (m < 2) {
                    VO2W_RUN(conv(f32(s.tt), idx, Bi(idx), xin, nullptr, f32(s.rr), (int)s.L, s.C, s.C, k, 1, 0.1f, 1.f));
                } else {            // xs (+)= resblock_j(x); after the last one x = xs / num_kernels
                    const bool last = j == g->nk - 1 && g->nk > 1;
                    VO2W_RUN(conv(f32(s.tt), idx, Bi(idx), xin, j == 0 ? nullptr : f32(s.xs), f32(s.xs), (int)s.L, s.C, s.C, k, 1, 0.1f,
                                  last ? 1.0f / (float)g->nk : 1.f));
                }
            }
        }
        x_in = f32(s.xs);
        L_in = s.L;
    }
    {
        // conv_post (Cout = 1): leaky-relu 0.01 in, tanh out, straight from the time-major tensor
        const int idx = idx_conv_post(g);
        const LayerW& l = g->lw[idx];
        VO2W_REQUIRE(l.d0 == 1 && l.d1 <= 64 && l.k <= 11, "fp32 mode: conv_post %d -> %d, k %d unsupported", l.d1, l.d0, l.k);
        const size_t smem = ((size_t)(256 + l.k - 1) * (l.d1 + 1) + (size_t)l.k * l.d1) * 4;
        static bool attr[VO2W_MAX_DEV];
        const int cur = vo2w_cur_dev();
        if (!attr[cur]) {
            VO2W_CUDA_OK(cudaFuncSetAttribute(post_conv_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
            attr[cur] = true;
        }
        dim3 grid((unsigned)vo2w_ceil_div(L_in, 256), (unsigned)B);
        post_conv_rows_kernel<<<grid, 256, smem, st>>>(x_in, g->w_dev + l.w_off, Bi(idx), wav, (int)L_in, l.d1, l.k, 0.01f);
        VO2W_RUN(cudaGetLastError() == cudaSuccess ? VO2W_OK : VO2W_E_CUDA);
    }
    return VO2W_OK;
}

}  // namespace

namespace {
__global__ void wav_to_i16_kernel(const float* __restrict__ w, short* __restrict__ o, float scale, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[i] = (short)(int)(w[i] * scale);   // numpy astype("int16") of in-range floats truncates toward zero
}
}  // namespace

// Generator.forward with an optional int16 copy of the waveform, (int16)(wav * i16_scale) as utils/model.py:83-86 makes
// it: written by the last launch itself where conv_post rides the MRF launch, else by a conversion kernel.
static int forward_impl(vo2w_generator* g, const float* mel, int B, int T, float* wav, short* wav_i16, float i16_scale,
                        void* workspace, size_t workspace_bytes, int ws_fresh, void* stream) {
    VO2W_REQUIRE(g && mel && wav, "generator_forward: null argument");
    VO2W_REQUIRE(B > 0 && T > 0, "generator_forward: empty batch (B=%d T=%d)", B, T);
    VO2W_REQUIRE(B <= 65535, "generator_forward: batch %d exceeds grid.z", B);
    NvtxRange range(g->precision == VO2W_PREC_BF16 ? "vo2w.generator.forward.bf16" : "vo2w.generator.forward.fp32");
    const Plan p = make_plan(g, B, T);
    if (workspace_bytes < p.total || !workspace) {
        vo2w_set_error("generator_forward: workspace %zu bytes, need %zu", workspace_bytes, p.total);
        return VO2W_E_WORKSPACE;
    }
    VO2W_REQUIRE(((uintptr_t)workspace & 255) == 0, "generator_forward: workspace must be 256-byte aligned");
    VO2W_REQUIRE(((uintptr_t)mel & 15) == 0 && ((uintptr_t)wav & 15) == 0, "generator_forward: mel/wav must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    g->last = p;
    g->last_ws = workspace;
    g->last_launches = 0;
    if (g->timing) {
        if (g->ev.empty()) {
            cudaEvent_t e;
            VO2W_CUDA_OK(cudaEventCreate(&e));
            g->ev.push_back(e);
        }
        VO2W_CUDA_OK(cudaEventRecord(g->ev[0], st));
    }
    int rc;
    if (ws_fresh && g->precision == VO2W_PREC_BF16)   // pad / tail rows of the operand buffers must read as zero
        if ((rc = zero_pad_rows(g, p, workspace, st)) != VO2W_OK) return rc;
    if (g->precision == VO2W_PREC_BF16) {
        rc = forward_bf16(g, p, mel, wav, wav_i16, i16_scale, (uint8_t*)workspace, st);
        if (rc != VO2W_OK || g->last_fused_post) return rc;
    } else {
        g->last_fused_post = false;
        rc = forward_fp32(g, p, mel, wav, (uint8_t*)workspace, st);
        if (rc != VO2W_OK) return rc;
    }
    if (wav_i16) {
        long long hop = 1;
        for (int i = 0; i < g->n_stage; ++i) hop *= g->cfg.upsample_rates[i];
        const long long n_out = (long long)B * T * hop;
        wav_to_i16_kernel<<<(unsigned)vo2w_ceil_div(n_out, 256), 256, 0, st>>>(wav, wav_i16, i16_scale, n_out);
        VO2W_CUDA_OK(cudaGetLastError());
        ++g->last_launches;
    }
    return VO2W_OK;
}

extern "C" int vo2w_generator_forward(vo2w_generator* g, const float* mel, int B, int T, float* wav, void* workspace,
                                      size_t workspace_bytes, int ws_fresh, void* stream) {
    return forward_impl(g, mel, B, T, wav, nullptr, 0.f, workspace, workspace_bytes, ws_fresh, stream);
}

static int forward_host_impl(vo2w_generator* g, const float* mel_host, int B, int T, float* wav_host,
                             int16_t* wav_i16_host, float max_wav_value, void* workspace, size_t workspace_bytes,
                             int ws_fresh, void* mel_dev, void* wav_dev, void* stream, bool sync) {
    VO2W_REQUIRE(g && mel_host && mel_dev && wav_dev, "generator_forward_host: null argument");
    VO2W_REQUIRE(wav_host || wav_i16_host, "generator_forward_host: no output buffer");
    VO2W_REQUIRE(B > 0 && T > 0, "generator_forward_host: empty batch");
    cudaStream_t st = (cudaStream_t)stream;
    long long up = 1;
    for (int i = 0; i < g->n_stage; ++i) up *= g->cfg.upsample_rates[i];
    const long long n_out = (long long)B * T * up;
    VO2W_CUDA_OK(cudaMemcpyAsync(mel_dev, mel_host, (size_t)B * 80 * T * 4, cudaMemcpyHostToDevice, st));
    // int16 samples are staged in their own region of the workspace (Plan::i16), never over the operand buffers
    short* i16_dev = wav_i16_host ? (short*)((uint8_t*)workspace + make_plan(g, B, T).i16) : nullptr;
    int rc = forward_impl(g, (const float*)mel_dev, B, T, (float*)wav_dev, i16_dev, max_wav_value, workspace, workspace_bytes,
                          ws_fresh, stream);
    if (rc != VO2W_OK) return rc;
    if (wav_i16_host) VO2W_CUDA_OK(cudaMemcpyAsync(wav_i16_host, i16_dev, (size_t)n_out * 2, cudaMemcpyDeviceToHost, st));
    if (wav_host) VO2W_CUDA_OK(cudaMemcpyAsync(wav_host, wav_dev, (size_t)n_out * 4, cudaMemcpyDeviceToHost, st));
    if (sync) VO2W_CUDA_OK(cudaStreamSynchronize(st));
    return VO2W_OK;
}

extern "C" int vo2w_generator_forward_host(vo2w_generator* g, const float* mel_host, int B, int T, float* wav_host,
                                           int16_t* wav_i16_host, float max_wav_value, void* workspace,
                                           size_t workspace_bytes, int ws_fresh, void* mel_dev, void* wav_dev,
                                           void* stream) {
    return forward_host_impl(g, mel_host, B, T, wav_host, wav_i16_host, max_wav_value, workspace, workspace_bytes, ws_fresh,
                             mel_dev, wav_dev, stream, true);
}

extern "C" int vo2w_generator_forward_host_async(vo2w_generator* g, const float* mel_host, int B, int T, float* wav_host,
                                                 int16_t* wav_i16_host, float max_wav_value, void* workspace,
                                                 size_t workspace_bytes, int ws_fresh, void* mel_dev, void* wav_dev,
                                                 void* stream) {
    return forward_host_impl(g, mel_host, B, T, wav_host, wav_i16_host, max_wav_value, workspace, workspace_bytes, ws_fresh,
                             mel_dev, wav_dev, stream, false);
}

namespace {
struct TapRef { size_t off = 0; int C = 0; long long L = 0; bool is_bf16_buf = false; float inv_slope = 1.0f; };
// where the activation `tap` of the last forward lives in the workspace; false = this mode / schedule never writes it
bool find_tap(const vo2w_generator* g, const std::string& name, TapRef& t) {
    const Plan& p = g->last;
    const bool bf = g->precision == VO2W_PREC_BF16;
    if (name == "conv_pre" && !bf) {
        t.off = p.c0_f32; t.C = g->cfg.upsample_initial_channel; t.L = p.T;
        return true;
    }
    for (int i = 0; i < g->n_stage; ++i) {
        bool found = false;
        // "ups.i" in bf16 mode: the value recovered from the stream
        if (name == "ups." + std::to_string(i)) { t.off = bf ? p.st[i].xa : p.st[i].x_f32; found = true; t.is_bf16_buf = bf; t.inv_slope = bf ? 10.0f : 1.0f; }
        else if (name == "mrf." + std::to_string(i) && !bf) { t.off = p.st[i].xs; found = true; }
        else if (name == "mrf_act." + std::to_string(i) && bf) {
            if (i + 1 == g->n_stage && g->last_fused_post) return false;   // conv_post consumed it inside the MRF launch
            t.off = p.st[i].mrf; found = true; t.is_bf16_buf = true;
        }
        if (found) { t.C = p.st[i].C; t.L = p.st[i].L; return true; }
    }
    return false;
}
}  // namespace

extern "C" int vo2w_generator_has_tap(const vo2w_generator* g, const char* tap) {
    if (!g || !tap || !g->last_ws) return 0;
    TapRef t;
    return find_tap(g, std::string(tap), t) ? 1 : 0;
}

extern "C" long long vo2w_generator_debug_tap(vo2w_generator* g, const char* tap, float* out_host, size_t capacity) {
    VO2W_REQUIRE(g && tap && out_host, "debug_tap: null argument");
    VO2W_REQUIRE(g->last_ws, "debug_tap: no forward has run");
    const Plan& p = g->last;
    uint8_t* ws = (uint8_t*)g->last_ws;
    TapRef t;
    VO2W_REQUIRE(find_tap(g, std::string(tap), t), "debug_tap: unknown tap '%s' for this precision mode / schedule", tap);
    const size_t off = t.off;
    const int C = t.C;
    const long long L = t.L;
    const bool is_bf16_buf = t.is_bf16_buf;
    const float inv_slope = t.inv_slope;
    const size_t n = (size_t)p.B * C * L;
    VO2W_REQUIRE(n <= capacity, "debug_tap: need %zu floats, capacity %zu", n, capacity);
    VO2W_CUDA_OK(cudaDeviceSynchronize());
    if (is_bf16_buf) {
        float* tmp = nullptr;
        VO2W_CUDA_OK(cudaMalloc(&tmp, n * 4));
        int rc = unchunk_bf16((const __nv_bfloat16*)(ws + off), tmp, p.B, C, (int)L, (int)vo2w_padded_len(L), 0, inv_slope);
        if (rc == VO2W_OK && cudaMemcpy(out_host, tmp, n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = VO2W_E_CUDA;
        cudaFree(tmp);
        if (rc != VO2W_OK) return rc;
    } else {                       // fp32 mode keeps [B,L,C]: hand out the reference's [B,C,L]
        float* tmp = nullptr;
        VO2W_CUDA_OK(cudaMalloc(&tmp, n * 4));
        int rc = rows_transpose_add((const float*)(ws + off), nullptr, tmp, p.B, (int)L, C, 0);
        if (rc == VO2W_OK && cudaMemcpy(out_host, tmp, n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = VO2W_E_CUDA;
        cudaFree(tmp);
        if (rc != VO2W_OK) return rc;
    }
    return (long long)n;
}

// ---------------------------------------------------------------------------------------
// Known-answer probe: ONE reference layer through either kernel family, NCL fp32 in/out.
// transposed = 0: nn.Conv1d(Cin,Cout,k,dilation=d,padding=d(k-1)/2); 1: nn.ConvTranspose1d(Cin,Cout,k=2s,stride=s,padding=s/2)
// y = layer(leaky_relu(x, in_slope)) [+ res]
// ---------------------------------------------------------------------------------------
extern "C" int vo2w_debug_layer(int transposed, int use_umma, const float* w_host, const float* bias_host, int Cin,
                                int Cout, int k, int dil_or_stride, const float* x_dev, int B, int L, float in_slope,
                                const float* res_dev, float* out_dev, void* stream) {
    VO2W_REQUIRE(w_host && bias_host && x_dev && out_dev, "debug_layer: null argument");
    int rc = vo2w_device_check();
    if (rc != VO2W_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int s = transposed ? dil_or_stride : 1;
    const long long Lo = (long long)L * s;
    const size_t wn = (size_t)Cin * Cout * k;
    if (!use_umma) {
        float *w = nullptr, *b = nullptr;
        VO2W_CUDA_OK(cudaMalloc(&w, wn * 4));
        VO2W_CUDA_OK(cudaMalloc(&b, (size_t)Cout * 4));
        VO2W_CUDA_OK(cudaMemcpyAsync(w, w_host, wn * 4, cudaMemcpyHostToDevice, st));
        VO2W_CUDA_OK(cudaMemcpyAsync(b, bias_host, (size_t)Cout * 4, cudaMemcpyHostToDevice, st));
        SimtConv q;
        memset(&q, 0, sizeof(q));
        q.x = x_dev; q.w = w; q.bias = b; q.res = res_dev; q.out = out_dev;
        q.B = B; q.Cin = Cin; q.Cout = Cout; q.L = L; q.k = k; q.in_slope = in_slope; q.div = 1.f;
        if (transposed) { q.stride = s; q.dil = 1; q.pad = (k - s) / 2; rc = simt_convt1d(q, st); }
        else { q.stride = 1; q.dil = dil_or_stride; q.pad = q.dil * (k - 1) / 2; rc = simt_conv1d(q, st); }
        cudaStreamSynchronize(st);
        cudaFree(w);
        cudaFree(b);
        return rc;
    }
    VO2W_REQUIRE(res_dev == nullptr || !transposed, "debug_layer: residual only with Conv1d");
    if (use_umma == 2) {
        // stream_umma_kernel (persistent CTA pairs): the kernel behind conv_pre and the upsamplers of the bf16 mode
        VO2W_REQUIRE(res_dev == nullptr, "debug_layer: the stream kernel has no residual input");
        SuLayer S;
        rc = transposed ? su_pack_convT(S, w_host, bias_host, Cin, Cout, k, s) : su_pack_conv(S, w_host, bias_host, Cout, Cin, k, dil_or_stride);
        if (rc != VO2W_OK) return rc;
        rc = su_upload(S);
        if (rc != VO2W_OK) return rc;
        const long long Lp = vo2w_padded_len(L);
        __nv_bfloat16* xin = nullptr;
        float* of = nullptr;
        const size_t in_b = (size_t)B * S.Cin * Lp * 2, out_b = (size_t)B * S.Cout * Lo * 4;
        VO2W_CUDA_OK(cudaMalloc(&xin, in_b));
        VO2W_CUDA_OK(cudaMalloc(&of, out_b));
        VO2W_CUDA_OK(cudaMemsetAsync(xin, 0, in_b, st));
        rc = pack_ncl_to_operand(x_dev, xin, B, Cin, L, (int)Lp, in_slope, st);
        if (rc == VO2W_OK) rc = su_launch(S, xin, nullptr, of, 1.0f, B, L, st);
        if (rc == VO2W_OK) rc = unchunk_f32(of, out_dev, B, S.Cout, Cout, (int)Lo, st);
        cudaError_t e = cudaStreamSynchronize(st);
        if (rc == VO2W_OK && e != cudaSuccess) {
            vo2w_set_error("debug_layer: stream kernel failed: %s", cudaGetErrorString(e));
            rc = VO2W_E_CUDA;
        }
        cudaFree(xin);
        cudaFree(of);
        su_free(S);
        return rc;
    }
    PackedLayer Lr;
    rc = transposed ? pack_convT(Lr, w_host, bias_host, Cin, Cout, k, s) : pack_conv(Lr, w_host, bias_host, Cout, Cin, k, dil_or_stride);
    if (rc != VO2W_OK) return rc;
    rc = upload_layer(Lr);
    if (rc != VO2W_OK) return rc;
    const int Cin_p = Lr.Cin, Cout_p = (Cout + 15) / 16 * 16;
    const long long Lp = vo2w_padded_len(L);
    __nv_bfloat16* xin = nullptr;
    float *of = nullptr, *rf = nullptr;
    const size_t in_b = (size_t)B * Cin_p * Lp * 2, out_b = (size_t)B * Cout_p * Lo * 4;
    VO2W_CUDA_OK(cudaMalloc(&xin, in_b));
    VO2W_CUDA_OK(cudaMalloc(&of, out_b));
    VO2W_CUDA_OK(cudaMemsetAsync(xin, 0, in_b, st));
    rc = pack_ncl_to_operand(x_dev, xin, B, Cin, L, (int)Lp, in_slope, st);
    EpilogueSpec ep;
    ep.out_f32 = of;
    if (rc == VO2W_OK && res_dev) {
        VO2W_CUDA_OK(cudaMalloc(&rf, out_b));
        rc = chunk_f32(res_dev, rf, B, Cout, Cout_p, (int)Lo, st);
        ep.res[0] = rf;
        ep.n_res = 1;
    }
    // the packed layer reports padded Cout through n_cols*n_tiles; the epilogue indexes by the true Cout rounded to 16
    Lr.Cout = transposed ? Cout : Cout_p;
    const PackedLayer* Ls[1] = {&Lr};
    const __nv_bfloat16* ins[1] = {xin};
    if (rc == VO2W_OK) rc = launch_layers(Ls, ins, 1, B, L, ep, st);
    if (rc == VO2W_OK) rc = unchunk_f32(of, out_dev, B, Cout_p, Cout, (int)Lo, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (rc == VO2W_OK && e != cudaSuccess) {
        vo2w_set_error("debug_layer: kernel failed: %s", cudaGetErrorString(e));
        rc = VO2W_E_CUDA;
    }
    cudaFree(xin);
    cudaFree(of);
    if (rf) cudaFree(rf);
    free_layer(Lr);
    return rc;
}
