/*
 * vo2w.h -- C ABI of libvo2w.so: the B200 (sm_100a) synthesis hot path of
 * sarulab-speech/visual-onoma-to-wave (HiFi-GAN V1 Generator + LengthRegulator).
 *
 * The reference is pure Python/PyTorch and has no FFI of its own; each entry point
 * below names the reference interface it replaces (file:line under /root/reference).
 * INTEGRATION.md shows the ctypes binding a reference maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types.  `stream` is a cudaStream_t passed
 *     as void* (0 = legacy default stream).
 *   - device pointers unless the name says host.  The caller owns every buffer; the
 *     library owns only the packed weights inside a vo2w_generator handle.
 *   - return 0 on success, a negative VO2W_E_* code otherwise (the Python mirror raises
 *     RuntimeError, which is the reference's only error convention).  vo2w_last_error()
 *     returns a thread-local message for the last failure.
 *   - there is no CPU fallback: without a CUDA device every compute call fails.
 *   - threads: a handle's packed weights are read-only after create, but forward() also records
 *     per-call state in the handle (launch count, timing events, the plan debug_tap reads), and a
 *     workspace belongs to one in-flight forward.  So calls on ONE handle must be serialised by the
 *     caller (the reference's callers are single-threaded Python); different
 *     handles, and different devices, are independent.  Kernel attributes are set per device.
 */
#ifndef VO2W_H
#define VO2W_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VO2W_OK            0
#define VO2W_E_INVALID    -1   /* bad argument (shape, dtype, alignment, null pointer) */
#define VO2W_E_CUDA       -2   /* a CUDA call or kernel launch failed                */
#define VO2W_E_WORKSPACE  -3   /* workspace too small                                */
#define VO2W_E_NODEVICE   -4   /* no sm_100 device                                   */

/* duration dtypes accepted by the LengthRegulator (reference: float32 from the duration
 * predictor, scripts/model/modules.py:110-113; float/int64 targets, utils/tools.py:48-49) */
#define VO2W_DUR_I32 0
#define VO2W_DUR_I64 1
#define VO2W_DUR_F32 2
#define VO2W_DUR_F64 3

/* precision modes of the Generator */
#define VO2W_PREC_FP32 0   /* tcgen05 kind::tf32, hi/lo operand split: max-abs-err <= 1e-4 */
#define VO2W_PREC_BF16 1   /* tcgen05 kind::f16, bf16 operands, fp32 accumulate: SNR >= 40 dB */

/* HiFi-GAN hyper-parameters read by Generator.__init__ (scripts/hifigan/models.py:116-143,
 * values from scripts/hifigan/config.json:11-15). */
typedef struct vo2w_gen_config {
    int32_t num_upsamples;              /* <= 4 */
    int32_t upsample_rates[4];
    int32_t upsample_kernel_sizes[4];
    int32_t upsample_initial_channel;
    int32_t num_kernels;                /* <= 3 */
    int32_t resblock_kernel_sizes[3];
    int32_t resblock_dilation_sizes[3][3];
    int32_t num_mels;                   /* conv_pre hard-codes 80 (models.py:119) */
} vo2w_gen_config;

typedef struct vo2w_generator vo2w_generator;

const char* vo2w_last_error(void);
int  vo2w_version(void);
/* 0 when a compute-capability-10.x device is current, else VO2W_E_NODEVICE. */
int  vo2w_device_check(void);

/* ---- HiFi-GAN Generator ------------------------------------------------------------
 * Replaces hifigan.Generator.__init__ + load_state_dict + remove_weight_norm + .to(device)
 * (scripts/hifigan/models.py:112-147,167-174; scripts/utils/model.py:56-68).
 * `weights_host` is the FOLDED state-dict (weight_norm already removed, 156 tensors)
 * concatenated as fp32 in state-dict order: conv_pre.{weight,bias}, ups.i.{weight,bias},
 * resblocks.n.convs1.m.{weight,bias} (m=0..2), resblocks.n.convs2.m.{weight,bias},
 * conv_post.{weight,bias} -- i.e. for every layer weight first, then bias, layers in
 * the order Generator registers them.  `n_floats` is checked against the config. */
int vo2w_generator_create(const vo2w_gen_config* cfg, const float* weights_host, size_t n_floats,
                          int precision, vo2w_generator** out);
int vo2w_generator_destroy(vo2w_generator* g);

/* Bytes of device scratch forward() needs for a [B,80,T] batch. */
size_t vo2w_generator_workspace_bytes(const vo2w_generator* g, int B, int T);

/* Replaces Generator.forward (scripts/hifigan/models.py:149-165):
 * mel fp32 [B,80,T] contiguous  ->  wav fp32 [B,1,T*prod(upsample_rates)].
 * `ws_fresh` != 0 tells the library the workspace has not been used with this (B,T)
 * before (it then zeroes the halo rows it relies on). */
int vo2w_generator_forward(vo2w_generator* g, const float* mel, int B, int T, float* wav,
                           void* workspace, size_t workspace_bytes, int ws_fresh, void* stream);

/* Same call with HOST buffers (the vocoder_infer seam, scripts/utils/model.py:73-98:
 * host mel in, `.cpu().numpy()` out).  Copies through the caller's pinned or pageable
 * memory on `stream` and synchronises it before returning.  wav_i16 (optional) receives
 * (wav * max_wav_value) truncated to int16 as utils/model.py:83-86 does. */
int vo2w_generator_forward_host(vo2w_generator* g, const float* mel_host, int B, int T,
                                float* wav_host, int16_t* wav_i16_host, float max_wav_value,
                                void* workspace, size_t workspace_bytes, int ws_fresh,
                                void* mel_dev, void* wav_dev, void* stream);

/* Same, but returns once the copies and kernels are queued on `stream` (no synchronisation): the caller
 * overlaps the host<->device copies of one sub-batch with the kernels of another by alternating two
 * streams, each with its own workspace and staging buffers, and synchronises the stream before reading
 * wav_host / reusing mel_host.  Host buffers should be pinned for the copies to be asynchronous. */
int vo2w_generator_forward_host_async(vo2w_generator* g, const float* mel_host, int B, int T,
                                      float* wav_host, int16_t* wav_i16_host, float max_wav_value,
                                      void* workspace, size_t workspace_bytes, int ws_fresh,
                                      void* mel_dev, void* wav_dev, void* stream);

/* Per-layer probe used by the known-answer tests: copies the fp32 activation named
 * `tap` ("conv_pre", "ups.0", "resblocks.3", "mrf.1", ...) of the LAST forward into
 * out_host as NCL [B,C,L].  Returns the element count, or a negative error. */
long long vo2w_generator_debug_tap(vo2w_generator* g, const char* tap, float* out_host,
                                   size_t capacity);
/* 1 when the last forward left the activation `tap` in its workspace (a tensor that a fused launch never
 * writes to HBM has no tap), else 0. */
int vo2w_generator_has_tap(const vo2w_generator* g, const char* tap);

/* Known-answer probe: ONE reference layer through one kernel family, NCL fp32 in/out.
 * transposed = 0: nn.Conv1d(Cin, Cout, k, dilation=d, padding=d*(k-1)/2) (models.py:26-92)
 * transposed = 1: nn.ConvTranspose1d(Cin, Cout, k=2s, stride=s, padding=s/2) (models.py:123-135)
 * out = layer(leaky_relu(x, in_slope)) [+ res];  use_umma selects fp32 CUDA cores (0), the tcgen05 single-CTA kernel
 * (1) or the persistent tcgen05 CTA-pair kernel of the bf16 hot path (2; no residual).
 * w_host/bias_host in the torch layout, x/res/out on the device.  Synchronises `stream`. */
int vo2w_debug_layer(int transposed, int use_umma, const float* w_host, const float* bias_host, int Cin,
                     int Cout, int k, int dil_or_stride, const float* x_dev, int B, int L, float in_slope,
                     const float* res_dev, float* out_dev, void* stream);

/* Known-answer probe of the fused ResBlock-iteration kernel (scripts/hifigan/models.py:97-103, 154-160):
 *   y   = scale * sum_{s < n_seg} [ x'_s + conv2_s(leaky_relu(conv1_s(a_s), 0.1)) ],  a_s = bf16(leaky_relu(x_s, 0.1))
 *   act = bf16(leaky_relu(y, out_slope))   (the only tensor the kernel leaves in HBM)
 *   out = act >= 0 ? act : act / out_slope (optional: the value a consumer recovers from the stream)
 * where x'_s = a_s >= 0 ? a_s : 10 a_s is the residual as the kernel recovers it from its input stream.
 * conv1_s = Conv1d(C,C,ks[s],dilation=dils[s]), conv2_s = Conv1d(C,C,ks[s]), both "same"-padded, weights and
 * biases in the torch layout on the host (arrays of n_seg pointers); x_s / out / act NCL fp32 on the device.
 * n_seg = 1, scale = 1 is one ResBlock iteration; n_seg = num_kernels, scale = 1/num_kernels the MRF
 * average of a stage.
 * cg = CTAs per tcgen05 group (1, or 2 for cta_group::2 pairs). */
int vo2w_debug_pair(int n_seg, const float* const* w1_host, const float* const* b1_host,
                    const float* const* w2_host, const float* const* b2_host, int C, const int* ks,
                    const int* dils, int cg, const float* const* x_dev, int B, int L, float scale,
                    float out_slope, float* out_dev, float* act_dev, void* stream);

/* The same launch with conv_post + tanh on board, as the last stage of Generator.forward runs it
 * (scripts/hifigan/models.py:154-163): C = 32, CTA pairs;
 *   wav = tanh(Conv1d(32, 1, 7, padding=3)(leaky_relu(y, out_slope)))  with y as above, in fp32 (y never reaches HBM);
 * post_w_host [1,32,7] (torch layout).  wav_dev fp32 [B,L] and / or wav_i16_dev = (int16)(wav * i16_scale)
 * (scripts/utils/model.py:83-86). */
int vo2w_debug_pair_post(int n_seg, const float* const* w1_host, const float* const* b1_host,
                         const float* const* w2_host, const float* const* b2_host, const int* ks, const int* dils,
                         const float* const* x_dev, int B, int L, float scale, float out_slope,
                         const float* post_w_host, float post_b, float* wav_dev, int16_t* wav_i16_dev,
                         float i16_scale, void* stream);

/* Host-only probe of the packing of stream_umma_kernel (conv_pre and the upsamplers of the bf16 mode, use_umma = 2 in
 * vo2w_debug_layer): meta[0..15] = {Cin, Cout, up_s, KB, n_kb, NB, n_cols, n_tiles, n_items, halo_l, halo_r,
 * stage_bytes, w_elems, bias_elems, 0, 0}, then item_shift[12], tile_shift[16]; w_out receives the packed bf16 bits
 * [rank][tile][item][kb][KB/8][n_cols/2][8].  Pass w_out = NULL to query sizes. */
int vo2w_debug_stream_pack(int transposed, const float* w, const float* bias, int Cin, int Cout, int k,
                           int dil_or_stride, int32_t* meta, uint16_t* w_out, size_t w_capacity,
                           float* bias_out, size_t bias_capacity);

/* Host-only probe of the weight packing (no device needed): meta[0..11] = {Cin, Cout, n_cols,
 * n_tiles, phases_per_tile, halo_l, halo_r, n_items, w_tile_stride, up_s, w_elems, bias_elems},
 * then per item {row_shift, n_off, n_size, zero_init, w_off}; w_out receives the packed bf16 bits.
 * Pass w_out = NULL to query sizes.  The CPU tests replay the kernel's indexing from these. */
int vo2w_debug_pack(int transposed, const float* w, const float* bias, int Cin, int Cout, int k,
                    int dil_or_stride, int32_t* meta, uint16_t* w_out, size_t w_capacity,
                    float* bias_out, size_t bias_capacity);

/* Kernel launches issued by the last forward (bench.py's gpu_launches). */
int vo2w_generator_last_launches(const vo2w_generator* g);

/* Per-launch device timing of a forward (measurement aid of bench.py's roofline): with timing on, every forward
 * records a CUDA event on its stream around each launch.  get_timing waits for the last forward and fills
 * ms[i] = duration of launch i and kind[i] = 0 layout helper / 1 single convolution / 2 fused ResBlock
 * iterations (pair_umma_kernel); returns the number of launches or a negative error. */
int vo2w_generator_set_timing(vo2w_generator* g, int on);
int vo2w_generator_get_timing(vo2w_generator* g, float* ms, int32_t* kind, int capacity);

/* ---- mel_linear + PostNet -------------------------------------
 * Replaces PostNet.forward (scripts/transformer/Layers.py:126-137) in eval mode: x [B,L,n_mel] fp32 ->
 * 4 x tanh(BatchNorm1d(Conv1d k)) -> BatchNorm1d(Conv1d k) -> [B,L,n_mel].  `folded_host` holds, per convolution,
 * weight [Cout,Cin,k] then bias [Cout] with the eval-mode BatchNorm already folded in by the caller
 * (w' = w * g / sqrt(var + eps), b' = (b - mean) * g / sqrt(var + eps) + beta).  add_input != 0 returns
 * postnet(x) + x, the `postnet_output` of vTTS.forward (scripts/model/vtts.py:107).  Tensor pipe with fp32-level
 * accuracy (conv_tf32.cu) when n_mel % 16 == 0 and dim % 64 == 0, else fp32 CUDA cores. */
typedef struct vo2w_postnet vo2w_postnet;
int vo2w_postnet_create(int n_mel, int dim, int k, int n_conv, const float* folded_host, size_t n_floats,
                        vo2w_postnet** out);
int vo2w_postnet_destroy(vo2w_postnet* p);
size_t vo2w_postnet_workspace_bytes(const vo2w_postnet* p, int B, int L);
int vo2w_postnet_forward(vo2w_postnet* p, const float* x, int B, int L, float* out, int add_input,
                         void* workspace, size_t workspace_bytes, void* stream);

/* nn.Linear on the last axis (mel_linear, scripts/model/vtts.py:106): y[r,:] = W x[r,:] + b for `rows` rows;
 * x [rows,K], w [N,K], bias [N] or NULL, y [rows,N], all fp32 on the device.  `scratch` (2*N*K floats on the device,
 * or NULL) lets shapes the tensor-pipe kernel tiles (K % 8 == 0; N % 64 == 0 or N <= 128 and N % 16 == 0) run
 * there; without it, or for other shapes, an fp32 CUDA-core kernel does the work. */
int vo2w_linear_forward(const float* x, const float* w, const float* bias, long long rows, int K, int N,
                        float* y, float* scratch, void* stream);

/* ---- mel Decoder: the FFT-block stack ----------------------------
 * Replaces Decoder.forward (scripts/transformer/Models.py:168-197) in eval mode:
 *   dec = x + pos[:L];  per layer FFTBlock.forward (scripts/transformer/Layers.py:21-30) =
 *   MultiHeadAttention (SubLayers.py:31-62; key padding mask, softmax(q k^T / sqrt(d_k)) v, LayerNorm(fc + x)),
 *   masked_fill(mask, 0), PositionwiseFeedForward (SubLayers.py:95-104; LayerNorm(w_2(relu(w_1 x)) + x)),
 *   masked_fill(mask, 0).
 * x, out [B,L,d_model] fp32; pos [L,d_model] = the first L rows of the sinusoid table (position_enc, or a longer
 * table built the same way when L > max_seq_len, Models.py:175-181); mask [B,L] bytes, 1 = padding frame, or NULL.
 * Head size d_model / n_head must be 128 (the reference's 256 / 2).  `blob_host`, per layer: [w_qs|w_ks|w_vs]
 * [3D,D], their biases [3D], fc [D,D], bias, attention LayerNorm gamma, beta, w_1 [d_inner,D,k1], bias,
 * w_2 [D,d_inner,k2], bias, feed-forward LayerNorm gamma, beta.  d_inner % 64 == 0.  Linear / Conv1d layers on the
 * tensor pipe with fp32-level accuracy (conv_tf32.cu), attention and LayerNorm on fp32 CUDA cores. */
typedef struct vo2w_decoder vo2w_decoder;
int vo2w_decoder_create(int d_model, int n_head, int d_inner, int k1, int k2, int n_layers, const float* blob_host,
                        size_t n_floats, vo2w_decoder** out);
int vo2w_decoder_destroy(vo2w_decoder* d);
size_t vo2w_decoder_workspace_bytes(const vo2w_decoder* d, int B, int L);
int vo2w_decoder_forward(vo2w_decoder* d, const float* x, const float* pos, const uint8_t* mask, int B, int L,
                         float* out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- VarianceAdaptor pieces in front of LengthRegulator -----------
 * vo2w_varpred_*: replaces VariancePredictor.forward (scripts/model/modules.py:161-213) in eval mode:
 *   x [B,T,d_in] fp32 -> LayerNorm(relu(Conv1d k3)) x 2 -> Linear(F,1).squeeze(-1) -> masked_fill(mask, 0) -> out [B,T].
 *   mask [B,T] bytes (1 = padding token) or NULL.  kernel_size must be 3 (the reference pads the second convolution
 *   by 1 whatever the kernel size, modules.py:189).  `blob_host`: conv1d_1 weight [F,d_in,k], bias, layer_norm_1 gamma,
 *   beta, conv1d_2 weight [F,F,k], bias, layer_norm_2 gamma, beta, linear_layer weight [F], bias [1].
 * vo2w_variance_embed: replaces get_energy_embedding / get_kurtosis_embedding (modules.py:53-78) after the predictor:
 *   denorm != 0 (no target given): v = ((value * std + mean) * control - mean) / std, each step rounded to fp32 as
 *   the reference's tensor-scalar ops are; denorm == 0 (a target is given): v = value.  Then
 *   out[r,:] = x[r,:] + table[bucketize(v, bins)]  with torch.bucketize's right=False rule (first edge >= v);
 *   value_out (optional) receives v.  bins [n_edges] ascending, table [n_edges+1, D].
 * vo2w_duration_round: out = clamp(round(exp(log_duration) - 1) * d_control, min=0), round half to even
 *   (modules.py:110-113); the float durations feed vo2w_length_regulate (VO2W_DUR_F32) unchanged. */
typedef struct vo2w_varpred vo2w_varpred;
int vo2w_varpred_create(int d_in, int d_filter, int k, const float* blob_host, size_t n_floats, vo2w_varpred** out);
int vo2w_varpred_destroy(vo2w_varpred* p);
size_t vo2w_varpred_workspace_bytes(const vo2w_varpred* p, int B, int T);
int vo2w_varpred_forward(vo2w_varpred* p, const float* x, const uint8_t* mask, int B, int T, float* out,
                         void* workspace, size_t workspace_bytes, void* stream);
/* vo2w_varpred_forward_static: vo2w_varpred_forward without a caller workspace and without any allocation.  It has no
 * counterpart in the reference; it was added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The predictor's
 * intermediates live in a device array built into the library (VO2W_VARPRED_STATIC_BYTES, held by every GPU that loads
 * libvo2w.so until the process exits); the call makes no cudaMalloc / cudaFree and uses no allocating library routine.
 * Same arguments, results (bit-identical: the same kernels run) and errors as vo2w_varpred_forward, plus:
 *   - VO2W_E_WORKSPACE when vo2w_varpred_workspace_bytes(p, B, T) > VO2W_VARPRED_STATIC_BYTES, i.e. when
 *     B * (T + 2) * (4 * (d_in + 2 * d_filter) + 1) (each term rounded up to 256 bytes) exceeds 8 MiB -- about 2700
 *     tokens for the reference's d_in = d_filter = 256 (B = 16 x T = 168);
 *   - the first call on a device looks the array's address up (cudaGetSymbolAddress, with the caller's current device):
 *     make it outside graph capture;
 *   - every call on a device shares the one array, which keeps the previous call's contents (nothing in it has to
 *     start at zero: every intermediate is written in full before it is read).  Two calls that overlap on the GPU --
 *     different streams, or different handles -- race on it: serialise them, or use vo2w_varpred_forward with one
 *     workspace per stream. */
#define VO2W_VARPRED_STATIC_BYTES (8u << 20)
int vo2w_varpred_forward_static(vo2w_varpred* p, const float* x, const uint8_t* mask, int B, int T, float* out,
                                void* stream);
int vo2w_variance_embed(const float* x, const float* value, const float* bins, int n_edges, const float* table,
                        float mean, float stdv, float control, int denorm, float* value_out, float* out,
                        long long rows, int D, void* stream);
int vo2w_duration_round(const float* log_duration, float d_control, float* out, long long n, void* stream);

/* Test / profiling entry for the fp32-accurate tensor-pipe convolution of the acoustic tail (conv_tf32.cu):
 * out[b,l,:] = act(bias + res[b,l,:] + sum_j W[:,:,j] x[b,l+j-(k-1)/2,:]), time-major x [B,L,Cin], out/res [B,L,Cout],
 * w [Cout,Cin,k] (Conv1d layout; k=1 is nn.Linear), act 0 none / 1 relu / 2 tanh; bias, res may be NULL.
 * `scratch` (2*Cout*Cin*k floats) receives the packed weights.  Cin % 8 == 0; Cout % 64 == 0, or Cout <= 128 and
 * Cout % 16 == 0. */
int vo2w_debug_tf32_conv(const float* x, const float* w, const float* bias, const float* res, float* out, int B, int L,
                         int Cin, int Cout, int k, int act, float* scratch, void* stream);
/* The same with the fusions the Generator's fp32 mode uses (scripts/hifigan/models.py:96-103,156-163):
 * out = act(out_scale * (bias + res + acc + sum_j W[:,:,j] leaky_relu(x[l + (j - (k-1)/2) * dil], in_slope))),
 * dilation `dil` ((k-1)*dil <= 50), acc may alias out (the MRF sum), in_slope 1 = identity. */
int vo2w_debug_tf32_conv_ex(const float* x, const float* w, const float* bias, const float* res, const float* acc,
                            float* out, int B, int L, int Cin, int Cout, int k, int dil, float in_slope, float out_scale,
                            int act, float* scratch, void* stream);

/* ---- LengthRegulator -----------------------------------------------------------------
 * Replaces LengthRegulator.LR/expand (scripts/model/modules.py:132-155) + pad
 * (scripts/utils/tools.py:669-687).
 *
 * vo2w_lr_lengths: mel_len[b] = sum_i max(trunc(duration[b,i]), 0)  (int64, uncropped)
 * and *max_len_dev = max_b mel_len[b].  One launch; the caller reads max_len back only
 * when the reference would (max_len is None). */
int vo2w_lr_lengths(const void* duration, int dur_dtype, int B, int T_src,
                    int64_t* mel_len_dev, int64_t* max_len_dev, void* stream);

/* out[b,p,:] = x[b,i,:] for cum[i] <= p < cum[i+1], zeros for p >= mel_len[b], rows
 * beyond Lmax cropped.  x is [B,T_src,row_bytes] (any element type, copied bitwise),
 * out is [B,Lmax,row_bytes].  mask (optional, uint8 [B,Lmax]) = p >= mel_len[b]
 * (get_mask_from_lengths, scripts/utils/tools.py:164-171).  mel_len (optional) is
 * rewritten with the uncropped lengths. */
int vo2w_length_regulate(const void* x, const void* duration, int dur_dtype, int B, int T_src,
                         size_t row_bytes, void* out, long long Lmax, int64_t* mel_len_dev,
                         uint8_t* mask, void* stream);

/* mask[b,p] = p >= lengths[b], p < max_len (scripts/utils/tools.py:164-171). */
int vo2w_mask_from_lengths(const int64_t* lengths_dev, int B, long long max_len, uint8_t* mask,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VO2W_H */
