// Host-side weight packing and launch assembly for umma_conv_kernel.  See pack.cuh.
#include "pack.cuh"

#include <algorithm>

uint16_t f32_to_bf16_rn(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);   // NaN
    u += 0x7fffu + ((u >> 16) & 1u);                                            // round to nearest even
    return (uint16_t)(u >> 16);
}

static int round_up(int a, int b) { return (a + b - 1) / b * b; }

// Write W_tap[n][ci] for one item into the packed block [Cin/8][n_size][8].
// get(n, ci) returns the fp32 weight or 0 outside the real tensor.
template <typename F>
static void pack_item(uint16_t* dst, int Cin_p, int n_size, F get) {
    for (int c = 0; c < Cin_p / 8; ++c)
        for (int n = 0; n < n_size; ++n)
            for (int e = 0; e < 8; ++e) dst[((size_t)c * n_size + n) * 8 + e] = f32_to_bf16_rn(get(n, c * 8 + e));
}

int pack_conv(PackedLayer& L, const float* w, const float* bias, int Cout, int Cin, int k, int dil) {
    VO2W_REQUIRE(k >= 1 && k <= UMMA_MAX_ITEMS && (k & 1), "pack_conv: kernel size %d unsupported", k);
    VO2W_REQUIRE(dil * (k - 1) / 2 <= VO2W_PAD_ROWS, "pack_conv: halo %d exceeds the %d pad rows", dil * (k - 1) / 2,
                 VO2W_PAD_ROWS);
    L = PackedLayer();
    L.Cin = round_up(Cin, 16);
    L.Cout = Cout;
    L.k = k;
    L.dil = dil;
    const int Cout_p = round_up(Cout, 16);
    L.n_cols = std::min(Cout_p, 256);
    VO2W_REQUIRE(Cout_p % L.n_cols == 0, "pack_conv: Cout %d does not tile by %d", Cout, L.n_cols);
    L.n_tiles = Cout_p / L.n_cols;
    L.halo_l = L.halo_r = dil * (k - 1) / 2;
    L.n_items = k;
    const size_t item_elems = (size_t)L.n_cols * L.Cin;
    L.w_tile_stride = (uint32_t)(item_elems * k * 2);
    L.w_host.assign(item_elems * k * L.n_tiles, 0);
    L.bias_host.assign((size_t)L.n_tiles * L.n_cols, 0.f);
    for (int j = 0; j < k; ++j) {
        UItem& it = L.items[j];
        it.row_shift = j * dil;
        it.n_off = 0;
        it.n_size = L.n_cols;
        it.zero_init = j == 0;
        it.w_off = (uint32_t)(item_elems * j * 2);
    }
    for (int t = 0; t < L.n_tiles; ++t) {
        for (int j = 0; j < k; ++j)
            pack_item(L.w_host.data() + ((size_t)t * k + j) * item_elems, L.Cin, L.n_cols, [&](int n, int ci) {
                const int co = t * L.n_cols + n;
                return (co < Cout && ci < Cin) ? w[((size_t)co * Cin + ci) * k + j] : 0.f;
            });
        for (int n = 0; n < L.n_cols; ++n) {
            const int co = t * L.n_cols + n;
            L.bias_host[(size_t)t * L.n_cols + n] = co < Cout ? bias[co] : 0.f;
        }
    }
    return VO2W_OK;
}

int pack_convT(PackedLayer& L, const float* w, const float* bias, int Cin, int Cout, int k, int s) {
    VO2W_REQUIRE(k == 2 * s && s % 2 == 0, "pack_convT: needs kernel == 2*stride and even stride (k=%d s=%d)", k, s);
    VO2W_REQUIRE(Cin % 16 == 0 && Cout % 16 == 0, "pack_convT: channels must be multiples of 16");
    L = PackedLayer();
    L.Cin = Cin;
    L.Cout = Cout;
    L.k = k;
    L.up_s = s;
    const int pad = s / 2;
    L.n_cols = std::min(s * Cout, 256);
    VO2W_REQUIRE(L.n_cols % Cout == 0, "pack_convT: Cout %d does not divide the %d-column tile", Cout, L.n_cols);
    L.phases_per_tile = L.n_cols / Cout;
    VO2W_REQUIRE(s % L.phases_per_tile == 0, "pack_convT: %d phases per tile do not divide stride %d", L.phases_per_tile, s);
    L.n_tiles = s / L.phases_per_tile;
    L.halo_l = L.halo_r = 1;
    // phase r of out[s*q + r]:  x[q+e] . w[:,:,j0]  +  x[q+e-1] . w[:,:,j0+s],  j0 = (r+pad) % s, e = (r+pad) / s.
    // The staged tile starts at input sample l0-1, so x[q+e-1] is row shift e and x[q+e] is e+1.
    // When every phase of a tile shares e the kernel adds it per tile (extra shift); otherwise
    // (single tile holding all phases) it is folded into the items here.
    const bool shared_e = L.n_tiles > 1;
    if (shared_e)
        VO2W_REQUIRE((s / 2) % L.phases_per_tile == 0, "pack_convT: tile straddles the phase halves");
    const int groups = shared_e ? 1 : L.phases_per_tile;      // column groups with their own shifts
    const int g_cols = L.n_cols / groups;
    L.n_items = 2 * groups;
    VO2W_REQUIRE(L.n_items <= UMMA_MAX_ITEMS, "pack_convT: too many items");
    const size_t item_elems = (size_t)g_cols * Cin;
    L.w_tile_stride = (uint32_t)(item_elems * L.n_items * 2);
    L.w_host.assign(item_elems * L.n_items * L.n_tiles, 0);
    L.bias_host.assign((size_t)L.n_tiles * L.n_cols, 0.f);
    for (int g = 0; g < groups; ++g)
        for (int half = 0; half < 2; ++half) {          // half 0: x[q+e-1] with tap j0+s; half 1: x[q+e] with tap j0
            UItem& it = L.items[2 * g + half];
            const int e = shared_e ? 0 : (g + pad) / s;
            it.row_shift = e + half;
            it.n_off = g * g_cols;
            it.n_size = g_cols;
            it.zero_init = half == 0;
            it.w_off = (uint32_t)(item_elems * (2 * g + half) * 2);
        }
    for (int t = 0; t < L.n_tiles; ++t) {
        for (int g = 0; g < groups; ++g)
            for (int half = 0; half < 2; ++half)
                pack_item(L.w_host.data() + ((size_t)t * L.n_items + 2 * g + half) * item_elems, Cin, g_cols,
                          [&](int n, int ci) {
                              const int col = g * g_cols + n;
                              const int r = t * L.phases_per_tile + col / Cout, co = col % Cout;
                              const int j = (r + pad) % s + (half == 0 ? s : 0);
                              return w[((size_t)ci * Cout + co) * k + j];
                          });
        for (int n = 0; n < L.n_cols; ++n) L.bias_host[(size_t)t * L.n_cols + n] = bias[n % Cout];
    }
    return VO2W_OK;
}

int upload_layer(PackedLayer& L) {
    VO2W_CUDA_OK(cudaMalloc(&L.w_dev, L.w_host.size() * 2));
    VO2W_CUDA_OK(cudaMemcpy(L.w_dev, L.w_host.data(), L.w_host.size() * 2, cudaMemcpyHostToDevice));
    VO2W_CUDA_OK(cudaMalloc(&L.bias_dev, L.bias_host.size() * 4));
    VO2W_CUDA_OK(cudaMemcpy(L.bias_dev, L.bias_host.data(), L.bias_host.size() * 4, cudaMemcpyHostToDevice));
    std::vector<uint16_t>().swap(L.w_host);
    return VO2W_OK;
}

void free_layer(PackedLayer& L) {
    if (L.w_dev) cudaFree(L.w_dev);
    if (L.bias_dev) cudaFree(L.bias_dev);
    L.w_dev = nullptr;
    L.bias_dev = nullptr;
}

int launch_layers(const PackedLayer* const* layers, const __nv_bfloat16* const* ins, int n_seg, int B, int L_in,
                  const EpilogueSpec& ep, cudaStream_t st) {
    VO2W_REQUIRE(n_seg >= 1 && n_seg <= UMMA_MAX_SEGS, "launch_layers: %d segments", n_seg);
    const PackedLayer& L0 = *layers[0];
    UConv p;
    memset(&p, 0, sizeof(p));
    p.n_seg = n_seg;
    p.Cin = L0.Cin;
    p.n_cols = L0.n_cols;
    int n_max = 0, halo_max = 0;
    for (int s = 0; s < n_seg; ++s) {
        const PackedLayer& L = *layers[s];
        VO2W_REQUIRE(L.Cin == L0.Cin && L.n_cols == L0.n_cols && L.n_tiles == L0.n_tiles && L.up_s == L0.up_s &&
                         L.Cout == L0.Cout,
                     "launch_layers: fused segments must share their shape");
        for (int i = 0; i < L.n_items; ++i) n_max = std::max(n_max, L.items[i].n_size);
        halo_max = std::max(halo_max, L.halo_l + L.halo_r);
    }
    // Row blocks per CTA: every weight stage then serves n_blk x 128 rows.  Measured on B200: it pays only where
    // the per-CTA work is tiny (conv_post: 16 accumulator columns, 14 MMAs per 128 rows -> 142 us became 103 us);
    // for the upsamplers the many small CTAs per SM (TMEM and shared memory allow up to 8) hide more latency than
    // the larger tiles save, so they keep n_blk = 1.
    const int steps = p.Cin / 16;
    auto stage_for = [&](int n_blk, size_t limit) {
        const size_t a_bytes = (size_t)p.Cin * (VO2W_TILE_M * n_blk + halo_max) * 2;
        for (int c = std::min(steps, 8); c >= 1; --c) {
            if (steps % c) continue;
            const size_t stage = (size_t)n_max * 32 * c;
            if (stage <= 48 * 1024 && a_bytes + UMMA_STAGES * stage <= limit) return c;
        }
        return 0;
    };
    int n_blk = 1, ks = 0;
    for (int nb = 8; nb >= 1; nb >>= 1) {
        if (nb * p.n_cols > 512 || (nb > 1 && p.n_cols > 16)) continue;
        const long long ctas = (long long)vo2w_ceil_div(L_in, VO2W_TILE_M * nb) * L0.n_tiles * B;
        if (nb > 1 && ctas < 2 * 148) continue;
        const int c = stage_for(nb, 110 * 1024);
        if (c) { n_blk = nb; ks = c; break; }
    }
    if (!ks) ks = stage_for(1, UMMA_SMEM_LIMIT);
    p.n_blk = n_blk;
    int rows_max = 0;
    for (int s = 0; s < n_seg; ++s) {
        const PackedLayer& L = *layers[s];
        USeg& sg = p.seg[s];
        sg.in = ins[s];
        sg.w = L.w_dev;
        sg.w_tile_stride = L.w_tile_stride;
        sg.n_items = L.n_items;
        sg.halo_l = L.halo_l;
        sg.rows = VO2W_TILE_M * n_blk + L.halo_l + L.halo_r;
        for (int i = 0; i < L.n_items; ++i) {
            sg.items[i] = L.items[i];
            if (s > 0) sg.items[i].zero_init = 0;     // later segments accumulate on top
        }
        rows_max = std::max(rows_max, sg.rows);
    }
    p.a_bytes = (uint32_t)(p.Cin * rows_max * 2);
    VO2W_REQUIRE(ks > 0, "launch_layers: tile does not fit shared memory (Cin %d rows %d)", p.Cin, rows_max);
    p.kb_steps = ks;
    p.n_kblocks = steps / ks;
    p.stage_bytes = (uint32_t)(n_max * 32 * ks);
    p.Lp_in = (int)vo2w_padded_len(L_in);
    p.M_valid = L_in;
    int tc = 32;
    while (tc < p.n_blk * p.n_cols) tc <<= 1;
    p.tmem_cols = tc;
    p.bias = ep.bias ? ep.bias : L0.bias_dev;
    p.n_res = ep.n_res;
    for (int i = 0; i < 3; ++i) p.res[i] = ep.res[i];
    p.div = ep.div;
    p.out_f32 = ep.out_f32;
    p.out_bf16 = ep.out_bf16;
    p.out_slope = ep.out_slope;
    p.out_wave = ep.out_wave;
    p.Cout = L0.Cout;
    p.up_s = L0.up_s;
    p.up_pad = L0.up_s / 2;
    p.phases_per_tile = L0.phases_per_tile;
    p.L_out = L_in * L0.up_s;
    p.Lp_out = (int)vo2w_padded_len(p.L_out);
    return umma_conv_launch(p, vo2w_ceil_div(L_in, VO2W_TILE_M * p.n_blk), L0.n_tiles, B, st);
}

// Host-only probe of the packing logic (no device needed): lets the CPU test-suite replay the
// kernel's indexing in numpy against F.conv1d / F.conv_transpose1d.
extern "C" int vo2w_debug_pack(int transposed, const float* w, const float* bias, int Cin, int Cout, int k,
                               int dil_or_stride, int32_t* meta, uint16_t* w_out, size_t w_capacity, float* bias_out,
                               size_t bias_capacity) {
    VO2W_REQUIRE(w && bias && meta, "debug_pack: null argument");
    PackedLayer L;
    int rc = transposed ? pack_convT(L, w, bias, Cin, Cout, k, dil_or_stride) : pack_conv(L, w, bias, Cout, Cin, k, dil_or_stride);
    if (rc != VO2W_OK) return rc;
    int32_t head[12] = {L.Cin, L.Cout, L.n_cols, L.n_tiles, L.phases_per_tile, L.halo_l, L.halo_r, L.n_items,
                        (int32_t)L.w_tile_stride, L.up_s, (int32_t)L.w_host.size(), (int32_t)L.bias_host.size()};
    memcpy(meta, head, sizeof(head));
    for (int i = 0; i < L.n_items; ++i) {
        int32_t* m = meta + 12 + 5 * i;
        m[0] = L.items[i].row_shift; m[1] = L.items[i].n_off; m[2] = L.items[i].n_size;
        m[3] = L.items[i].zero_init; m[4] = (int32_t)L.items[i].w_off;
    }
    if (w_out) {
        VO2W_REQUIRE(w_capacity >= L.w_host.size() && bias_capacity >= L.bias_host.size(), "debug_pack: buffers too small");
        memcpy(w_out, L.w_host.data(), L.w_host.size() * 2);
        memcpy(bias_out, L.bias_host.data(), L.bias_host.size() * 4);
    }
    return VO2W_OK;
}
