// LengthRegulator for sm_100a: prefix sum over durations + bandwidth-bound row gather.
//
// Replaces LengthRegulator.LR / expand (scripts/model/modules.py:132-155) and pad
// (scripts/utils/tools.py:669-687) of the reference, which walk the batch in Python with
// one .item() device->host sync per token.  Here: one launch computes the (uncropped)
// lengths, one launch expands.  HBM-bound: bytes = read x + read durations + write out
// ("LengthRegulator kernel").
#include "common.cuh"

namespace {

constexpr int LR_THREADS = 256;
constexpr int LR_ROWS = 64;   // output rows per CTA

// n = max(int(d), 0): python int() truncates toward zero; negatives and -0.0 give 0
// (modules.py:152-153).  Saturates instead of overflowing on absurd values.
__device__ __forceinline__ long long dur_count(const void* dur, int dtype, long long idx) {
    switch (dtype) {
        case VO2W_DUR_I32: { int v = ((const int*)dur)[idx]; return v > 0 ? v : 0; }
        case VO2W_DUR_I64: { long long v = ((const long long*)dur)[idx]; return v > 0 ? v : 0; }
        case VO2W_DUR_F32: {
            float v = ((const float*)dur)[idx];
            if (!(v >= 1.0f)) return 0;                       // also NaN
            return v >= 9.0e18f ? (long long)9.0e18 : (long long)v;   // cast truncates
        }
        default: {
            double v = ((const double*)dur)[idx];
            if (!(v >= 1.0)) return 0;
            return v >= 9.0e18 ? (long long)9.0e18 : (long long)v;
        }
    }
}

__device__ __forceinline__ long long warp_inclusive_scan(long long v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}

// Block-wide inclusive scan of the row's counts into cum[1..T] (cum[0] = 0); returns total.
// Warp-level shuffles inside a warp, one shared-memory hop across the 8 warps.
__device__ long long block_scan_row(const void* dur, int dtype, long long row_base, int T, long long* cum,
                                    long long* warp_tot) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long carry = 0;
    if (tid == 0) cum[0] = 0;
    for (int base = 0; base < T; base += LR_THREADS) {
        const int i = base + tid;
        long long v = i < T ? dur_count(dur, dtype, row_base + i) : 0;
        long long s = warp_inclusive_scan(v, lane);
        if (lane == 31) warp_tot[warp] = s;
        __syncthreads();
        long long off = carry;
        for (int w = 0; w < warp; ++w) off += warp_tot[w];
        if (i < T) cum[i + 1] = s + off;
        long long tot = 0;
        for (int w = 0; w < LR_THREADS / 32; ++w) tot += warp_tot[w];
        carry += tot;
        __syncthreads();
    }
    return carry;
}

__global__ void __launch_bounds__(LR_THREADS) lr_lengths_kernel(const void* dur, int dtype, int T, long long* mel_len,
                                                                unsigned long long* max_len) {
    __shared__ long long warp_tot[LR_THREADS / 32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    long long acc = 0;
    for (int i = tid; i < T; i += LR_THREADS) acc += dur_count(dur, dtype, (long long)b * T + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) warp_tot[warp] = acc;
    __syncthreads();
    if (tid == 0) {
        long long tot = 0;
        for (int w = 0; w < LR_THREADS / 32; ++w) tot += warp_tot[w];
        mel_len[b] = tot;
        if (max_len) atomicMax(max_len, (unsigned long long)tot);
    }
}

// VEC = bytes moved per lane per access (16 when row_bytes and both pointers allow it).
template <typename VecT>
__global__ void __launch_bounds__(LR_THREADS)
lr_expand_kernel(const unsigned char* __restrict__ x, const void* __restrict__ dur, int dtype, int T, size_t row_bytes,
                 unsigned char* __restrict__ out, long long Lmax, long long* __restrict__ mel_len,
                 unsigned char* __restrict__ mask) {
    extern __shared__ long long lr_smem[];          // cum[T+1]
    __shared__ long long warp_tot[LR_THREADS / 32];
    __shared__ int src_tok[LR_ROWS];                // token index per output row, -1 = zero row
    long long* cum = lr_smem;
    const int b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long p0 = (long long)blockIdx.x * LR_ROWS;

    const long long total = block_scan_row(dur, dtype, (long long)b * T, T, cum, warp_tot);
    if (blockIdx.x == 0 && tid == 0 && mel_len) mel_len[b] = total;

    // one binary search per output row: largest i with cum[i] <= p  (cum non-decreasing)
    if (tid < LR_ROWS) {
        const long long p = p0 + tid;
        int tok = -1;
        if (p < total && p < Lmax) {
            int lo = 0, hi = T;                      // invariant: cum[lo] <= p < cum[hi]
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (cum[mid] <= p) lo = mid; else hi = mid;
            }
            tok = lo;
        }
        src_tok[tid] = tok;
        if (mask && p < Lmax) mask[(long long)b * Lmax + p] = (unsigned char)(p >= total);
    }
    __syncthreads();

    const int vec_per_row = (int)(row_bytes / sizeof(VecT));
    const VecT* xb = (const VecT*)(x + (size_t)b * T * row_bytes);
    VecT* ob = (VecT*)(out + ((size_t)b * Lmax + p0) * row_bytes);
    VecT zero;
    memset(&zero, 0, sizeof(VecT));
    // each warp owns rows warp, warp+8, ...; two rows in flight per iteration for MLP
    for (int r = warp; r < LR_ROWS; r += 2 * (LR_THREADS / 32)) {
        const int r2 = r + LR_THREADS / 32;
        const bool v1 = p0 + r < Lmax, v2 = r2 < LR_ROWS && p0 + r2 < Lmax;
        const int t1 = v1 ? src_tok[r] : -1, t2 = v2 ? src_tok[r2] : -1;
        for (int c = lane; c < vec_per_row; c += 32) {
            VecT a = zero, d = zero;
            if (t1 >= 0) a = __ldg(xb + (size_t)t1 * vec_per_row + c);
            if (t2 >= 0) d = __ldg(xb + (size_t)t2 * vec_per_row + c);
            if (v1) __stcs(ob + (size_t)r * vec_per_row + c, a);     // streaming store: written once
            if (v2) __stcs(ob + (size_t)r2 * vec_per_row + c, d);
        }
    }
}

__global__ void mask_kernel(const long long* lengths, long long max_len, unsigned char* mask, int B) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * max_len) return;
    const int b = (int)(i / max_len);
    mask[i] = (unsigned char)((i - (long long)b * max_len) >= lengths[b]);
}

}  // namespace

extern "C" int vo2w_lr_lengths(const void* duration, int dur_dtype, int B, int T_src, int64_t* mel_len_dev,
                               int64_t* max_len_dev, void* stream) {
    VO2W_REQUIRE(B >= 0 && T_src >= 0, "lr_lengths: negative shape");
    VO2W_REQUIRE(dur_dtype >= 0 && dur_dtype <= 3, "lr_lengths: bad duration dtype %d", dur_dtype);
    VO2W_REQUIRE(mel_len_dev != nullptr, "lr_lengths: mel_len is null");
    cudaStream_t st = (cudaStream_t)stream;
    if (max_len_dev) VO2W_CUDA_OK(cudaMemsetAsync(max_len_dev, 0, sizeof(int64_t), st));
    if (B == 0) return VO2W_OK;
    lr_lengths_kernel<<<B, LR_THREADS, 0, st>>>(duration, dur_dtype, T_src, (long long*)mel_len_dev,
                                                (unsigned long long*)max_len_dev);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

extern "C" int vo2w_length_regulate(const void* x, const void* duration, int dur_dtype, int B, int T_src,
                                    size_t row_bytes, void* out, long long Lmax, int64_t* mel_len_dev, uint8_t* mask,
                                    void* stream) {
    VO2W_REQUIRE(B >= 0 && T_src >= 0 && Lmax >= 0, "length_regulate: negative shape");
    VO2W_REQUIRE(dur_dtype >= 0 && dur_dtype <= 3, "length_regulate: bad duration dtype %d", dur_dtype);
    VO2W_REQUIRE(row_bytes > 0, "length_regulate: row_bytes is 0");
    VO2W_REQUIRE(B < 65536, "length_regulate: batch %d exceeds grid.y", B);
    cudaStream_t st = (cudaStream_t)stream;
    if (B == 0) return VO2W_OK;
    if (Lmax == 0) {   // all-zero durations: out is [B,0,H]; lengths still have to be reported
        if (mel_len_dev) return vo2w_lr_lengths(duration, dur_dtype, B, T_src, mel_len_dev, nullptr, stream);
        return VO2W_OK;
    }
    VO2W_REQUIRE(x != nullptr && out != nullptr && duration != nullptr, "length_regulate: null pointer");
    const size_t smem = (size_t)(T_src + 1) * sizeof(long long);
    VO2W_REQUIRE(smem <= 200 * 1024, "length_regulate: T_src %d too long for the shared-memory scan", T_src);
    dim3 grid((unsigned)vo2w_ceil_div(Lmax, LR_ROWS), (unsigned)B);
    const uintptr_t align = (uintptr_t)x | (uintptr_t)out | (uintptr_t)row_bytes;
#define VO2W_LR_LAUNCH(VT)                                                                                      \
    do {                                                                                                        \
        if (smem > 48 * 1024)                                                                                   \
            VO2W_CUDA_OK(cudaFuncSetAttribute(lr_expand_kernel<VT>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                              (int)smem));                                                      \
        lr_expand_kernel<VT><<<grid, LR_THREADS, smem, st>>>((const unsigned char*)x, duration, dur_dtype, T_src, \
                                                             row_bytes, (unsigned char*)out, Lmax,              \
                                                             (long long*)mel_len_dev, mask);                    \
    } while (0)
    if (align % 16 == 0) VO2W_LR_LAUNCH(uint4);
    else if (align % 8 == 0) VO2W_LR_LAUNCH(uint2);
    else if (align % 4 == 0) VO2W_LR_LAUNCH(unsigned int);
    else if (align % 2 == 0) VO2W_LR_LAUNCH(unsigned short);
    else VO2W_LR_LAUNCH(unsigned char);
#undef VO2W_LR_LAUNCH
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

extern "C" int vo2w_mask_from_lengths(const int64_t* lengths_dev, int B, long long max_len, uint8_t* mask,
                                      void* stream) {
    VO2W_REQUIRE(B >= 0 && max_len >= 0, "mask_from_lengths: negative shape");
    if (B == 0 || max_len == 0) return VO2W_OK;
    VO2W_REQUIRE(lengths_dev && mask, "mask_from_lengths: null pointer");
    const long long n = (long long)B * max_len;
    mask_kernel<<<(unsigned)vo2w_ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>((const long long*)lengths_dev, max_len,
                                                                                   mask, B);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}
