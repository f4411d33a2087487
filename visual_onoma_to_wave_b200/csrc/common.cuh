// Shared host/device helpers for libvo2w (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>

#include "../../include/vo2w.h"

// ---------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------
void vo2w_set_error(const char* fmt, ...);

#define VO2W_CUDA_OK(expr)                                                              \
    do {                                                                                \
        cudaError_t _e = (expr);                                                        \
        if (_e != cudaSuccess) {                                                        \
            vo2w_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr,                 \
                           cudaGetErrorString(_e));                                     \
            return VO2W_E_CUDA;                                                         \
        }                                                                               \
    } while (0)

#define VO2W_REQUIRE(cond, ...)                                                         \
    do {                                                                                \
        if (!(cond)) {                                                                  \
            vo2w_set_error(__VA_ARGS__);                                                \
            return VO2W_E_INVALID;                                                      \
        }                                                                               \
    } while (0)

static inline int vo2w_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// One-time per-DEVICE setup (cudaFuncSetAttribute, SM count): function attributes belong to the device's context, so a
// process that drives several GPUs must set them once per device, not once per process.
// Programmatic dependent launch of the tcgen05 kernels (pair_umma.cu, stream_umma.cu, conv_tf32.cu); VO2W_PDL=0 disables it.
static inline bool vo2w_pdl_enabled() {
    static const bool on = !(getenv("VO2W_PDL") && getenv("VO2W_PDL")[0] == '0');
    return on;
}
#ifdef __CUDACC__
// Programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// (barrier set-up, TMEM allocation, weight loads) while the previous kernel of the stream is draining; it must not
// touch that kernel's results before griddep_wait() returns (= the previous grid has completed and flushed).
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
#define VO2W_MAX_DEV 64
static inline int vo2w_cur_dev() {
    int d = 0;
    cudaGetDevice(&d);
    return (d >= 0 && d < VO2W_MAX_DEV) ? d : 0;
}
int vo2w_sm_count();   // errors.cu: multiprocessors of the current device (cached per device)

// ---------------------------------------------------------------------------------------
// activation layouts (DESIGN.md "Data layout in HBM")
//
//  operand buffers  bf16 [B][C/8][Lp][8]   Lp = PAD_ROWS + roundup(L,128) + PAD_ROWS
//      16-byte channel chunks, time-major inside a chunk: a tile of `rows` consecutive
//      samples of one chunk is one contiguous span (one 1-D bulk copy) and lands in
//      shared memory already in the UMMA K-major no-swizzle core-matrix order.  The pad
//      rows stay zero and implement the convolutions' "same" padding.
//  residual buffers fp32 [B][C/4][L][4]     16-byte chunks as well, no padding.
// ---------------------------------------------------------------------------------------
#define VO2W_PAD_ROWS 32
#define VO2W_TILE_M   128

static inline long long vo2w_padded_len(long long L) {
    return VO2W_PAD_ROWS + ((L + VO2W_TILE_M - 1) / VO2W_TILE_M) * VO2W_TILE_M + VO2W_PAD_ROWS;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------
// PTX wrappers: mbarrier, 1-D bulk async copy (TMA engine), tcgen05 / TMEM
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// suspendTimeHint of mbarrier.try_wait: the thread may stay suspended this long (it is woken when the phase completes).
// Without it an attempt comes back after ~100 - 200 cycles, and the retry loops of the waiting roles -- epilogue warps
// wait 50 - 90 % of a launch -- were a quarter of all instructions the fused kernel issued (ncu source page, C = 32, k = 3):
// instructions that buy nothing and, at the board's power cap, cost clock.
#define VO2W_MBAR_SUSPEND_NS 100000u
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(VO2W_MBAR_SUSPEND_NS)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug becomes a launch failure ("unspecified launch failure"),
// never a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > 4000000000LL) __trap();
    }
}

// global -> shared::cta, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
        : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// warp-collective; writes the TMEM base address to *dst_smem
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T, bf16 x bf16 -> fp32, issued by ONE thread (SASS: UTCHMMA)
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all previously issued tcgen05.mma of this thread arrive (once) on the mbarrier when done
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                 : "memory");
}

// TMEM -> registers: lane i of the warp reads TMEM lane (base + i), 8 consecutive columns (SASS: LDTM)
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory matrix descriptor, K-major, SWIZZLE_NONE ("interleave"):
//   in 16-byte units the operand is ((8,n),2):((1,SBO),LBO): 8 rows 16 B apart form a
//   core matrix, SBO = byte step between 8-row groups, LBO = byte step between the two
//   16-byte K chunks of one MMA.  No swizzle => any 16-byte aligned start address, which
//   is what lets a convolution tap be a plain row offset into the staged tile.
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version 1 (sm_100)
    return d;                  // base_offset 0, lbo_mode 0, layout_type 0 (SWIZZLE_NONE)
}
// instruction descriptor for kind::f16: D fp32, A/B bf16, both K-major, M x N
__device__ __forceinline__ uint32_t umma_idesc_bf16(uint32_t M, uint32_t N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// one lane of a converged warp (the form ptxas recognises as warp-uniform for tcgen05 issue)
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ float lrelu(float v, float slope) { return v > 0.f ? v : v * slope; }
#endif  // __CUDACC__
