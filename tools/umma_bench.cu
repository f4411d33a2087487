// tcgen05.mma issue-rate microbenchmark for sm_100a (design input for the conv kernels).
//
// One CTA (or CTA pair) per SM, one thread issues a long stream of MMAs on operands that are already
// in shared memory / TMEM; we measure MACs per cycle per SM for the operand shapes the convolution
// stages need:  M = 64/128(/256 with cta_group::2),  N = 32..256,  A from shared memory (SS) or from
// TMEM (TS), SWIZZLE_NONE or SWIZZLE_128B K-major layouts, kind::f16 (bf16) or kind::tf32.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/umma_bench tools/umma_bench.cu
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../visual_onoma_to_wave_b200/csrc/common.cuh"

struct BenchP {
    int batches;    // MMA batches of 8 instructions
};

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

template <int CG, bool TF32>
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
    if (CG == 1) {
        if (TF32) asm volatile("tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, 1;" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
        else asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, 1;" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
    } else {
        if (TF32) asm volatile("tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, 1;" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
        else asm volatile("tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, 1;" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
    }
}
template <int CG>
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc) {
    if (CG == 1) asm volatile("tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, 1;" ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
    else asm volatile("tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, 1;" ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc) : "memory");
}
template <int CG>
__device__ __forceinline__ void commit(uint32_t bar) {
    if (CG == 1)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    else
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// CG cta_group, M/N instruction shape, TS: A from TMEM, TF32, NACC independent accumulators (round-robin),
// SHIFT: row shift of every second A descriptor (conv taps on a SWIZZLE_NONE K-major tile)
// AROWS > 0: the fused kernel's operand pattern -- A tile of AROWS rows per 16-byte channel chunk (LBO = AROWS * 16 B),
// NACC row blocks of 128 rows, 8 / NACC K=16 steps each, a tap shift of (batch % 3) rows
template <int CG, int M, int N, bool TS, bool TF32, int NACC, int SHIFT, int AROWS = 0>
__global__ void __launch_bounds__(128) bench_kernel(const BenchP p, unsigned long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[4];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    uint32_t cta_rank = 0;
    if (CG == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));

    for (int i = threadIdx.x; i < (96 * 1024) / 4; i += blockDim.x)
        ((uint32_t*)smem)[i] = 0x3c003c00u ^ ((uint32_t)(i * 2654435761u) & 0x007f007fu);
    fence_proxy_async();
    const uint32_t bar0 = smem_u32(bars);
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) mbar_init(bar0 + 8u * i, 1);
        fence_mbar_init();
    }
    if (warp == 0) {
        if (CG == 1) {
            tmem_alloc(smem_u32(&tmem_base_s), 512);
        } else {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    constexpr int m_cta = M / CG, n_cta = N / CG;
    constexpr uint32_t a_lbo = AROWS ? (uint32_t)AROWS * 16u : (uint32_t)(m_cta + 32) * 16u, b_lbo = (uint32_t)n_cta * 16u;
    const uint32_t a_smem = smem_u32(smem);                 // 32 KB (64 KB with AROWS)
    const uint32_t b_smem = a_smem + (AROWS ? 64 : 32) * 1024;
    if (warp == 1 && cta_rank == 0) {
        constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24) |
                                   (TF32 ? ((2u << 7) | (2u << 10)) : ((1u << 7) | (1u << 10)));
        const uint64_t a0 = umma_smem_desc(a_smem, a_lbo, 128u), b0 = umma_smem_desc(b_smem, b_lbo, 128u);
        const long long t0 = clock64();
        for (int b = 0; b < p.batches; ++b) {
            if (b >= 4) mbar_wait(bar0 + 8u * (b & 3), (uint32_t)(((b >> 2) - 1) & 1));
            if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    const uint32_t d = tmem_base + (uint32_t)((ks % NACC) * N);
                    const uint64_t bd = b0 + (uint64_t)(((uint32_t)ks * 2u * b_lbo) >> 4);
                    if (AROWS) {
                        constexpr int KSTEPS = 8 / NACC;
                        const int blk = ks / KSTEPS, st = ks % KSTEPS;
                        const uint64_t ad = a0 + (uint64_t)(((uint32_t)blk * 128u * 16u + (uint32_t)st * 2u * a_lbo + (uint32_t)(b % 3) * 16u) >> 4);
                        const uint64_t bd2 = b0 + (uint64_t)(((uint32_t)st * 2u * b_lbo) >> 4);
                        mma_ss<CG, TF32>(tmem_base + (uint32_t)(blk * N), ad, bd2, idesc);
                    } else if (TS) {
                        mma_ts<CG>(d, tmem_base + 448u + (uint32_t)ks * 8u, bd, idesc);
                    } else {
                        const uint64_t ad = a0 + (uint64_t)((((uint32_t)(ks & 3) * 2u * a_lbo) + ((ks & 1) ? SHIFT * 16u : 0u)) >> 4);
                        mma_ss<CG, TF32>(d, ad, bd, idesc);
                    }
                }
                commit<CG>(bar0 + 8u * (b & 3));
            }
            __syncwarp();
        }
        for (int b = p.batches > 4 ? p.batches - 4 : 0; b < p.batches; ++b)
            mbar_wait(bar0 + 8u * (b & 3), (uint32_t)((b >> 2) & 1));
        const long long t1 = clock64();
        if ((threadIdx.x & 31) == 0) cycles[blockIdx.x / CG] = (unsigned long long)(t1 - t0);
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) {
        if (CG == 1) tmem_dealloc(tmem_base, 512);
        else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// Queue-depth probe: one thread issues `n` MMAs (cg=2, N=32, 4 accumulators) back to back and reads the clock right
// after the last issue (before any completion wait), then waits for completion.  issue_cycles ~ 0 for small n
// if the tensor pipe queues instructions; the knee is the queue depth.
__global__ void __launch_bounds__(128) queue_kernel(int n, unsigned long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[2];
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    uint32_t cta_rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
    for (int i = threadIdx.x; i < (96 * 1024) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u;
    fence_proxy_async();
    const uint32_t bar0 = smem_u32(bars);
    if (threadIdx.x == 0) { mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); fence_mbar_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    if (warp == 1 && cta_rank == 0) {
        constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24) | (1u << 7) | (1u << 10);
        const uint64_t a0 = umma_smem_desc(smem_u32(smem), 514u * 16u, 128u), b0 = umma_smem_desc(smem_u32(smem) + 64 * 1024, 16u * 16u, 128u);
        if (elect_one()) {
            const long long t0 = clock64();
            for (int i = 0; i < n; ++i) mma_ss<2, false>(tmem_base + (uint32_t)((i & 3) * 32), a0 + (uint64_t)((i & 3) * 128), b0, idesc);
            const long long t1 = clock64();
            commit<2>(bar0);
            mbar_wait(bar0, 0);
            const long long t2 = clock64();
            // cost of polling barriers that are already complete, and of a cluster-mapped remote arrive
            for (int i = 0; i < 8; ++i) mbar_wait(bar0, 0);
            const long long t3 = clock64();
            for (int i = 0; i < 8; ++i) commit<2>(bar0 + 8);
            const long long t4 = clock64();
            if (blockIdx.x == 0) { out[0] = (unsigned long long)(t1 - t0); out[1] = (unsigned long long)(t2 - t0);
                                   out[2] = (unsigned long long)(t3 - t2); out[3] = (unsigned long long)(t4 - t3); }
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

template <int CG, int M, int N, bool TS, bool TF32, int NACC, int SHIFT, int AROWS = 0>
static void run(const char* name, int batches, int n_sm) {
    BenchP p{batches};
    auto kern = bench_kernel<CG, M, N, TS, TF32, NACC, SHIFT, AROWS>;
    unsigned long long* d_cyc;
    const int n_units = n_sm / CG;
    CK(cudaMalloc(&d_cyc, sizeof(unsigned long long) * n_units));
    const size_t smem = 96 * 1024 + 1024;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    float best_ms = 1e30f;
    std::vector<unsigned long long> h(n_units);
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(e0));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(n_units * CG);
        cfg.blockDim = dim3(128);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        CK(cudaLaunchKernelEx(&cfg, kern, p, d_cyc));
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        best_ms = std::min(best_ms, ms);
    }
    CK(cudaMemcpy(h.data(), d_cyc, sizeof(unsigned long long) * n_units, cudaMemcpyDeviceToHost));
    std::sort(h.begin(), h.end());
    const double K = TF32 ? 8.0 : 16.0;
    const double macs_unit = (double)M * N * K * 8.0 * batches;
    const double cyc_med = (double)h[n_units / 2];
    const double per_mma = cyc_med / (8.0 * batches);
    const double mac_cyc_sm = macs_unit / cyc_med / CG;
    const double tflops = 2.0 * macs_unit * n_units / (best_ms * 1e-3) / 1e12;
    printf("%-14s cg=%d M=%3d N=%3d %s %s nacc=%d shift=%d | cyc/MMA %6.1f (ideal %5.1f) | MAC/cyc/SM %7.1f (%5.1f%% of %d) | %7.1f TFLOP/s wall\n",
           name, CG, M, N, TS ? "TS" : "SS", TF32 ? "tf32" : "bf16", NACC, SHIFT, per_mma, (double)(M / CG < 128 ? 128 : M / CG) * N / 256.0,
           mac_cyc_sm, 100.0 * mac_cyc_sm / (TF32 ? 2048.0 : 4096.0), TF32 ? 2048 : 4096, tflops);
    fflush(stdout);
    cudaFree(d_cyc);
}

int main(int argc, char** argv) {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int n_sm = prop.multiProcessorCount;
    printf("device %s, %d SMs, cc %d.%d\n", prop.name, n_sm, prop.major, prop.minor);
    const int bt = argc > 1 ? atoi(argv[1]) : 2048;
    if (argc > 2 && argv[2][0] == 'q') {   // tensor-pipe queue depth
        unsigned long long* d;
        CK(cudaMalloc(&d, 32));
        CK(cudaFuncSetAttribute(queue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 97 * 1024));
        for (int n : {1, 2, 4, 8, 12, 16, 24, 32, 48, 64, 128}) {
            unsigned long long h[4];
            for (int rep = 0; rep < 2; ++rep) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 97 * 1024;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                cfg.attrs = at; cfg.numAttrs = 1;
                CK(cudaLaunchKernelEx(&cfg, queue_kernel, n, d));
                CK(cudaDeviceSynchronize());
            }
            CK(cudaMemcpy(h, d, 32, cudaMemcpyDeviceToHost));
            printf("queue probe: %3d MMAs (cg=2 N=32): issue loop %6llu cycles, until complete %6llu cycles; 8 polls of a complete barrier %llu, 8 commits %llu\n", n, h[0], h[1], h[2], h[3]);
        }
        return 0;
    }
    if (argc > 2) {   // the fused kernel's operand patterns only
        run<2, 256, 32, false, false, 1, 0>("2CTA SS", bt, n_sm);
        run<2, 256, 32, false, false, 4, 0>("2CTA SS 4acc", bt, n_sm);
        run<2, 256, 32, false, false, 4, 0, 514>("pair C32 conv1", bt, n_sm);
        run<2, 256, 32, false, false, 4, 0, 512>("pair C32 conv2", bt, n_sm);
        run<2, 256, 32, false, false, 4, 0, 160>("C32 lbo160", bt, n_sm);
        run<2, 256, 64, false, false, 2, 0, 258>("pair C64 conv1", bt, n_sm);
        run<2, 256, 64, false, false, 2, 0, 256>("pair C64 conv2", bt, n_sm);
        run<2, 256, 128, false, false, 2, 0, 258>("pair C128 conv1", bt, n_sm);
        return 0;
    }
    run<1, 128, 32, false, false, 1, 0>("SS", bt, n_sm);
    run<1, 128, 64, false, false, 1, 0>("SS", bt, n_sm);
    run<1, 128, 128, false, false, 1, 0>("SS", bt, n_sm);
    run<1, 128, 256, false, false, 1, 0>("SS", bt, n_sm);
    run<1, 128, 32, false, false, 4, 3>("SS taps", bt, n_sm);
    run<1, 128, 64, false, false, 4, 3>("SS taps", bt, n_sm);
    run<1, 128, 128, false, false, 2, 3>("SS taps", bt, n_sm);
    run<1, 128, 256, false, false, 1, 3>("SS taps", bt, n_sm);
    run<1, 128, 32, true, false, 1, 0>("TS", bt, n_sm);
    run<1, 128, 64, true, false, 1, 0>("TS", bt, n_sm);
    run<1, 128, 128, true, false, 1, 0>("TS", bt, n_sm);
    run<1, 128, 256, true, false, 1, 0>("TS", bt, n_sm);
    run<1, 128, 32, true, false, 4, 0>("TS", bt, n_sm);
    run<1, 128, 64, true, false, 4, 0>("TS", bt, n_sm);
    run<1, 64, 128, false, false, 1, 0>("SS M64", bt, n_sm);
    run<1, 64, 256, false, false, 1, 0>("SS M64", bt, n_sm);
    run<2, 256, 32, false, false, 1, 0>("2CTA SS", bt, n_sm);
    run<2, 256, 64, false, false, 1, 0>("2CTA SS", bt, n_sm);
    run<2, 256, 128, false, false, 1, 0>("2CTA SS", bt, n_sm);
    run<2, 256, 256, false, false, 1, 0>("2CTA SS", bt, n_sm);
    run<2, 256, 64, false, false, 4, 3>("2CTA SS taps", bt, n_sm);
    run<2, 256, 128, true, false, 1, 0>("2CTA TS", bt, n_sm);
    run<1, 128, 64, false, true, 1, 0>("SS tf32", bt, n_sm);
    run<1, 128, 128, false, true, 1, 0>("SS tf32", bt, n_sm);
    run<1, 128, 256, false, true, 1, 0>("SS tf32", bt, n_sm);
    run<2, 256, 256, false, true, 1, 0>("2CTA SS tf32", bt, n_sm);
    return 0;
}
