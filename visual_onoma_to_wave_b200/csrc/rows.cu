// fp32 helpers on time-major rows (rows.cuh): tiled transpose, nn.Linear, residual + LayerNorm + mask, table add.
#include "rows.cuh"

namespace {

// [B,R,C] -> [B,C,R] (32 x 32 tiles through shared memory); optional `add` [B,C,R] is added to the result
__global__ void __launch_bounds__(256) transpose_add_kernel(const float* __restrict__ in, const float* __restrict__ add,
                                                            float* __restrict__ out, int R, int C) {
    __shared__ float tile[32][33];
    const int b = blockIdx.z, r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const float* ib = in + (size_t)b * R * C;
    for (int i = ty; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + tx;
        tile[i][tx] = (r < R && c < C) ? ib[(size_t)r * C + c] : 0.f;
    }
    __syncthreads();
    float* ob = out + (size_t)b * R * C;
    const float* ab = add ? add + (size_t)b * R * C : nullptr;
    for (int i = ty; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + tx;
        if (r < R && c < C) {
            float v = tile[tx][i];
            if (ab) v += ab[(size_t)c * R + r];
            ob[(size_t)c * R + r] = v;
        }
    }
}


// y[r, n] = sum_k x[r, k] * w[n, k] + b[n]   (nn.Linear on the last axis; rows = every leading axis flattened).
// One CTA: 32 rows x all N outputs; x rows staged in shared memory, each warp owns 4 rows, lanes walk n.
constexpr int LIN_ROWS = 32;
__global__ void __launch_bounds__(256) linear_rows_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          const float* __restrict__ bias, float* __restrict__ y,
                                                          long long rows, int K, int N) {
    extern __shared__ float xs[];   // [LIN_ROWS][K]
    const long long r0 = (long long)blockIdx.x * LIN_ROWS;
    const int n_rows = (int)min((long long)LIN_ROWS, rows - r0);
    for (int e = threadIdx.x; e < n_rows * K; e += 256) xs[e] = x[(size_t)r0 * K + e];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int n = lane; n < N; n += 32) {
        const float* wn = w + (size_t)n * K;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int kk = 0; kk < K; ++kk) {
            const float wv = __ldg(wn + kk);
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[i] = fmaf(xs[(warp * 4 + i) * K + kk], wv, acc[i]);
        }
        const float bv = bias ? __ldg(bias + n) : 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = warp * 4 + i;
            if (r < n_rows) y[(size_t)(r0 + r) * N + n] = acc[i] + bv;
        }
    }
}



__global__ void __launch_bounds__(256) add_table_kernel(const float* __restrict__ x, const float* __restrict__ table,
                                                        float* __restrict__ y, long long n, long long per_b) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i < n) y[i] = x[i] + table[i % per_b];
}

}  // namespace

int rows_transpose_add(const float* in, const float* add, float* out, int B, int R, int C, cudaStream_t st) {
    dim3 grid((unsigned)vo2w_ceil_div(C, 32), (unsigned)vo2w_ceil_div(R, 32), (unsigned)B);
    transpose_add_kernel<<<grid, 256, 0, st>>>(in, add, out, R, C);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int rows_linear(const float* x, const float* w, const float* bias, long long rows, int K, int N, float* y, cudaStream_t st) {
    VO2W_REQUIRE(rows > 0 && K > 0 && N > 0 && (size_t)LIN_ROWS * K * 4 <= 160 * 1024, "linear: bad shape (rows %lld K %d N %d)", rows, K, N);
    static bool attr_set[VO2W_MAX_DEV];
    const int cur = vo2w_cur_dev();
    if (!attr_set[cur]) {
        VO2W_CUDA_OK(cudaFuncSetAttribute(linear_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
        attr_set[cur] = true;
    }
    const long long blocks = (rows + LIN_ROWS - 1) / LIN_ROWS;
    VO2W_REQUIRE(blocks <= 0x7fffffffLL, "linear: too many rows");
    linear_rows_kernel<<<(unsigned)blocks, 256, (size_t)LIN_ROWS * K * 4, st>>>(x, w, bias, y, rows, K, N);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}

int rows_add_table(const float* x, const float* table, float* y, int B, int L, int D, cudaStream_t st) {
    const long long n = (long long)B * L * D;
    add_table_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, table, y, n, (long long)L * D);
    VO2W_CUDA_OK(cudaGetLastError());
    return VO2W_OK;
}
