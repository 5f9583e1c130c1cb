// The block products of the top-k eigensolver on sm_100a (eigsh(covmat, k), conan.py:1985-1993).
//
// A few eigenpairs of the P' x P' covariance are found by a block Krylov iteration (conan_b200/stage3.py).
// Its work is three kinds of products with tall, skinny blocks of vectors (b <= 16 columns at a time, at most
// kMaxCols in a basis), none of which has enough arithmetic per byte for the tensor cores:
//   sym_block_matvec_kernel   Y = A W, A the float32 covariance [n][n] read once per product (the whole cost:
//                             4 n^2 bytes from HBM), W and Y float64 [n][b]; float64 multiply-adds, so the
//                             iteration converges to float64 accuracy on the float32 matrix
//   block_gram_kernel         G = U^T V, [mu][mv] from two [n][m] blocks (orthogonalisation, Rayleigh-Ritz)
//   block_combine_kernel      X = U S, a [n][m] block times a small [m][c] matrix (rotations of the basis)
// Everything is float64 except A; the small dense problems in between (at most kMaxCols x kMaxCols) are
// solved by the caller.
#include <cuda_runtime.h>

#include <cstdint>

#include "common.h"

namespace {

constexpr int kMaxB = 16;          // columns of W per sym_block_matvec launch
constexpr int kMaxCols = 160;      // columns of a basis block in gram / combine
constexpr int kRowsPerWarp = 4;
constexpr int kWarps = 8;
constexpr int kRowsPerCta = kRowsPerWarp * kWarps;     // 32 rows of A per CTA
constexpr int kJTile = 256;                            // columns of A (rows of W) staged per step

template <int B>
__global__ void __launch_bounds__(kWarps * 32) sym_block_matvec_kernel(const float* __restrict__ A, int64_t lda, int n,
                                                                       const double* __restrict__ W, int64_t ldw,
                                                                       double* __restrict__ Y, int64_t ldy, int b) {
    // W tile in shared memory, [kJTile][B] with one double of padding per row (odd stride in 8-byte units: the 32
    // lanes of a warp read 32 different rows at the same column)
    __shared__ double s_w[kJTile][B + 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int row0 = blockIdx.x * kRowsPerCta + warp * kRowsPerWarp;
    double acc[kRowsPerWarp][B];
#pragma unroll
    for (int r = 0; r < kRowsPerWarp; r++)
#pragma unroll
        for (int c = 0; c < B; c++) acc[r][c] = 0.0;
    for (int j0 = 0; j0 < n; j0 += kJTile) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < kJTile * B; idx += kWarps * 32) {
            const int j = idx / B, c = idx - j * B;
            s_w[j][c] = (j0 + j < n && c < b) ? W[(int64_t)(j0 + j) * ldw + c] : 0.0;
        }
        __syncthreads();
#pragma unroll 2
        for (int t = 0; t < kJTile / 32; t++) {
            const int j = j0 + t * 32 + lane;
            float a[kRowsPerWarp];
#pragma unroll
            for (int r = 0; r < kRowsPerWarp; r++)
                a[r] = (j < n && row0 + r < n) ? __ldg(A + (int64_t)(row0 + r) * lda + j) : 0.f;
#pragma unroll
            for (int c = 0; c < B; c++) {
                const double w = s_w[t * 32 + lane][c];
#pragma unroll
                for (int r = 0; r < kRowsPerWarp; r++) acc[r][c] = fma((double)a[r], w, acc[r][c]);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < kRowsPerWarp; r++)
#pragma unroll
        for (int c = 0; c < B; c++) {
            double v = acc[r][c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && row0 + r < n && c < b) Y[(int64_t)(row0 + r) * ldy + c] = v;
        }
}

// G[u][v] += sum over a slab of rows of U[i][u] * V[i][v]; one CTA per slab of kGramRows rows.
constexpr int kGramRows = 16;
__global__ void __launch_bounds__(256) block_gram_kernel(const double* __restrict__ U, int64_t ldu, int mu,
                                                         const double* __restrict__ V, int64_t ldv, int mv, int n,
                                                         double* __restrict__ G) {
    extern __shared__ double s_uv[];                 // [kGramRows][mu] then [kGramRows][mv]
    double* su = s_uv;
    double* sv = s_uv + kGramRows * mu;
    const int i0 = blockIdx.x * kGramRows;
    const int rows = min(kGramRows, n - i0);
    for (int idx = threadIdx.x; idx < kGramRows * mu; idx += 256) {
        const int r = idx / mu, c = idx - r * mu;
        su[idx] = r < rows ? U[(int64_t)(i0 + r) * ldu + c] : 0.0;
    }
    for (int idx = threadIdx.x; idx < kGramRows * mv; idx += 256) {
        const int r = idx / mv, c = idx - r * mv;
        sv[idx] = r < rows ? V[(int64_t)(i0 + r) * ldv + c] : 0.0;
    }
    __syncthreads();
    for (int o = threadIdx.x; o < mu * mv; o += 256) {
        const int a = o / mv, b = o - a * mv;
        double s = 0.0;
#pragma unroll
        for (int r = 0; r < kGramRows; r++) s = fma(su[r * mu + a], sv[r * mv + b], s);
        atomicAdd(G + o, s);
    }
}

// X[i][c] = sum_m U[i][m] S[m][c]; S in shared memory, one thread per output element of a slab of rows.
__global__ void __launch_bounds__(256) block_combine_kernel(const double* __restrict__ U, int64_t ldu, int m,
                                                            const double* __restrict__ S, int c, int n,
                                                            double* __restrict__ X, int64_t ldx) {
    extern __shared__ double s_s[];                  // [m][c]
    for (int idx = threadIdx.x; idx < m * c; idx += 256) s_s[idx] = S[idx];
    __syncthreads();
    const int64_t total = (int64_t)n * c;
    for (int64_t o = (int64_t)blockIdx.x * 256 + threadIdx.x; o < total; o += (int64_t)gridDim.x * 256) {
        const int64_t i = o / c;
        const int col = (int)(o - i * c);
        const double* u = U + i * ldu;
        double s = 0.0;
        for (int k = 0; k < m; k++) s = fma(u[k], s_s[k * c + col], s);
        X[i * ldx + col] = s;
    }
}

}  // namespace

extern "C" int conan_sym_block_matvec(const float* d_A, int64_t lda, int32_t n, const double* d_W, int64_t ldw, int32_t b,
                                      double* d_Y, int64_t ldy, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_A || !d_W || !d_Y || n <= 0 || b <= 0 || lda < n || ldw < b || ldy < b)
        return conan::fail(CONAN_ERR_ARG, "conan_sym_block_matvec: bad arguments");
    if (b > kMaxB) return conan::fail(CONAN_ERR_SIZE, "conan_sym_block_matvec: b=%d columns, at most %d per call", b, kMaxB);
    const unsigned grid = (unsigned)((n + kRowsPerCta - 1) / kRowsPerCta);
    if (b <= 4) sym_block_matvec_kernel<4><<<grid, kWarps * 32, 0, stream>>>(d_A, lda, n, d_W, ldw, d_Y, ldy, b);
    else if (b <= 8) sym_block_matvec_kernel<8><<<grid, kWarps * 32, 0, stream>>>(d_A, lda, n, d_W, ldw, d_Y, ldy, b);
    else if (b <= 12) sym_block_matvec_kernel<12><<<grid, kWarps * 32, 0, stream>>>(d_A, lda, n, d_W, ldw, d_Y, ldy, b);
    else sym_block_matvec_kernel<16><<<grid, kWarps * 32, 0, stream>>>(d_A, lda, n, d_W, ldw, d_Y, ldy, b);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

extern "C" int conan_block_gram(const double* d_U, int64_t ldu, int32_t mu, const double* d_V, int64_t ldv, int32_t mv,
                                int32_t n, double* d_G, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_U || !d_V || !d_G || n <= 0 || mu <= 0 || mv <= 0 || ldu < mu || ldv < mv)
        return conan::fail(CONAN_ERR_ARG, "conan_block_gram: bad arguments");
    if (mu > kMaxCols || mv > kMaxCols)
        return conan::fail(CONAN_ERR_SIZE, "conan_block_gram: %d x %d columns, at most %d each", mu, mv, kMaxCols);
    CONAN_CUDA_TRY(cudaMemsetAsync(d_G, 0, sizeof(double) * (size_t)mu * mv, stream));
    const size_t smem = sizeof(double) * (size_t)kGramRows * (mu + mv);
    block_gram_kernel<<<(unsigned)((n + kGramRows - 1) / kGramRows), 256, smem, stream>>>(d_U, ldu, mu, d_V, ldv, mv, n, d_G);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

extern "C" int conan_block_combine(const double* d_U, int64_t ldu, int32_t m, const double* d_S, int32_t c, int32_t n,
                                   double* d_X, int64_t ldx, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_U || !d_S || !d_X || n <= 0 || m <= 0 || c <= 0 || ldu < m || ldx < c)
        return conan::fail(CONAN_ERR_ARG, "conan_block_combine: bad arguments");
    if (m > kMaxCols || c > kMaxCols)
        return conan::fail(CONAN_ERR_SIZE, "conan_block_combine: %d x %d, at most %d each", m, c, kMaxCols);
    if (d_X == d_U) return conan::fail(CONAN_ERR_ARG, "conan_block_combine: the product cannot be formed in place");
    const size_t smem = sizeof(double) * (size_t)m * c;
    static bool raised = false;
    if (smem > 48 * 1024 && !raised) {
        CONAN_CUDA_TRY(cudaFuncSetAttribute(block_combine_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        raised = true;
    }
    const int64_t total = (int64_t)n * c;
    const unsigned grid = (unsigned)((total + 255) / 256 > 148 * 16 ? 148 * 16 : (total + 255) / 256);
    block_combine_kernel<<<grid, 256, smem, stream>>>(d_U, ldu, m, d_S, c, n, d_X, ldx);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
