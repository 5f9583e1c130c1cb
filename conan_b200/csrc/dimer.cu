// DIMER mode of stage 2 on sm_100a (conan.py:553-565, :1420-1428).
//
// In a homodimer run the X and the Y residues carry the same labels, so the n x n block of a frame and its
// transpose describe the same physical contact pattern with the two chains exchanged.  The reference
// orients every frame as it is read: if the weighted contacts (trunc - r) / trunc below the diagonal
// outweigh those above it, the frame is transposed (read_frame, :553-565); all per-pair statistics then
// run on the oriented frames.  Its rmsd_frame additionally takes the smaller of the distance to a frame
// and to that frame's transpose (:1420-1428).
//   dimer_orient_kernel   one CTA per frame: the two weighted sums as exact integers in level units
//                         (maxlevel - k), then an in-place transpose of the frames that need it.
//   dimer_ssd_kernel      one CTA per frame: sum (k_f[x][y] - k_ref[y][x])^2 against the previous frame and
//                         against global frame 0, the transposed twins of conan_pair_stats' per-frame sums.
#include <cuda_runtime.h>

#include <cstdint>

#include "common.h"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ unsigned long long block_sum(unsigned long long v, unsigned long long* s_red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) s_red[warp] = v;
    __syncthreads();
    unsigned long long t = 0;
    for (int w = 0; w < kThreads / 32; w++) t += s_red[w];
    return t;
}

__global__ void __launch_bounds__(kThreads) dimer_orient_kernel(uint16_t* __restrict__ levels, int64_t stride, int n, int maxlevel,
                                                                uint8_t* __restrict__ flipped) {
    __shared__ unsigned long long s_red[kThreads / 32];
    uint16_t* fr = levels + (int64_t)blockIdx.x * stride;
    unsigned long long above = 0, below = 0;
    for (int idx = threadIdx.x; idx < n * n; idx += kThreads) {
        const int x = idx / n, y = idx - x * n;
        const unsigned w = (unsigned)(maxlevel - (int)fr[idx]);        // 0 for a pair beyond trunc
        if (x < y) above += w;
        else if (x > y) below += w;
    }
    above = block_sum(above, s_red);
    below = block_sum(below, s_red);
    const bool flip = above < below;
    if (threadIdx.x == 0 && flipped) flipped[blockIdx.x] = flip ? 1 : 0;
    if (!flip) return;                                                   // (block-uniform)
    for (int idx = threadIdx.x; idx < n * n; idx += kThreads) {
        const int x = idx / n, y = idx - x * n;
        if (x < y) {
            const uint16_t a = fr[idx], b = fr[y * n + x];
            fr[idx] = b;
            fr[y * n + x] = a;
        }
    }
}

__global__ void __launch_bounds__(kThreads) dimer_ssd_kernel(const uint16_t* __restrict__ levels, int64_t stride, int n,
                                                             const uint16_t* __restrict__ first_row,
                                                             const uint16_t* __restrict__ prev_row,
                                                             long long* __restrict__ ssd_prev_t,
                                                             long long* __restrict__ ssd_first_t) {
    __shared__ unsigned long long s_red[kThreads / 32];
    const int64_t f = blockIdx.x;
    const uint16_t* fr = levels + f * stride;
    const uint16_t* pv = f > 0 ? fr - stride : prev_row;               // null: no frame before this one
    unsigned long long sp = 0, s0 = 0;
    for (int idx = threadIdx.x; idx < n * n; idx += kThreads) {
        const int x = idx / n, y = idx - x * n;
        const int k = fr[idx];
        const int d0 = k - (int)first_row[y * n + x];
        s0 += (unsigned long long)((long long)d0 * d0);
        if (pv) {
            const int dp = k - (int)pv[y * n + x];
            sp += (unsigned long long)((long long)dp * dp);
        }
    }
    sp = block_sum(sp, s_red);
    s0 = block_sum(s0, s_red);
    if (threadIdx.x == 0) { ssd_prev_t[f] = (long long)sp; ssd_first_t[f] = (long long)s0; }
}

}  // namespace

extern "C" int conan_dimer_orient(uint16_t* d_levels, int64_t levels_stride, int64_t nf, int32_t n, int32_t nlevels,
                                  uint8_t* d_flipped, void* stream) {
    if (!d_levels || nf < 0 || n <= 0 || (int64_t)n * n > levels_stride)
        return conan::fail(CONAN_ERR_ARG, "conan_dimer_orient: bad arguments (the stack must hold n x n blocks)");
    if (nlevels < 2 || nlevels > 7744) return conan::fail(CONAN_ERR_SIZE, "conan_dimer_orient: nlevels=%d outside [2, 7744]", nlevels);
    if (nf == 0) return CONAN_OK;
    if (nf > 2147483647ll) return conan::fail(CONAN_ERR_SIZE, "conan_dimer_orient: too many frames for one launch");
    dimer_orient_kernel<<<(unsigned)nf, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(d_levels, levels_stride, n, nlevels - 1, d_flipped);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

extern "C" int conan_dimer_ssd(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int32_t n,
                               const uint16_t* d_first_row, const uint16_t* d_prev_row, int64_t* d_ssd_prev_t,
                               int64_t* d_ssd_first_t, void* stream) {
    if (!d_levels || !d_ssd_prev_t || !d_ssd_first_t || nf < 0 || n <= 0 || (int64_t)n * n > levels_stride)
        return conan::fail(CONAN_ERR_ARG, "conan_dimer_ssd: bad arguments (the stack must hold n x n blocks)");
    if (nf == 0) return CONAN_OK;
    if (nf > 2147483647ll) return conan::fail(CONAN_ERR_SIZE, "conan_dimer_ssd: too many frames for one launch");
    dimer_ssd_kernel<<<(unsigned)nf, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        d_levels, levels_stride, n, d_first_row ? d_first_row : d_levels, d_prev_row,
        reinterpret_cast<long long*>(d_ssd_prev_t), reinterpret_cast<long long*>(d_ssd_first_t));
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
