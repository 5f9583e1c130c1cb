// Stage 3 feeders on sm_100a: the streaming Pearson sums and the operand builders of the GEMMs.
//
// The reference keeps `vectors`, a float16 [pairs x frames] array of legend values
// (conan.py:731, :821-826), and runs scipy.stats.linregress once per residue pair against time
// (:1888-1919) and against every observable (:1922-1971), np.cov for the PCA (:1983) and
// scipy pdist for the frame-frame distances (:2063).  Here the float16 values are looked up from
// the uint16 level stack on the fly (a table of at most 7744 entries), so `vectors` is never
// materialised:
//   pearson_sums_kernel   for each pair p and each of J frame weights / centred observables O_j:
//                         T[j][p] = sum_f O_j[f] * (y_pf - c_p) and Q[p] = sum_f O_0[f] * (y_pf - c_p)^2
//                         (O_0 is the frame mask w).  A streaming pass, 2 B per pair-frame, HBM-bound;
//                         float32 inside a 32-frame window, float64 across windows; the shift c_p
//                         (about the pair's mean) removes the cancellation from the moments.
//   operand_pairs_kernel  rows = selected pairs, K = frames: centred float32 split into TF32 hi / lo
//                         (the A = B operand of the covariance SYRK), a tiled transpose of the stack.
//   operand_frames_kernel rows = frames, K = selected pairs: centred on the mean frame (distances
//                         are translation invariant) and scaled by 1/norm (the operand of the
//                         frame-frame Gram matrix), plus the squared row norms.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "common.h"

namespace {

constexpr int kThreads = 256;
constexpr int kVec = 4;
constexpr int kWin = 32;
constexpr int kMaxJ = 8;      // frame weights + observables per pass
constexpr int kMaxLevels = 7744;

__device__ __forceinline__ uint2 ld_stream(const uint16_t* p) {
    uint2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}

struct PearsonParams {
    const uint16_t* levels;
    int64_t stride, nf, P;
    const float* lut;          // [nlevels] float16 legend values widened to float
    int nlevels;
    const float* obs;          // [J][nf] O_0 = frame weights (0/1), O_j = centred observables * weight
    int J;
    const float* shift;        // [P] c_p
    double* T;                 // [J][P]
    double* Q;                 // [P]
    int64_t chunk_len;
};

// Packed FP32 pairs (FFMA2 / FMUL2): two fused multiply-adds per issue slot, each rounded like the scalar one.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// The kernel is bound by instruction issue before it is bound by HBM (a dozen instructions per 2-byte level with
// four weights), so the sums of a thread's four pairs run as two packed pairs: one FFMA2 per weight and pair of
// pairs; the weights of a frame sit together in shared memory ([frame][8]) and come with one or two 16-byte
// reads; eight frames are in flight per thread.
template <int J>
__global__ void __launch_bounds__(kThreads) pearson_sums_kernel(const PearsonParams prm) {
    extern __shared__ __align__(16) float s_mem[];
    float* s_obs = s_mem;                      // [kWin][8]
    float* s_lut = s_mem + kWin * 8;           // [nlevels]
    const int tid = threadIdx.x;
    for (int i = tid; i < prm.nlevels; i += kThreads) s_lut[i] = prm.lut[i];
    const int64_t p0 = ((int64_t)blockIdx.x * kThreads + tid) * kVec;
    const bool in_range = p0 < prm.stride;
    const int64_t fbeg = (int64_t)blockIdx.y * prm.chunk_len;
    const int64_t fend = min(prm.nf, fbeg + prm.chunk_len);
    float c[kVec];
#pragma unroll
    for (int q = 0; q < kVec; q++) c[q] = (p0 + q < prm.P) ? prm.shift[p0 + q] : 0.f;
    const f32x2 nc01 = pack2(-c[0], -c[1]), nc23 = pack2(-c[2], -c[3]), one2 = pack2(1.f, 1.f);
    double accT[J][kVec], accQ[kVec];
#pragma unroll
    for (int q = 0; q < kVec; q++) {
        accQ[q] = 0.0;
#pragma unroll
        for (int j = 0; j < J; j++) accT[j][q] = 0.0;
    }
    const uint16_t* col = prm.levels + (in_range ? p0 : 0);
    const int maxk = prm.nlevels - 1;
    for (int64_t fb = fbeg; fb < fend; fb += kWin) {
        const int nblk = (int)min((int64_t)kWin, fend - fb);
        __syncthreads();
        for (int i = tid; i < kWin * 8; i += kThreads) {
            const int fl = i >> 3, j = i & 7;
            s_obs[i] = (j < J && fl < nblk) ? prm.obs[(size_t)j * prm.nf + fb + fl] : 0.f;
        }
        __syncthreads();
        f32x2 wT01[J], wT23[J], wQ01 = 0ull, wQ23 = 0ull;      // (the bit pattern of two +0.0f)
#pragma unroll
        for (int j = 0; j < J; j++) { wT01[j] = 0ull; wT23[j] = 0ull; }
#pragma unroll 8
        for (int fl = 0; fl < nblk; fl++) {
            const uint2 w = ld_stream(col + (fb + fl) * prm.stride);
            const float4 oa = *reinterpret_cast<const float4*>(s_obs + fl * 8);
            float4 ob = make_float4(0.f, 0.f, 0.f, 0.f);
            if (J > 4) ob = *reinterpret_cast<const float4*>(s_obs + fl * 8 + 4);
            const float o[8] = {oa.x, oa.y, oa.z, oa.w, ob.x, ob.y, ob.z, ob.w};
            const float v0 = s_lut[min((int)(w.x & 0xffffu), maxk)], v1 = s_lut[min((int)(w.x >> 16), maxk)];
            const float v2 = s_lut[min((int)(w.y & 0xffffu), maxk)], v3 = s_lut[min((int)(w.y >> 16), maxk)];
            // d = value - c, as one packed add (x * 1 + (-c) is exact in the product and rounds once, like the subtraction)
            const f32x2 d01 = fma2(pack2(v0, v1), one2, nc01), d23 = fma2(pack2(v2, v3), one2, nc23);
            const f32x2 O0 = pack2(o[0], o[0]);
            wQ01 = fma2(mul2(O0, d01), d01, wQ01);
            wQ23 = fma2(mul2(O0, d23), d23, wQ23);
#pragma unroll
            for (int j = 0; j < J; j++) {
                const f32x2 Oj = pack2(o[j], o[j]);
                wT01[j] = fma2(Oj, d01, wT01[j]);
                wT23[j] = fma2(Oj, d23, wT23[j]);
            }
        }
        float q0, q1, q2, q3;
        unpack2(wQ01, q0, q1); unpack2(wQ23, q2, q3);
        accQ[0] += (double)q0; accQ[1] += (double)q1; accQ[2] += (double)q2; accQ[3] += (double)q3;
#pragma unroll
        for (int j = 0; j < J; j++) {
            float t0, t1, t2, t3;
            unpack2(wT01[j], t0, t1); unpack2(wT23[j], t2, t3);
            accT[j][0] += (double)t0; accT[j][1] += (double)t1; accT[j][2] += (double)t2; accT[j][3] += (double)t3;
        }
    }
#pragma unroll
    for (int q = 0; q < kVec; q++) {
        const int64_t p = p0 + q;
        if (p < prm.P) {
            atomicAdd(&prm.Q[p], accQ[q]);
#pragma unroll
            for (int j = 0; j < J; j++) atomicAdd(&prm.T[(size_t)j * prm.P + p], accT[j][q]);
        }
    }
}

// Smallest and largest level of every pair over the frames with mask != 0 (all frames if mask is
// null): decides len(set(rvec)) > 1 exactly (conan.py:1902, :1949).
__global__ void __launch_bounds__(kThreads) level_range_kernel(const uint16_t* __restrict__ levels, int64_t stride,
                                                               int64_t nf, int64_t P, const uint8_t* __restrict__ mask,
                                                               int64_t chunk_len, unsigned* __restrict__ kmin,
                                                               unsigned* __restrict__ kmax) {
    const int64_t p0 = ((int64_t)blockIdx.x * kThreads + threadIdx.x) * kVec;
    if (p0 >= stride) return;
    const int64_t fbeg = (int64_t)blockIdx.y * chunk_len, fend = min(nf, fbeg + chunk_len);
    unsigned mnx = 0xffffffffu, mny = 0xffffffffu, mxx = 0, mxy = 0;
    const uint16_t* col = levels + p0;
    for (int64_t f = fbeg; f < fend; f++) {
        if (mask && !mask[f]) continue;
        const uint2 w = ld_stream(col + f * stride);
        mnx = __vminu2(mnx, w.x); mny = __vminu2(mny, w.y);
        mxx = __vmaxu2(mxx, w.x); mxy = __vmaxu2(mxy, w.y);
    }
    const unsigned mn[kVec] = {mnx & 0xffffu, mnx >> 16, mny & 0xffffu, mny >> 16};
    const unsigned mx[kVec] = {mxx & 0xffffu, mxx >> 16, mxy & 0xffffu, mxy >> 16};
#pragma unroll
    for (int q = 0; q < kVec; q++)
        if (p0 + q < P) { atomicMin(&kmin[p0 + q], mn[q]); atomicMax(&kmax[p0 + q], mx[q]); }
}

// Per frame and residue: sum over the residue's pairs of (maxlevel - level), i.e. the weighted number
// of contacts (trunc - r)/trunc of conan.py:19-21, :843-850 in level units (exact integers).
// One CTA per frame; each pair is read once and credited to both of its residues in shared memory.
__global__ void residue_weights_kernel(const uint16_t* __restrict__ levels, int64_t stride, int symmetric, int xres,
                                       int yres, int maxlevel, int* __restrict__ out /* [nf][xres (+ yres)] */) {
    extern __shared__ int s_w[];
    const int nout = symmetric ? xres : xres + yres;
    for (int i = threadIdx.x; i < nout; i += blockDim.x) s_w[i] = 0;
    __syncthreads();
    const uint16_t* row = levels + (size_t)blockIdx.x * stride;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int i = warp; i < xres; i += nwarp) {
        int64_t base;
        int j0, j1;
        if (symmetric) { base = (int64_t)i * xres - (int64_t)i * (i + 1) / 2 - i - 1; j0 = i + 1; j1 = xres; }
        else { base = (int64_t)i * yres; j0 = 0; j1 = yres; }
        int acc = 0;
        for (int j = j0 + lane; j < j1; j += 32) {
            const int w = maxlevel - min((int)row[base + j], maxlevel);
            if (w) {
                acc += w;
                atomicAdd(&s_w[symmetric ? j : xres + j], w);
            }
        }
        acc = __reduce_add_sync(0xffffffffu, acc);
        if (lane == 0 && acc) atomicAdd(&s_w[i], acc);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nout; i += blockDim.x) out[(size_t)blockIdx.x * nout + i] = s_w[i];
}

// x = hi + lo with both parts binary16: hi carries the leading 11 significant bits, lo the next 11 (the caller
// scales the operand so that |x| stays near 2^9 and neither part leaves the binary16 range where it matters).
__device__ __forceinline__ void split_half(float x, __half& h, __half& l) {
    h = __float2half_rn(x);
    l = __float2half_rn(x - __half2float(h));
}

// rows = selected pairs, K = frames.  32 x 32 tiles through shared memory (a transpose of the stack).
__global__ void operand_pairs_kernel(const uint16_t* __restrict__ levels, int64_t stride, int64_t nf,
                                     const int* __restrict__ sel, int nsel, const float* __restrict__ lut, int nlevels,
                                     const float* __restrict__ mean, float scale, __half* __restrict__ hi,
                                     __half* __restrict__ lo, int64_t ld /* elements per output row, multiple of 64 */) {
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8 threads
    const int64_t f0 = (int64_t)blockIdx.x * 32;
    const int r0 = blockIdx.y * 32;
    for (int y = ty; y < 32; y += 8) {                        // y: frame inside the tile, tx: pair
        const int64_t f = f0 + y;
        const int r = r0 + tx;
        float v = 0.f;
        if (f < nf && r < nsel) v = (lut[min((int)levels[f * stride + sel[r]], nlevels - 1)] - mean[r]) * scale;
        tile[y][tx] = v;
    }
    __syncthreads();
    for (int y = ty; y < 32; y += 8) {                        // y: pair inside the tile, tx: frame
        const int r = r0 + y;
        const int64_t f = f0 + tx;
        if (r < nsel && f < ld) {
            const float v = f < nf ? tile[tx][y] : 0.f;
            __half h, l;
            split_half(v, h, l);
            hi[(size_t)r * ld + f] = h;
            lo[(size_t)r * ld + f] = l;
        }
    }
}

// rows = frames, K = selected pairs; also the squared norm of every row (float64).
__global__ void operand_frames_kernel(const uint16_t* __restrict__ levels, int64_t stride, int64_t nf,
                                      const int* __restrict__ sel, int nsel, const float* __restrict__ lut, int nlevels,
                                      const float* __restrict__ mean, float scale, __half* __restrict__ hi,
                                      __half* __restrict__ lo, int64_t ld, double* __restrict__ norm2) {
    const int64_t f = blockIdx.x;
    double acc = 0.0;
    for (int64_t c = threadIdx.x; c < ld; c += blockDim.x) {
        float v = 0.f;
        if (c < nsel) v = (lut[min((int)levels[f * stride + sel[c]], nlevels - 1)] - mean[c]) * scale;
        __half h, l;
        split_half(v, h, l);
        hi[(size_t)f * ld + c] = h;
        lo[(size_t)f * ld + c] = l;
        const double e = (double)__half2float(h) + (double)__half2float(l);   // the value the GEMM effectively uses
        acc += e * e;
    }
    __shared__ double red[32];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += red[w];
        norm2[f] = s;
    }
}

}  // namespace

extern "C" int conan_pearson_sums(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                  const float* d_lut, int32_t nlevels, const float* d_obs, int32_t J,
                                  const float* d_shift, double* d_T, double* d_Q, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_levels || !d_lut || !d_obs || !d_shift || !d_T || !d_Q)
        return conan::fail(CONAN_ERR_ARG, "conan_pearson_sums: null pointer argument");
    if (nf <= 0 || P <= 0 || levels_stride < P || levels_stride % kVec != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_pearson_sums: bad sizes nf=%lld P=%lld stride=%lld", (long long)nf,
                           (long long)P, (long long)levels_stride);
    if ((reinterpret_cast<uintptr_t>(d_levels) & 7) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_pearson_sums: level rows must be 8-byte aligned");
    if (nlevels < 2 || nlevels > kMaxLevels)
        return conan::fail(CONAN_ERR_SIZE, "conan_pearson_sums: nlevels=%d outside [2, %d]", nlevels, kMaxLevels);
    if (J < 1 || J > kMaxJ)
        return conan::fail(CONAN_ERR_SIZE, "conan_pearson_sums: J=%d outside [1, %d]; split the observables", J, kMaxJ);
    CONAN_CUDA_TRY(cudaMemsetAsync(d_T, 0, sizeof(double) * (size_t)J * P, stream));
    CONAN_CUDA_TRY(cudaMemsetAsync(d_Q, 0, sizeof(double) * (size_t)P, stream));
    PearsonParams prm;
    prm.levels = d_levels; prm.stride = levels_stride; prm.nf = nf; prm.P = P; prm.lut = d_lut; prm.nlevels = nlevels;
    prm.obs = d_obs; prm.J = J; prm.shift = d_shift; prm.T = d_T; prm.Q = d_Q;
    const int64_t tiles = (P + kThreads * kVec - 1) / (kThreads * kVec);
    int64_t nchunks = (148 * 16 + tiles - 1) / tiles;
    const int64_t maxc = nf / 64 > 0 ? nf / 64 : 1;
    if (nchunks > maxc) nchunks = maxc;
    if (nchunks < 1) nchunks = 1;
    if (nchunks > 65535) nchunks = 65535;
    int64_t clen = (nf + nchunks - 1) / nchunks;
    clen = (clen + kWin - 1) / kWin * kWin;
    nchunks = (nf + clen - 1) / clen;
    prm.chunk_len = clen;
    const size_t smem = sizeof(float) * ((size_t)nlevels + (size_t)8 * kWin);
    dim3 grid((unsigned)tiles, (unsigned)nchunks);
#define CONAN_LAUNCH_PEARSON(JJ)                                                                              \
    case JJ:                                                                                                   \
        CONAN_CUDA_TRY(cudaFuncSetAttribute(pearson_sums_kernel<JJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                            (int)smem));                                                       \
        pearson_sums_kernel<JJ><<<grid, kThreads, smem, stream>>>(prm);                                        \
        break;
    switch (J) {
        CONAN_LAUNCH_PEARSON(1) CONAN_LAUNCH_PEARSON(2) CONAN_LAUNCH_PEARSON(3) CONAN_LAUNCH_PEARSON(4)
        CONAN_LAUNCH_PEARSON(5) CONAN_LAUNCH_PEARSON(6) CONAN_LAUNCH_PEARSON(7) CONAN_LAUNCH_PEARSON(8)
    }
#undef CONAN_LAUNCH_PEARSON
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

extern "C" int conan_build_operand(const uint16_t* d_levels, int64_t levels_stride, int64_t nf,
                                   const int32_t* d_sel, int32_t nsel, const float* d_lut, int32_t nlevels,
                                   const float* d_mean, float scale, int32_t rows_are_frames, uint16_t* d_hi,
                                   uint16_t* d_lo, int64_t ld, double* d_norm2, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_levels || !d_sel || !d_lut || !d_mean || !d_hi || !d_lo)
        return conan::fail(CONAN_ERR_ARG, "conan_build_operand: null pointer argument");
    if (nf <= 0 || nsel <= 0 || nlevels < 2 || nlevels > kMaxLevels)
        return conan::fail(CONAN_ERR_ARG, "conan_build_operand: bad sizes");
    const int64_t k = rows_are_frames ? nsel : nf;
    if (ld < k || ld % 64 != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_build_operand: ld=%lld must be a multiple of 64 and >= %lld", (long long)ld, (long long)k);
    if (!(scale > 0.f)) return conan::fail(CONAN_ERR_ARG, "conan_build_operand: scale must be positive");
    __half* hi = reinterpret_cast<__half*>(d_hi);
    __half* lo = reinterpret_cast<__half*>(d_lo);
    if (rows_are_frames) {
        if (!d_norm2) return conan::fail(CONAN_ERR_ARG, "conan_build_operand: frame rows need d_norm2");
        operand_frames_kernel<<<(unsigned)nf, 256, 0, stream>>>(d_levels, levels_stride, nf, d_sel, nsel, d_lut, nlevels,
                                                               d_mean, scale, hi, lo, ld, d_norm2);
    } else {
        dim3 grid((unsigned)(ld / 32), (unsigned)((nsel + 31) / 32));
        operand_pairs_kernel<<<grid, 256, 0, stream>>>(d_levels, levels_stride, nf, d_sel, nsel, d_lut, nlevels, d_mean,
                                                       scale, hi, lo, ld);
    }
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

extern "C" int conan_level_range(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                 const uint8_t* d_frame_mask, uint32_t* d_kmin, uint32_t* d_kmax, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_levels || !d_kmin || !d_kmax) return conan::fail(CONAN_ERR_ARG, "conan_level_range: null pointer argument");
    if (nf <= 0 || P <= 0 || levels_stride < P || levels_stride % kVec != 0 || (reinterpret_cast<uintptr_t>(d_levels) & 7) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_level_range: bad sizes or alignment");
    CONAN_CUDA_TRY(cudaMemsetAsync(d_kmin, 0xff, sizeof(uint32_t) * (size_t)P, stream));
    CONAN_CUDA_TRY(cudaMemsetAsync(d_kmax, 0, sizeof(uint32_t) * (size_t)P, stream));
    const int64_t tiles = (P + kThreads * kVec - 1) / (kThreads * kVec);
    int64_t nchunks = (148 * 16 + tiles - 1) / tiles;
    if (nchunks > nf / 64) nchunks = nf / 64;
    if (nchunks < 1) nchunks = 1;
    if (nchunks > 65535) nchunks = 65535;
    const int64_t clen = (nf + nchunks - 1) / nchunks;
    nchunks = (nf + clen - 1) / clen;
    dim3 grid((unsigned)tiles, (unsigned)nchunks);
    level_range_kernel<<<grid, kThreads, 0, stream>>>(d_levels, levels_stride, nf, P, d_frame_mask, clen, d_kmin, d_kmax);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

extern "C" int conan_residue_weights(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int32_t symmetric,
                                     int32_t xres, int32_t yres, int32_t nlevels, int32_t* d_out, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_levels || !d_out) return conan::fail(CONAN_ERR_ARG, "conan_residue_weights: null pointer argument");
    if (nf <= 0 || xres <= 0 || (!symmetric && yres <= 0) || nlevels < 2 || nlevels > kMaxLevels)
        return conan::fail(CONAN_ERR_ARG, "conan_residue_weights: bad sizes");
    const int64_t P = symmetric ? (int64_t)xres * (xres - 1) / 2 : (int64_t)xres * yres;
    if (levels_stride < P) return conan::fail(CONAN_ERR_ARG, "conan_residue_weights: levels_stride < P");
    const int nout = symmetric ? xres : xres + yres;
    if (nout > 48000 || nf > 2147483647ll)
        return conan::fail(CONAN_ERR_SIZE, "conan_residue_weights: more than 48000 residues");
    const size_t smem = sizeof(int) * (size_t)nout;
    CONAN_CUDA_TRY(cudaFuncSetAttribute(residue_weights_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    residue_weights_kernel<<<(unsigned)nf, 256, smem, stream>>>(d_levels, levels_stride, symmetric ? 1 : 0, xres, yres,
                                                               nlevels - 1, d_out);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
