// Stage 2 on sm_100a: streaming per-pair statistics over the uint16 level stack.
//
// Replaces the per-frame Python loop of read_process_frames (conan.py:782-874): update_aggregate
// (:599-611) with the hysteresis of update_interaction (:24-45), the native / non-native contact
// counts (:841-854) and the two rmsd_frame calls of plot_frames (:643-644, :1414-1432).  All
// accumulators are integers over level indices, so results are exact and independent of the
// order in which frames, chunks or GPUs are merged; the host turns them into the reference's
// float64 tables.
//
// Decomposition: the stack is [nf][stride] frame-major, so a warp reading consecutive pairs of
// one frame is coalesced and a thread walking down the frames of its own pairs sees the time
// series in order (which the hysteresis needs).  The grid is (pair tiles) x (frame chunks).  A
// chunk does not know the interaction status entering it, so it runs under the hypothesis
// "entering status = off" and records how many leading frames sat inside the hysteresis buffer
// (`prefix`): those are exactly the frames whose status would differ under the other hypothesis.
// combine_kernel then walks the chunk records of each pair in order, applies the carry, and adds
// the per-frame corrections through a difference array.  The same record format carries the
// status across GPUs (frame shards).
//
// The kernel is HBM-bound in principle (2 B read per residue-pair-frame); see DESIGN.md for the
// instruction budget that decides how close it gets.
#include <cuda_runtime.h>

#include <cstdint>

#include "common.h"

namespace {

constexpr int kThreads = 256;
constexpr int kVec = 4;                       // pairs per thread (one 8-byte load per frame)
constexpr int kPairsPerCta = kThreads * kVec;
constexpr int kFrameBlock = 32;               // frames between flushes of the per-frame accumulators
constexpr int kUnroll = 8;                    // loads in flight per thread

// One record per (chunk, pair), struct-of-arrays in the workspace.
struct Records {
    unsigned long long* sum_k;    // [nchunks][Ppad]
    unsigned long long* sum_k2;
    int* n_int;
    int* n_enc;
    int* first_f;                 // global frame index, -1 = never on (under "enter off")
    int* last_f;
    int* prefix;                  // leading frames inside the buffer (== chunk length if undecided)
    uint16_t* kmin;
    uint16_t* kmax;
    uint8_t* flags;               // bit0: status leaving the chunk (enter off), bit1: decided
    int64_t Ppad;
};

struct StatsParams {
    const uint16_t* levels;
    int64_t stride;               // elements between frames
    int64_t nf, P;
    int64_t frame0;               // global index of local frame 0
    int64_t chunk_len;
    int lo_cnt, hi_first;
    const uint16_t* first_row;    // levels of global frame 0
    const uint16_t* prev_row;     // levels of the frame before local frame 0 (may be null)
    // per-frame outputs, accumulated with atomics (zeroed by the caller-side memset)
    int* native;
    int* nonnative;
    unsigned long long* ssd_prev;
    unsigned long long* ssd_first;
    Records rec;
};

__device__ __forceinline__ uint2 ld_stream(const uint16_t* p) {
    uint2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}

__device__ __forceinline__ unsigned warp_sum_u32_wide(unsigned v, unsigned* hi) {
    // exact 32-lane sum of values < 2^32 returned as lo + (hi << 16)
    unsigned lo = __reduce_add_sync(0xffffffffu, v & 0xffffu);
    *hi = __reduce_add_sync(0xffffffffu, v >> 16);
    return lo;
}

__global__ void __launch_bounds__(kThreads, 2) pair_stats_kernel(const StatsParams prm) {
    __shared__ unsigned s_cnt[kFrameBlock];                 // (native << 16) | total on, per frame
    __shared__ unsigned long long s_sp[kFrameBlock];        // sum (k - k_prev)^2
    __shared__ unsigned long long s_s0[kFrameBlock];        // sum (k - k_first)^2

    const int tid = threadIdx.x, lane = tid & 31;
    const int64_t p0 = ((int64_t)blockIdx.x * kThreads + tid) * kVec;
    const int chunk = blockIdx.y;
    const int64_t fbeg = (int64_t)chunk * prm.chunk_len;
    const int64_t fend = min(prm.nf, fbeg + prm.chunk_len);
    const bool in_range = p0 < prm.stride;   // the row pitch is a multiple of kVec: loads stay inside the row

    if (tid < kFrameBlock) { s_cnt[tid] = 0; s_sp[tid] = 0; s_s0[tid] = 0; }
    __syncthreads();

    unsigned valid = 0;   // bit q set: pair p0+q exists
#pragma unroll
    for (int q = 0; q < kVec; q++) valid |= (p0 + q < prm.P) ? (1u << q) : 0u;
    // columns between P and the row pitch hold no data: read them as level 0 everywhere
    const unsigned mask_x = ((valid & 1u) ? 0xffffu : 0u) | ((valid & 2u) ? 0xffff0000u : 0u);
    const unsigned mask_y = ((valid & 4u) ? 0xffffu : 0u) | ((valid & 8u) ? 0xffff0000u : 0u);

    // reference rows
    unsigned k0[kVec], kp[kVec];
    {
        uint2 w = make_uint2(0, 0);
        if (in_range) w = ld_stream(prm.first_row + p0);
        w.x &= mask_x; w.y &= mask_y;
        k0[0] = w.x & 0xffffu; k0[1] = w.x >> 16; k0[2] = w.y & 0xffffu; k0[3] = w.y >> 16;
        const uint16_t* prow = fbeg > 0 ? prm.levels + (fbeg - 1) * prm.stride : prm.prev_row;
        uint2 v = make_uint2(0, 0);
        if (in_range && prow) v = ld_stream(prow + p0);
        v.x &= mask_x; v.y &= mask_y;
        kp[0] = v.x & 0xffffu; kp[1] = v.x >> 16; kp[2] = v.y & 0xffffu; kp[3] = v.y >> 16;
    }
    const bool have_prev = fbeg > 0 || prm.prev_row != nullptr;
    unsigned native_bits = 0;
#pragma unroll
    for (int q = 0; q < kVec; q++) native_bits |= ((int)k0[q] < prm.lo_cnt) ? (1u << q) : 0u;
    native_bits &= valid;

    unsigned long long sk[kVec], sk2[kVec];
    unsigned a1[kVec], a2[kVec];        // 32-bit running parts, flushed every kFrameBlock frames
    int n_int[kVec], n_enc[kVec], first_f[kVec], last_f[kVec], prefix[kVec];
    unsigned kmn[kVec], kmx[kVec];
    unsigned status = 0, decided = 0;   // bit per pair
#pragma unroll
    for (int q = 0; q < kVec; q++) {
        sk[q] = 0; sk2[q] = 0; a1[q] = 0; a2[q] = 0; n_int[q] = 0; n_enc[q] = 0;
        first_f[q] = -1; last_f[q] = -1; prefix[q] = 0; kmn[q] = 0xffffu; kmx[q] = 0;
    }

    const uint16_t* col = prm.levels + p0;
    for (int64_t fb = fbeg; fb < fend; fb += kFrameBlock) {
        const int nblk = (int)min((int64_t)kFrameBlock, fend - fb);
        for (int u0 = 0; u0 < nblk; u0 += kUnroll) {
            uint2 w[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; u++) {
                w[u] = make_uint2(0, 0);
                if (in_range && u0 + u < nblk) w[u] = ld_stream(col + (fb + u0 + u) * prm.stride);
            }
#pragma unroll
            for (int u = 0; u < kUnroll; u++) {
                if (u0 + u < nblk) {
                    const int fl = u0 + u;                      // frame inside the block
                    const int fg = (int)(prm.frame0 + fb + fl); // global frame index
                    const unsigned wx = w[u].x & mask_x, wy = w[u].y & mask_y;
                    unsigned k[kVec] = {wx & 0xffffu, wx >> 16, wy & 0xffffu, wy >> 16};
                    unsigned sp = 0, s0 = 0, onbits = 0;
#pragma unroll
                    for (int q = 0; q < kVec; q++) {
                        const unsigned kk = k[q];
                        a1[q] += kk;
                        a2[q] += kk * kk;
                        kmn[q] = min(kmn[q], kk);
                        kmx[q] = max(kmx[q], kk);
                        const bool lo = (int)kk < prm.lo_cnt;
                        const bool hi = (int)kk >= prm.hi_first;
                        const bool was = (status >> q) & 1u;
                        const bool now = lo || (was && !hi);
                        const bool dec = (decided >> q) & 1u;
                        if (!dec && !(lo || hi)) prefix[q]++;
                        decided |= (lo || hi) ? (1u << q) : 0u;
                        n_int[q] += now ? 1 : 0;
                        n_enc[q] += (now && !was) ? 1 : 0;
                        last_f[q] = now ? fg : last_f[q];
                        first_f[q] = (now && first_f[q] < 0) ? fg : first_f[q];
                        onbits |= now ? (1u << q) : 0u;
                        const int dp = (int)kk - (int)kp[q];
                        const int d0 = (int)kk - (int)k0[q];
                        sp += (unsigned)(dp * dp);
                        s0 += (unsigned)(d0 * d0);
                        kp[q] = kk;
                    }
                    status = onbits;
                    onbits &= valid;
                    // per-frame reductions over the warp's 128 pairs (padding columns read as
                    // level 0 in every row, so their squared differences vanish)
                    unsigned packed = (__popc(onbits & native_bits) << 16) | __popc(onbits);
                    packed = __reduce_add_sync(0xffffffffu, packed);
                    const bool use_prev = have_prev || (fb + fl) > fbeg;
                    unsigned sp_hi, s0_hi;
                    const unsigned sp_lo = warp_sum_u32_wide(use_prev ? sp : 0u, &sp_hi);
                    const unsigned s0_lo = warp_sum_u32_wide(s0, &s0_hi);
                    if (lane == 0) {
                        if (packed) atomicAdd(&s_cnt[fl], packed);
                        const unsigned long long vp = (unsigned long long)sp_lo + ((unsigned long long)sp_hi << 16);
                        const unsigned long long v0 = (unsigned long long)s0_lo + ((unsigned long long)s0_hi << 16);
                        if (vp) atomicAdd(&s_sp[fl], vp);
                        if (v0) atomicAdd(&s_s0[fl], v0);
                    }
                }
            }
        }
        // widen the 32-bit partial sums: with nlevels <= 7744 (checked by the entry point)
        // 32 frames of k^2 stay below 2^31
#pragma unroll
        for (int q = 0; q < kVec; q++) { sk[q] += a1[q]; sk2[q] += a2[q]; a1[q] = 0; a2[q] = 0; }
        __syncthreads();
        if (tid < nblk) {
            const int64_t f = fb + tid;
            const unsigned c = s_cnt[tid];
            if (c) {
                const int nat = (int)(c >> 16), tot = (int)(c & 0xffffu);
                if (nat) atomicAdd(&prm.native[f], nat);
                if (tot - nat) atomicAdd(&prm.nonnative[f], tot - nat);
            }
            if (s_sp[tid]) atomicAdd(&prm.ssd_prev[f], s_sp[tid]);
            if (s_s0[tid]) atomicAdd(&prm.ssd_first[f], s_s0[tid]);
            s_cnt[tid] = 0; s_sp[tid] = 0; s_s0[tid] = 0;
        }
        __syncthreads();
    }

    // chunk record
    const int64_t rbase = (int64_t)chunk * prm.rec.Ppad;
#pragma unroll
    for (int q = 0; q < kVec; q++) {
        const int64_t p = p0 + q;
        if (p < prm.P) {
            prm.rec.sum_k[rbase + p] = sk[q];
            prm.rec.sum_k2[rbase + p] = sk2[q];
            prm.rec.n_int[rbase + p] = n_int[q];
            prm.rec.n_enc[rbase + p] = n_enc[q];
            prm.rec.first_f[rbase + p] = first_f[q];
            prm.rec.last_f[rbase + p] = last_f[q];
            prm.rec.prefix[rbase + p] = prefix[q];
            prm.rec.kmin[rbase + p] = (uint16_t)kmn[q];
            prm.rec.kmax[rbase + p] = (uint16_t)kmx[q];
            prm.rec.flags[rbase + p] = (uint8_t)(((status >> q) & 1u) | (((decided >> q) & 1u) << 1));
        }
    }
}

struct CombineParams {
    Records rec;
    int nchunks;
    int64_t P, nf, frame0, chunk_len;
    int lo_cnt;
    const uint16_t* first_row;
    const uint8_t* status_in;
    conan_pair_stats_out out;
    int* diff_native;      // [nf + 1] difference arrays for the carry corrections
    int* diff_nonnative;
};

// One thread per pair: fold the chunk records in frame order, applying the entering status.
__global__ void combine_kernel(const CombineParams prm) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= prm.P) return;
    unsigned long long sk = 0, sk2 = 0;
    int n_int = 0, n_enc = 0, first_f = -1, last_f = -1;
    unsigned kmn = 0xffffu, kmx = 0;
    int status = prm.status_in ? (prm.status_in[p] ? 1 : 0) : 0;
    int decided_any = 0;
    const bool native = (int)prm.first_row[p] < prm.lo_cnt;
    int* diff = native ? prm.diff_native : prm.diff_nonnative;
    for (int c = 0; c < prm.nchunks; c++) {
        const int64_t i = (int64_t)c * prm.rec.Ppad + p;
        const int64_t cbeg = (int64_t)c * prm.chunk_len;
        const int64_t clen = min(prm.nf, cbeg + prm.chunk_len) - cbeg;
        sk += prm.rec.sum_k[i];
        sk2 += prm.rec.sum_k2[i];
        kmn = min(kmn, (unsigned)prm.rec.kmin[i]);
        kmx = max(kmx, (unsigned)prm.rec.kmax[i]);
        const int fl = prm.rec.flags[i];
        const int dec = (fl >> 1) & 1;
        const int pre = prm.rec.prefix[i];
        int c_int = prm.rec.n_int[i], c_enc = prm.rec.n_enc[i];
        int c_first = prm.rec.first_f[i], c_last = prm.rec.last_f[i];
        int c_out = fl & 1;
        if (status) {
            // entering "on": the buffered prefix stays on, and a first decisive "on" is no new encounter
            const int gbeg = (int)(prm.frame0 + cbeg);
            if (pre > 0) {
                c_int += pre;
                c_last = max(c_last, gbeg + pre - 1);
                atomicAdd(&diff[cbeg], 1);
                atomicAdd(&diff[cbeg + pre], -1);
            }
            if (dec && c_first == gbeg + pre) c_enc -= 1;
            if (pre > 0) c_first = gbeg;
            if (!dec) c_out = 1;
        }
        (void)clen;
        n_int += c_int;
        n_enc += c_enc;
        if (first_f < 0) first_f = c_first;
        last_f = max(last_f, c_last);
        status = c_out;
        decided_any |= dec;
    }
    prm.out.d_sum_k[p] = (int64_t)sk;
    prm.out.d_sum_k2[p] = (int64_t)sk2;
    prm.out.d_n_int[p] = n_int;
    prm.out.d_n_enc[p] = n_enc;
    prm.out.d_first_f[p] = first_f;
    prm.out.d_last_f[p] = last_f;
    prm.out.d_status[p] = (uint8_t)status;
    prm.out.d_kmin[p] = (uint16_t)kmn;
    prm.out.d_kmax[p] = (uint16_t)kmx;
    if (prm.out.d_decided) prm.out.d_decided[p] = (uint8_t)decided_any;
}

// Inclusive scan of the difference arrays added onto the per-frame counts (single CTA).
__global__ void apply_diff_kernel(const int* __restrict__ dn, const int* __restrict__ dnn, int64_t nf,
                                  int* __restrict__ native, int* __restrict__ nonnative) {
    __shared__ int carry[2];
    __shared__ int wsum[2][32];
    if (threadIdx.x == 0) { carry[0] = 0; carry[1] = 0; }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int64_t base = 0; base < nf; base += blockDim.x) {
        const int64_t f = base + threadIdx.x;
        int a = f < nf ? dn[f] : 0, b = f < nf ? dnn[f] : 0;
        for (int o = 1; o < 32; o <<= 1) {
            int ta = __shfl_up_sync(0xffffffffu, a, o), tb = __shfl_up_sync(0xffffffffu, b, o);
            if (lane >= o) { a += ta; b += tb; }
        }
        if (lane == 31) { wsum[0][warp] = a; wsum[1][warp] = b; }
        __syncthreads();
        int oa = carry[0], ob = carry[1];
        for (int w = 0; w < warp; w++) { oa += wsum[0][w]; ob += wsum[1][w]; }
        a += oa; b += ob;
        if (f < nf) { if (a) native[f] += a; if (b) nonnative[f] += b; }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) { carry[0] = a; carry[1] = b; }
        (void)nw;
        __syncthreads();
    }
}

struct Plan {
    int64_t Ppad, chunk_len;
    int nchunks;
    size_t off[10], off_diff_n, off_diff_nn, total;
};

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

bool make_plan(int64_t nf, int64_t P, Plan* pl) {
    if (nf <= 0 || P <= 0) return false;
    const int64_t tiles = (P + kPairsPerCta - 1) / kPairsPerCta;
    int64_t want = (148 * 16 + tiles - 1) / tiles;             // aim for ~16 CTAs per SM in flight / queued
    int64_t maxc = nf / 32 > 0 ? nf / 32 : 1;                   // chunks of at least 32 frames
    int64_t nchunks = want < 1 ? 1 : want;
    if (nchunks > maxc) nchunks = maxc;
    if (nchunks > 65535) nchunks = 65535;
    int64_t clen = (nf + nchunks - 1) / nchunks;
    clen = (clen + kFrameBlock - 1) / kFrameBlock * kFrameBlock;
    nchunks = (nf + clen - 1) / clen;
    pl->nchunks = (int)nchunks;
    pl->chunk_len = clen;
    pl->Ppad = (P + 63) / 64 * 64;
    const size_t n = (size_t)nchunks * pl->Ppad;
    const size_t sizes[10] = {8, 8, 4, 4, 4, 4, 4, 2, 2, 1};
    size_t o = 0;
    for (int i = 0; i < 10; i++) { pl->off[i] = o; o = align256(o + sizes[i] * n); }
    pl->off_diff_n = o;  o = align256(o + sizeof(int) * (size_t)(nf + 1));
    pl->off_diff_nn = o; o = align256(o + sizeof(int) * (size_t)(nf + 1));
    pl->total = o;
    return true;
}

}  // namespace

extern "C" size_t conan_pair_stats_workspace_bytes(int64_t nf, int64_t P) {
    Plan pl;
    if (!make_plan(nf, P, &pl)) return 0;
    return pl.total;
}

extern "C" int conan_pair_stats(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                int64_t frame0, int32_t nlevels, int32_t lo_cnt, int32_t hi_first,
                                int32_t equal_cut,
                                const uint16_t* d_first_row, const uint16_t* d_prev_row,
                                const uint8_t* d_status_in, const conan_pair_stats_out* out,
                                const conan_frame_stats_out* fout, void* d_work, size_t work_bytes,
                                void* stream_) {
    return conan_pair_stats_ex(d_levels, levels_stride, nf, P, frame0, nlevels, lo_cnt, hi_first, equal_cut,
                               d_first_row, d_prev_row, d_status_in, out, fout, d_work, work_bytes, stream_, 0);
}

extern "C" int conan_pair_stats_ex(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                   int64_t frame0, int32_t nlevels, int32_t lo_cnt, int32_t hi_first,
                                   int32_t equal_cut,
                                   const uint16_t* d_first_row, const uint16_t* d_prev_row,
                                   const uint8_t* d_status_in, const conan_pair_stats_out* out,
                                   const conan_frame_stats_out* fout, void* d_work, size_t work_bytes,
                                   void* stream_, int32_t recombine) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_levels || !out || !fout || !d_work)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: null pointer argument");
    if (!out->d_sum_k || !out->d_sum_k2 || !out->d_n_int || !out->d_n_enc || !out->d_first_f ||
        !out->d_last_f || !out->d_status || !out->d_kmin || !out->d_kmax)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: a per-pair output pointer is null");
    if (!fout->d_native || !fout->d_nonnative || !fout->d_ssd_prev || !fout->d_ssd_first)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: a per-frame output pointer is null");
    if (nf <= 0 || P <= 0) return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: nf=%lld P=%lld", (long long)nf, (long long)P);
    if (levels_stride < P || levels_stride % kVec != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: levels_stride=%lld must be >= P and a multiple of %d",
                           (long long)levels_stride, kVec);
    if ((reinterpret_cast<uintptr_t>(d_levels) & 7) != 0 ||
        (d_first_row && (reinterpret_cast<uintptr_t>(d_first_row) & 7) != 0) ||
        (d_prev_row && (reinterpret_cast<uintptr_t>(d_prev_row) & 7) != 0))
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: level rows must be 8-byte aligned");
    if (frame0 < 0 || frame0 + nf > 2147483647ll)
        return conan::fail(CONAN_ERR_SIZE, "conan_pair_stats: frame indices must fit in int32");
    if (nlevels < 2 || nlevels > 7744)
        return conan::fail(CONAN_ERR_SIZE, "conan_pair_stats: nlevels=%d outside [2, 7744] (the xpm two-character limit)", nlevels);
    if (lo_cnt < 0 || hi_first < 0) return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: negative threshold");
    if (equal_cut) hi_first = lo_cnt;
    if (hi_first < lo_cnt)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: hi_first=%d < lo_cnt=%d (TRUNC_INTER_HIGH < TRUNC_INTER)", hi_first, lo_cnt);
    Plan pl;
    make_plan(nf, P, &pl);
    if (work_bytes < pl.total)
        return conan::fail(CONAN_ERR_SIZE, "conan_pair_stats: workspace %zu bytes, need %zu", work_bytes, pl.total);
    if ((reinterpret_cast<uintptr_t>(d_work) & 255) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: workspace must be 256-byte aligned");

    char* w = static_cast<char*>(d_work);
    Records rec;
    rec.sum_k = reinterpret_cast<unsigned long long*>(w + pl.off[0]);
    rec.sum_k2 = reinterpret_cast<unsigned long long*>(w + pl.off[1]);
    rec.n_int = reinterpret_cast<int*>(w + pl.off[2]);
    rec.n_enc = reinterpret_cast<int*>(w + pl.off[3]);
    rec.first_f = reinterpret_cast<int*>(w + pl.off[4]);
    rec.last_f = reinterpret_cast<int*>(w + pl.off[5]);
    rec.prefix = reinterpret_cast<int*>(w + pl.off[6]);
    rec.kmin = reinterpret_cast<uint16_t*>(w + pl.off[7]);
    rec.kmax = reinterpret_cast<uint16_t*>(w + pl.off[8]);
    rec.flags = reinterpret_cast<uint8_t*>(w + pl.off[9]);
    rec.Ppad = pl.Ppad;
    int* diff_n = reinterpret_cast<int*>(w + pl.off_diff_n);
    int* diff_nn = reinterpret_cast<int*>(w + pl.off_diff_nn);

    if (!recombine) {
        CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_native, 0, sizeof(int) * (size_t)nf, stream));
        CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_nonnative, 0, sizeof(int) * (size_t)nf, stream));
        CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_ssd_prev, 0, sizeof(int64_t) * (size_t)nf, stream));
        CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_ssd_first, 0, sizeof(int64_t) * (size_t)nf, stream));
    }
    CONAN_CUDA_TRY(cudaMemsetAsync(diff_n, 0, sizeof(int) * (size_t)(nf + 1), stream));
    CONAN_CUDA_TRY(cudaMemsetAsync(diff_nn, 0, sizeof(int) * (size_t)(nf + 1), stream));

    StatsParams sp;
    sp.levels = d_levels; sp.stride = levels_stride; sp.nf = nf; sp.P = P; sp.frame0 = frame0;
    sp.chunk_len = pl.chunk_len; sp.lo_cnt = lo_cnt; sp.hi_first = hi_first;
    sp.first_row = d_first_row ? d_first_row : d_levels;
    sp.prev_row = d_prev_row;
    sp.native = fout->d_native; sp.nonnative = fout->d_nonnative;
    sp.ssd_prev = reinterpret_cast<unsigned long long*>(fout->d_ssd_prev);
    sp.ssd_first = reinterpret_cast<unsigned long long*>(fout->d_ssd_first);
    sp.rec = rec;
    const int64_t tiles = (P + kPairsPerCta - 1) / kPairsPerCta;
    if (tiles > 2147483647ll) return conan::fail(CONAN_ERR_SIZE, "conan_pair_stats: too many pairs");
    if (!recombine) {
        dim3 grid((unsigned)tiles, (unsigned)pl.nchunks);
        pair_stats_kernel<<<grid, kThreads, 0, stream>>>(sp);
        CONAN_CUDA_TRY(cudaGetLastError());
    }

    CombineParams cp;
    cp.rec = rec; cp.nchunks = pl.nchunks; cp.P = P; cp.nf = nf; cp.frame0 = frame0; cp.chunk_len = pl.chunk_len;
    cp.lo_cnt = lo_cnt; cp.first_row = sp.first_row; cp.status_in = d_status_in; cp.out = *out;
    cp.diff_native = diff_n; cp.diff_nonnative = diff_nn;
    combine_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(cp);
    CONAN_CUDA_TRY(cudaGetLastError());
    apply_diff_kernel<<<1, 1024, 0, stream>>>(diff_n, diff_nn, nf, fout->d_native, fout->d_nonnative);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
