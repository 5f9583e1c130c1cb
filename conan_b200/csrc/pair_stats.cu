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
// series in order (which the hysteresis needs).  The grid is (pair tiles) x (frame chunks).
//   entering_status_kernel  for every chunk boundary, scans backwards from the boundary to the
//                           most recent frame outside the hysteresis buffer; that frame alone
//                           decides the status entering the chunk (a handful of rows per boundary)
//   pair_stats_kernel       the streaming pass; each thread owns four adjacent pairs (one 8-byte
//                           load per frame, eight loads in flight).  The kernel is bound by
//                           instruction issue, not by HBM, so the loop is written for instruction
//                           count: cut-off tests on two packed levels at once (one add each, the
//                           verdict lands in bits 15 / 31), hysteresis as one LOP3, the on/off
//                           history shifted into a 32-frame bit mask per pair from which contact
//                           frames, encounters, first and last frame are taken with popc / clz /
//                           ffs once per 32 frames, per-frame sums reduced with REDUX per warp.
//   combine_kernel          adds the chunk records of each pair.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "common.h"

namespace {

constexpr int kThreads = 256;
constexpr int kVec = 4;                       // pairs per thread (one 8-byte load per frame)
constexpr int kPairsPerCta = kThreads * kVec;
constexpr int kWin = 32;                      // frames per history window / per-frame flush block
constexpr int kStageFrames = 8;               // frames per shared-memory stage
constexpr int kStagesS = 4;                   // stages in flight per CTA (4 x 8 frames x 2 KB = 64 KB)
constexpr int kConsumers = kThreads;          // 8 consumer warps; one more warp feeds the stages
constexpr int kSmallLevels = 2049;            // up to here a warp's squared-difference sum fits 32 bits

// One record per (chunk, pair), struct-of-arrays in the workspace.
struct Records {
    unsigned long long* sum_k;    // [nchunks][Ppad]
    unsigned long long* sum_k2;
    int* n_int;
    int* n_enc;
    int* first_f;                 // global frame index, -1 = never on in the chunk
    int* last_f;
    uint8_t* status_out;          // status after the chunk's last frame
    uint8_t* enter;               // [nchunks + 1][Ppad] status entering chunk c; row nchunks = after the last frame
    int64_t Ppad;
};

struct StatsParams {
    const uint16_t* levels;
    int64_t stride;               // elements between frames
    int64_t nf, P;
    int64_t frame0;               // global index of local frame 0
    int64_t chunk_len;
    int nchunks;
    int lo_cnt, hi_first;
    int one;                      // == 1, opaque to the compiler: k*one - kp becomes one IMAD on the FMA pipe
    unsigned quiet_word;          // QUIET kernels: two packed copies of the last level
    const uint16_t* first_row;    // levels of global frame 0
    const uint16_t* prev_row;     // levels of the frame before local frame 0 (may be null)
    const uint8_t* status_in;     // status entering local frame 0 (may be null = off)
    uint8_t* decided;             // optional output: some frame of the call was outside the buffer
    // per-frame outputs, accumulated with atomics (zeroed by the entry point)
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

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spin = 0; spin < (1u << 26); spin++) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"     // suspends up to the hinted time
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(0x989680u) : "memory");
        if (ok) return;
    }
    __trap();   // a copy that never lands must not hang the GPU
}
// One frame row of the tile, global -> shared, through the bulk-copy engine (no registers involved).
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------------------------------
// status entering every chunk
// ------------------------------------------------------------------------------------------
__global__ void entering_status_kernel(const StatsParams prm) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;   // boundary: the status entering chunk b (b = nchunks: after the last frame)
    const bool live = p < prm.P;
    int64_t t = min(prm.nf, (int64_t)b * prm.chunk_len) - 1;
    int status = live ? -1 : 0;
    // all lanes of a warp walk back together, four rows per step (independent loads), until each has
    // met a frame outside the buffer; the most recent such frame decides
    const uint16_t* col = prm.levels + (live ? p : 0);
    while (true) {
        int k[4];
#pragma unroll
        for (int u = 0; u < 4; u++) k[u] = (status < 0 && t - u >= 0) ? (int)col[(t - u) * prm.stride] : -1;
#pragma unroll
        for (int u = 0; u < 4; u++)
            if (status < 0 && k[u] >= 0) {
                if (k[u] < prm.lo_cnt) status = 1;
                else if (k[u] >= prm.hi_first) status = 0;
            }
        t -= 4;
        if (__all_sync(0xffffffffu, status >= 0 || t < 0)) break;
    }
    if (!live) return;
    const bool dec = status >= 0;
    if (!dec) status = prm.status_in ? (prm.status_in[p] ? 1 : 0) : 0;
    prm.rec.enter[(int64_t)b * prm.rec.Ppad + p] = (uint8_t)status;
    if (b == prm.nchunks && prm.decided) prm.decided[p] = dec ? 1 : 0;
}

// ------------------------------------------------------------------------------------------
// streaming pass
// ------------------------------------------------------------------------------------------
// Close a history window of n frames for one pair.  h holds the window's on/off bits, the oldest
// frame in bit n-1 and the newest in bit 0; `carry` is the status before the window's first frame.
__device__ __forceinline__ void close_window(unsigned h, int n, unsigned carry, int wbase, int& n_int, int& n_enc,
                                             int& first_f, int& last_f) {
    if (h == 0) return;
    n_int += __popc(h);
    const unsigned prev = (h >> 1) | (carry << (n - 1));
    n_enc += __popc(h & ~prev);
    last_f = wbase + (n - __ffs(h));                    // newest on-frame = lowest set bit
    if (first_f < 0) first_f = wbase + (n - 32 + __clz(h));   // oldest on-frame = highest set bit
}

// What one warp contributes to one frame: contact counts and the two squared-difference sums.
struct FramePart {
    unsigned cnt;        // (native << 16) | total pairs on
    unsigned sp_lo, s0_lo;
    unsigned sp_hi;      // WIDE only: carry word of the split ssd_prev sum (its ssd_first twin sits in s_hi0:
                         // two 16-bit fields in one word overflow above ~5800 levels)
};

// One frame for the thread's four pairs.  `wx`, `wy` are the two packed level words.
template <bool WIDE>
__device__ __forceinline__ void frame_step(unsigned wx, unsigned wy, unsigned c_lo, unsigned c_hi, unsigned valid_y,
                                           unsigned native_y, unsigned& on_x, unsigned& on_y, unsigned (&hist)[kVec],
                                           unsigned (&a1)[kVec], unsigned (&a2)[kVec], unsigned (&kp)[kVec],
                                           const unsigned (&k0)[kVec], FramePart* slot, unsigned* slot_hi0, bool lane0,
                                           int one) {
    // hysteresis for four pairs: on = below_lo | (on & ~above_hi), verdicts in bits 15 / 31
    const unsigned ge_lo_x = wx + c_lo, ge_hi_x = wx + c_hi;
    const unsigned ge_lo_y = wy + c_lo, ge_hi_y = wy + c_hi;
    on_x = ~ge_lo_x | (on_x & ~ge_hi_x);
    on_y = ~ge_lo_y | (on_y & ~ge_hi_y);
    hist[0] = __funnelshift_l(on_x << 16, hist[0], 1);
    hist[1] = __funnelshift_l(on_x, hist[1], 1);
    hist[2] = __funnelshift_l(on_y << 16, hist[2], 1);
    hist[3] = __funnelshift_l(on_y, hist[3], 1);
    const unsigned k[kVec] = {wx & 0xffffu, wx >> 16, wy & 0xffffu, wy >> 16};
    unsigned sp = 0, s0 = 0;
#pragma unroll
    for (int q = 0; q < kVec; q++) {
        a1[q] += k[q];
        a2[q] += k[q] * k[q];
        const int dp = (int)k[q] * one - (int)kp[q];     // IMAD: keeps the integer ALU free for the bit work
        const int d0 = (int)k[q] * one - (int)k0[q];
        sp += (unsigned)(dp * dp);
        s0 += (unsigned)(d0 * d0);
        kp[q] = k[q];
    }
    // per-frame sums over the warp's 128 pairs, one REDUX each, stored by lane 0 without atomics
    const unsigned flags = __byte_perm(on_x, on_y, 0x7531) & valid_y;
    FramePart part;
    part.cnt = __reduce_add_sync(0xffffffffu, (__popc(flags & native_y) << 16) + __popc(flags));
    unsigned s0_hi = 0;
    if (!WIDE) {
        part.sp_lo = __reduce_add_sync(0xffffffffu, sp);
        part.s0_lo = __reduce_add_sync(0xffffffffu, s0);
        part.sp_hi = 0;
    } else {   // 128 squares of up to 7743^2 do not fit 32 bits: 16-bit halves, every sum in a word of its own
        part.sp_lo = __reduce_add_sync(0xffffffffu, sp & 0xffffu);
        part.s0_lo = __reduce_add_sync(0xffffffffu, s0 & 0xffffu);
        part.sp_hi = __reduce_add_sync(0xffffffffu, sp >> 16);
        s0_hi = __reduce_add_sync(0xffffffffu, s0 >> 16);
    }
    if (lane0) {
        *reinterpret_cast<uint4*>(slot) = make_uint4(part.cnt, part.sp_lo, part.s0_lo, part.sp_hi);
        if (WIDE) *slot_hi0 = s0_hi;
    }
}

// QUIET: stages (8 frames) in which all of the warp's 128 pairs sit at the last level, as they did in the frame
// before, are only counted -- nobody is on, the differences to the previous frame are zero, the one to the first
// frame is the same every time -- and the history masks and level sums catch up at the next busy stage or when the
// window closes.  Pays on sparse, block-structured stacks (an X x Y interface, multi-domain proteins); on a stack
// whose every 128-pair stretch holds something the test is pure overhead, hence a separate instantiation.
template <bool MASKED, bool WIDE, bool QUIET>
__device__ __forceinline__ void stats_body(const StatsParams& prm, FramePart (*s_part)[kThreads / 32][kWin],
                                           unsigned (*s_hi0)[kThreads / 32][kWin], const uint16_t* s_rows,
                                           uint64_t* full_bar, uint64_t* empty_bar) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool lane0 = lane == 0;
    const int64_t p0 = ((int64_t)blockIdx.x * kThreads + tid) * kVec;
    const int chunk = blockIdx.y;
    const int64_t fbeg = (int64_t)chunk * prm.chunk_len;
    const int64_t fend = min(prm.nf, fbeg + prm.chunk_len);
    const bool in_range = p0 < prm.stride;   // the row pitch is a multiple of kVec: loads stay inside the row

    unsigned mask_x = 0xffffffffu, mask_y = 0xffffffffu, valid_y = 0x80808080u;
    if (MASKED) {   // columns between P and the row pitch hold no data: read them as level 0 everywhere
        mask_x = ((p0 + 0 < prm.P) ? 0xffffu : 0u) | ((p0 + 1 < prm.P) ? 0xffff0000u : 0u);
        mask_y = ((p0 + 2 < prm.P) ? 0xffffu : 0u) | ((p0 + 3 < prm.P) ? 0xffff0000u : 0u);
        // the same validity in the byte-gathered layout of the per-frame counts (bit 7 of each byte)
        valid_y = ((p0 + 0 < prm.P) ? 0x80u : 0u) | ((p0 + 1 < prm.P) ? 0x8000u : 0u) |
                  ((p0 + 2 < prm.P) ? 0x800000u : 0u) | ((p0 + 3 < prm.P) ? 0x80000000u : 0u);
    }

    // reference rows: global frame 0 and the frame before this chunk (the chunk's own first frame
    // when there is none, which makes the first squared difference vanish)
    unsigned k0[kVec], kp[kVec];
    {
        uint2 w = ld_stream(prm.first_row + (in_range ? p0 : 0));
        w.x &= mask_x; w.y &= mask_y;
        k0[0] = w.x & 0xffffu; k0[1] = w.x >> 16; k0[2] = w.y & 0xffffu; k0[3] = w.y >> 16;
        const uint16_t* prow = fbeg > 0 ? prm.levels + (fbeg - 1) * prm.stride
                                        : (prm.prev_row ? prm.prev_row : prm.levels);
        uint2 v = ld_stream(prow + (in_range ? p0 : 0));
        v.x &= mask_x; v.y &= mask_y;
        kp[0] = v.x & 0xffffu; kp[1] = v.x >> 16; kp[2] = v.y & 0xffffu; kp[3] = v.y >> 16;
    }
    unsigned native_y = 0;
#pragma unroll
    for (int q = 0; q < kVec; q++) native_y |= ((int)k0[q] < prm.lo_cnt) ? (0x80u << (8 * q)) : 0u;
    native_y &= valid_y;

    // cut-off tests on packed levels: half + (0x8000 - cut) has bit 15 set iff level >= cut
    const unsigned c_lo = (unsigned)(0x8000 - prm.lo_cnt) * 0x00010001u;
    const unsigned c_hi = (unsigned)(0x8000 - prm.hi_first) * 0x00010001u;

    // entering status -> flag bits 15 / 31 of the two status words
    unsigned carry = 0;   // bit q = status of pair q before the current window
    {
        const uint8_t* e = prm.rec.enter + (int64_t)chunk * prm.rec.Ppad + p0;
#pragma unroll
        for (int q = 0; q < kVec; q++)
            if (p0 + q < prm.P) carry |= (e[q] ? 1u : 0u) << q;
    }
    unsigned on_x = ((carry & 1u) ? 0x8000u : 0u) | ((carry & 2u) ? 0x80000000u : 0u);
    unsigned on_y = ((carry & 4u) ? 0x8000u : 0u) | ((carry & 8u) ? 0x80000000u : 0u);

    unsigned long long sk[kVec], sk2[kVec];
    unsigned a1[kVec], a2[kVec];        // 32-bit running sums, widened once per window
    unsigned hist[kVec];
    int n_int[kVec], n_enc[kVec], first_f[kVec], last_f[kVec];
#pragma unroll
    for (int q = 0; q < kVec; q++) {
        sk[q] = 0; sk2[q] = 0; a1[q] = 0; a2[q] = 0; hist[q] = 0;
        n_int[q] = 0; n_enc[q] = 0; first_f[q] = -1; last_f[q] = -1;
    }

    // quiet stages: what a frame of nothing but last levels adds to the per-frame sums of this warp
    const unsigned qw = prm.quiet_word, ql = qw & 0xffffu;
    uint4 quiet_part = make_uint4(0, 0, 0, 0);
    unsigned quiet_hi0 = 0;
    int nq = 0, nq_win = 0;
    if (QUIET && !MASKED) {
        unsigned s0 = 0;
#pragma unroll
        for (int q = 0; q < kVec; q++) { const int d0 = (int)ql - (int)k0[q]; s0 += (unsigned)(d0 * d0); }
        if (!WIDE) {
            quiet_part.z = __reduce_add_sync(0xffffffffu, s0);
        } else {
            quiet_part.z = __reduce_add_sync(0xffffffffu, s0 & 0xffffu);
            quiet_hi0 = __reduce_add_sync(0xffffffffu, s0 >> 16);
        }
    }

    // this thread's four levels of frame r of stage slot: 8 bytes at a fixed offset
    const uint16_t* my = s_rows + tid * kVec;
    int it = 0, buf = 0;      // it: stage iteration since the start of the chunk
    for (int64_t fb = fbeg; fb < fend; fb += kWin, buf ^= 1) {
        const int nblk = (int)min((int64_t)kWin, fend - fb);
        FramePart* slots = &s_part[buf][warp][0];
        unsigned* slots_hi0 = &s_hi0[buf][warp][0];
#pragma unroll 1
        for (int u0 = 0; u0 < nblk; u0 += kStageFrames, it++) {
            const int slot = it % kStagesS;
            mbar_wait(&full_bar[slot], (uint32_t)(it / kStagesS) & 1u);
            const uint16_t* rows = my + (size_t)slot * kStageFrames * kPairsPerCta;
            if (nblk - u0 >= kStageFrames) {
                bool all_quiet = false;
                if (QUIET && !MASKED) {
                    unsigned diff = (on_x | on_y) & 0x80008000u;
#pragma unroll
                    for (int q = 0; q < kVec; q++) diff |= kp[q] ^ ql;
#pragma unroll
                    for (int u = 0; u < kStageFrames; u++) {
                        const uint2 w = *reinterpret_cast<const uint2*>(rows + u * kPairsPerCta);
                        diff |= (w.x ^ qw) | (w.y ^ qw);
                    }
                    all_quiet = __all_sync(0xffffffffu, diff == 0);
                }
                if (QUIET && all_quiet) {
                    nq += kStageFrames;
                    if (lane0) {
#pragma unroll
                        for (int u = 0; u < kStageFrames; u++) {
                            *reinterpret_cast<uint4*>(slots + u0 + u) = quiet_part;
                            if (WIDE) slots_hi0[u0 + u] = quiet_hi0;
                        }
                    }
                } else {
                    if (QUIET && nq) {
#pragma unroll
                        for (int q = 0; q < kVec; q++) hist[q] = nq < 32 ? hist[q] << nq : 0u;
                        nq_win += nq;
                        nq = 0;
                    }
#pragma unroll
                    for (int u = 0; u < kStageFrames; u++) {
                        uint2 w = *reinterpret_cast<const uint2*>(rows + u * kPairsPerCta);
                        if (MASKED) { w.x &= mask_x; w.y &= mask_y; }
                        frame_step<WIDE>(w.x, w.y, c_lo, c_hi, valid_y, native_y, on_x, on_y, hist, a1, a2, kp, k0,
                                         slots + u0 + u, slots_hi0 + u0 + u, lane0, prm.one);
                    }
                }
            } else {   // the last, partial stage of the call
                if (QUIET && nq) {
#pragma unroll
                    for (int q = 0; q < kVec; q++) hist[q] = nq < 32 ? hist[q] << nq : 0u;
                    nq_win += nq;
                    nq = 0;
                }
                for (int u = 0; u < nblk - u0; u++) {
                    uint2 w = *reinterpret_cast<const uint2*>(rows + u * kPairsPerCta);
                    if (MASKED) { w.x &= mask_x; w.y &= mask_y; }
                    frame_step<WIDE>(w.x, w.y, c_lo, c_hi, valid_y, native_y, on_x, on_y, hist, a1, a2, kp, k0,
                                     slots + u0 + u, slots_hi0 + u0 + u, lane0, prm.one);
                }
            }
            __syncwarp();
            if (lane0) mbar_arrive(&empty_bar[slot]);      // this warp is done with the stage
        }
        // close the window: counts from the history masks, widen the 32-bit sums (nlevels <= 7744
        // keeps 32 frames of k^2 below 2^31)
        const int wbase = (int)(prm.frame0 + fb);
        if (QUIET) {
            if (nq) {
#pragma unroll
                for (int q = 0; q < kVec; q++) hist[q] = nq < 32 ? hist[q] << nq : 0u;
                nq_win += nq;
                nq = 0;
            }
            if (nq_win) {     // the quiet stages' share of the level sums
#pragma unroll
                for (int q = 0; q < kVec; q++) { a1[q] += (unsigned)nq_win * ql; a2[q] += (unsigned)nq_win * ql * ql; }
                nq_win = 0;
            }
        }
#pragma unroll
        for (int q = 0; q < kVec; q++) {
            close_window(hist[q], nblk, (carry >> q) & 1u, wbase, n_int[q], n_enc[q], first_f[q], last_f[q]);
            sk[q] += a1[q]; sk2[q] += a2[q]; a1[q] = 0; a2[q] = 0;
        }
        carry = (hist[0] & 1u) | ((hist[1] & 1u) << 1) | ((hist[2] & 1u) << 2) | ((hist[3] & 1u) << 3);
#pragma unroll
        for (int q = 0; q < kVec; q++) hist[q] = 0;
        // one barrier per window (consumer warps only): the buffers alternate, so this window's readers
        // are done before the buffer is written again two windows later
        asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory");
        if (tid < nblk) {
            unsigned cnt = 0;
            unsigned long long vp = 0, v0 = 0;
#pragma unroll
            for (int w8 = 0; w8 < kThreads / 32; w8++) {
                const FramePart fp = s_part[buf][w8][tid];
                cnt += fp.cnt;
                vp += (unsigned long long)fp.sp_lo + ((unsigned long long)fp.sp_hi << 16);
                v0 += (unsigned long long)fp.s0_lo;
                if (WIDE) v0 += (unsigned long long)s_hi0[buf][w8][tid] << 16;
            }
            const int64_t f = fb + tid;
            const int nat = (int)(cnt >> 16), tot = (int)(cnt & 0xffffu);
            if (nat) atomicAdd(&prm.native[f], nat);
            if (tot - nat) atomicAdd(&prm.nonnative[f], tot - nat);
            if (vp) atomicAdd(&prm.ssd_prev[f], vp);
            if (v0) atomicAdd(&prm.ssd_first[f], v0);
        }
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
            prm.rec.status_out[rbase + p] = (uint8_t)((carry >> q) & 1u);
        }
    }
}

template <bool WIDE, bool QUIET>
__global__ void __launch_bounds__(kThreads + 32, 2) pair_stats_kernel(const StatsParams prm) {
    extern __shared__ __align__(128) unsigned char s_dyn[];    // [kStagesS][kStageFrames][1024] levels
    __shared__ __align__(16) FramePart s_part[2][kThreads / 32][kWin];   // per-warp, per-frame partial sums
    __shared__ unsigned s_hi0[2][kThreads / 32][kWin];                   // WIDE only: carry words of the ssd_first sums
    __shared__ __align__(8) uint64_t full_bar[kStagesS], empty_bar[kStagesS];
    uint16_t* s_rows = reinterpret_cast<uint16_t*>(s_dyn);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kStagesS; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], kThreads / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x >= kConsumers) {
        // ---- producer warp: one lane keeps kStagesS stages of kStageFrames rows in flight ----
        if (threadIdx.x == kConsumers) {
            const int64_t tile_p0 = (int64_t)blockIdx.x * kPairsPerCta;
            const int64_t fbeg = (int64_t)blockIdx.y * prm.chunk_len;
            const int64_t fend = min(prm.nf, fbeg + prm.chunk_len);
            const uint32_t row_bytes = (uint32_t)min((int64_t)kPairsPerCta, prm.stride - tile_p0) * 2u;
            const uint16_t* src = prm.levels + tile_p0;
            int it = 0;
            for (int64_t f = fbeg; f < fend; f += kStageFrames, it++) {
                const int slot = it % kStagesS;
                mbar_wait(&empty_bar[slot], ((uint32_t)(it / kStagesS) & 1u) ^ 1u);
                const int nrows = (int)min((int64_t)kStageFrames, fend - f);
                mbar_expect_tx(&full_bar[slot], row_bytes * (uint32_t)nrows);
                uint16_t* dst = s_rows + (size_t)slot * kStageFrames * kPairsPerCta;
                for (int r = 0; r < nrows; r++)
                    bulk_load(dst + (size_t)r * kPairsPerCta, src + (f + r) * prm.stride, row_bytes, &full_bar[slot]);
            }
        }
        return;
    }
    // only the last pair tile can contain padding columns
    const bool tail = ((int64_t)(blockIdx.x + 1) * kPairsPerCta > prm.P);
    if (tail) stats_body<true, WIDE, QUIET>(prm, s_part, s_hi0, s_rows, full_bar, empty_bar);
    else stats_body<false, WIDE, QUIET>(prm, s_part, s_hi0, s_rows, full_bar, empty_bar);
}

struct CombineParams {
    Records rec;
    int nchunks;
    int64_t P;
    conan_pair_stats_out out;
};

// One thread per pair: add up the chunk records.
__global__ void combine_kernel(const CombineParams prm) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= prm.P) return;
    unsigned long long sk = 0, sk2 = 0;
    int n_int = 0, n_enc = 0, first_f = -1, last_f = -1, status = 0;
    for (int c = 0; c < prm.nchunks; c++) {
        const int64_t i = (int64_t)c * prm.rec.Ppad + p;
        sk += prm.rec.sum_k[i];
        sk2 += prm.rec.sum_k2[i];
        n_int += prm.rec.n_int[i];
        n_enc += prm.rec.n_enc[i];
        const int cf = prm.rec.first_f[i];
        if (first_f < 0) first_f = cf;
        last_f = max(last_f, prm.rec.last_f[i]);
        status = prm.rec.status_out[i];
    }
    prm.out.d_sum_k[p] = (int64_t)sk;
    prm.out.d_sum_k2[p] = (int64_t)sk2;
    prm.out.d_n_int[p] = n_int;
    prm.out.d_n_enc[p] = n_enc;
    prm.out.d_first_f[p] = first_f;
    prm.out.d_last_f[p] = last_f;
    prm.out.d_status[p] = (uint8_t)status;
}

struct Plan {
    int64_t Ppad, chunk_len;
    int nchunks;
    size_t off[8], total;
};

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

bool make_plan(int64_t nf, int64_t P, Plan* pl) {
    if (nf <= 0 || P <= 0) return false;
    const int64_t tiles = (P + kPairsPerCta - 1) / kPairsPerCta;
    int64_t want = (148 * 32 + tiles - 1) / tiles;             // aim for ~32 CTAs per SM over the launch
    int64_t maxc = nf / 64 > 0 ? nf / 64 : 1;                   // chunks of at least 64 frames
    int64_t nchunks = want < 1 ? 1 : want;
    if (nchunks > maxc) nchunks = maxc;
    if (nchunks > 65534) nchunks = 65534;
    int64_t clen = (nf + nchunks - 1) / nchunks;
    clen = (clen + kWin - 1) / kWin * kWin;
    nchunks = (nf + clen - 1) / clen;
    pl->nchunks = (int)nchunks;
    pl->chunk_len = clen;
    pl->Ppad = (P + 63) / 64 * 64;
    const size_t n = (size_t)nchunks * pl->Ppad;
    const size_t sizes[7] = {8, 8, 4, 4, 4, 4, 1};
    size_t o = 0;
    for (int i = 0; i < 7; i++) { pl->off[i] = o; o = align256(o + sizes[i] * n); }
    pl->off[7] = o; o = align256(o + (size_t)(nchunks + 1) * pl->Ppad);
    pl->total = o;
    return true;
}

int check_common(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P, int32_t nlevels,
                 int32_t lo_cnt, int32_t* hi_first, int32_t equal_cut, const char* who) {
    if (!d_levels) return conan::fail(CONAN_ERR_ARG, "%s: null level stack", who);
    if (nf <= 0 || P <= 0) return conan::fail(CONAN_ERR_ARG, "%s: nf=%lld P=%lld", who, (long long)nf, (long long)P);
    // rows are streamed with cp.async.bulk: 16-byte aligned addresses and byte counts
    if (levels_stride < P || levels_stride % 8 != 0)
        return conan::fail(CONAN_ERR_ARG, "%s: levels_stride=%lld must be >= P and a multiple of 8", who,
                           (long long)levels_stride);
    if ((reinterpret_cast<uintptr_t>(d_levels) & 15) != 0)
        return conan::fail(CONAN_ERR_ARG, "%s: level rows must be 16-byte aligned", who);
    if (nlevels < 2 || nlevels > 7744)
        return conan::fail(CONAN_ERR_SIZE, "%s: nlevels=%d outside [2, 7744] (the xpm two-character limit)", who, nlevels);
    if (lo_cnt < 0 || *hi_first < 0 || lo_cnt > nlevels || *hi_first > nlevels)
        return conan::fail(CONAN_ERR_ARG, "%s: cut-offs must lie in [0, nlevels]", who);
    if (equal_cut) *hi_first = lo_cnt;
    if (*hi_first < lo_cnt)
        return conan::fail(CONAN_ERR_ARG, "%s: hi_first=%d < lo_cnt=%d (TRUNC_INTER_HIGH < TRUNC_INTER)", who, *hi_first, lo_cnt);
    return CONAN_OK;
}

}  // namespace

extern "C" size_t conan_pair_stats_workspace_bytes(int64_t nf, int64_t P) {
    Plan pl;
    if (!make_plan(nf, P, &pl)) return 0;
    return pl.total;
}

extern "C" int conan_pair_stats(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                int64_t frame0, int32_t nlevels, int32_t lo_cnt, int32_t hi_first,
                                int32_t equal_cut, const uint16_t* d_first_row, const uint16_t* d_prev_row,
                                const uint8_t* d_status_in, const conan_pair_stats_out* out,
                                const conan_frame_stats_out* fout, void* d_work, size_t work_bytes,
                                void* stream_) {
    const char* e = getenv("CONAN_STATS_SPARSE");
    return conan_pair_stats_ex(d_levels, levels_stride, nf, P, frame0, nlevels, lo_cnt, hi_first, equal_cut, d_first_row,
                               d_prev_row, d_status_in, out, fout, d_work, work_bytes, (e && atoi(e) != 0) ? CONAN_STATS_SPARSE : 0,
                               stream_);
}

extern "C" int conan_pair_stats_ex(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                   int64_t frame0, int32_t nlevels, int32_t lo_cnt, int32_t hi_first,
                                   int32_t equal_cut, const uint16_t* d_first_row, const uint16_t* d_prev_row,
                                   const uint8_t* d_status_in, const conan_pair_stats_out* out,
                                   const conan_frame_stats_out* fout, void* d_work, size_t work_bytes,
                                   int32_t flags, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!out || !fout || !d_work) return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: null pointer argument");
    if (!out->d_sum_k || !out->d_sum_k2 || !out->d_n_int || !out->d_n_enc || !out->d_first_f ||
        !out->d_last_f || !out->d_status)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: a per-pair output pointer is null");
    if (!fout->d_native || !fout->d_nonnative || !fout->d_ssd_prev || !fout->d_ssd_first)
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: a per-frame output pointer is null");
    int rc = check_common(d_levels, levels_stride, nf, P, nlevels, lo_cnt, &hi_first, equal_cut, "conan_pair_stats");
    if (rc != CONAN_OK) return rc;
    if ((d_first_row && (reinterpret_cast<uintptr_t>(d_first_row) & 7) != 0) ||
        (d_prev_row && (reinterpret_cast<uintptr_t>(d_prev_row) & 7) != 0))
        return conan::fail(CONAN_ERR_ARG, "conan_pair_stats: level rows must be 8-byte aligned");
    if (frame0 < 0 || frame0 + nf > 2147483647ll)
        return conan::fail(CONAN_ERR_SIZE, "conan_pair_stats: frame indices must fit in int32");
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
    rec.status_out = reinterpret_cast<uint8_t*>(w + pl.off[6]);
    rec.enter = reinterpret_cast<uint8_t*>(w + pl.off[7]);
    rec.Ppad = pl.Ppad;

    CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_native, 0, sizeof(int) * (size_t)nf, stream));
    CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_nonnative, 0, sizeof(int) * (size_t)nf, stream));
    CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_ssd_prev, 0, sizeof(int64_t) * (size_t)nf, stream));
    CONAN_CUDA_TRY(cudaMemsetAsync(fout->d_ssd_first, 0, sizeof(int64_t) * (size_t)nf, stream));

    StatsParams sp;
    sp.levels = d_levels; sp.stride = levels_stride; sp.nf = nf; sp.P = P; sp.frame0 = frame0;
    sp.chunk_len = pl.chunk_len; sp.nchunks = pl.nchunks; sp.lo_cnt = lo_cnt; sp.hi_first = hi_first; sp.one = 1;
    sp.first_row = d_first_row ? d_first_row : d_levels;
    sp.prev_row = d_prev_row;
    sp.status_in = d_status_in;
    sp.decided = out->d_decided;
    sp.native = fout->d_native; sp.nonnative = fout->d_nonnative;
    sp.ssd_prev = reinterpret_cast<unsigned long long*>(fout->d_ssd_prev);
    sp.ssd_first = reinterpret_cast<unsigned long long*>(fout->d_ssd_first);
    sp.rec = rec;
    const int64_t tiles = (P + kPairsPerCta - 1) / kPairsPerCta;
    if (tiles > 2147483647ll) return conan::fail(CONAN_ERR_SIZE, "conan_pair_stats: too many pairs");

    dim3 egrid((unsigned)((P + 255) / 256), (unsigned)(pl.nchunks + 1));
    entering_status_kernel<<<egrid, 256, 0, stream>>>(sp);
    CONAN_CUDA_TRY(cudaGetLastError());
    dim3 grid((unsigned)tiles, (unsigned)pl.nchunks);
    const size_t smem = (size_t)kStagesS * kStageFrames * kPairsPerCta * sizeof(uint16_t);
    // The sparse variant skips stages of nothing but last levels; it needs a pair at the last level to be off
    // whatever it was before, i.e. the last level at or above both cut-offs (else it is not used).
    const bool sparse = (flags & CONAN_STATS_SPARSE) && nlevels - 1 >= hi_first && nlevels - 1 >= lo_cnt;
    sp.quiet_word = sparse ? (unsigned)(nlevels - 1) * 0x00010001u : 0u;
#define CONAN_LAUNCH_STATS(W, Q)                                                                                       \
    do {                                                                                                                \
        CONAN_CUDA_TRY(cudaFuncSetAttribute(pair_stats_kernel<W, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        pair_stats_kernel<W, Q><<<grid, kThreads + 32, smem, stream>>>(sp);                                           \
    } while (0)
    if (nlevels <= kSmallLevels) {
        if (sparse) CONAN_LAUNCH_STATS(false, true); else CONAN_LAUNCH_STATS(false, false);
    } else {
        if (sparse) CONAN_LAUNCH_STATS(true, true); else CONAN_LAUNCH_STATS(true, false);
    }
#undef CONAN_LAUNCH_STATS
    CONAN_CUDA_TRY(cudaGetLastError());

    CombineParams cp;
    cp.rec = rec; cp.nchunks = pl.nchunks; cp.P = P; cp.out = *out;
    combine_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(cp);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

namespace {
// The (decided, status) summary of a whole frame shard: the backward scan of entering_status_kernel
// from the shard's end, without the streaming pass.
__global__ void shard_summary_kernel(const uint16_t* __restrict__ levels, int64_t stride, int64_t nf, int64_t P,
                                     int lo_cnt, int hi_first, uint8_t* __restrict__ decided,
                                     uint8_t* __restrict__ status) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = p < P;
    int64_t t = nf - 1;
    int st = live ? -1 : 0;
    while (true) {
        if (st < 0 && t >= 0) {
            const int k = levels[t * stride + p];
            if (k < lo_cnt) st = 1;
            else if (k >= hi_first) st = 0;
        }
        t--;
        if (__all_sync(0xffffffffu, st >= 0 || t < 0)) break;
    }
    if (!live) return;
    decided[p] = st >= 0 ? 1 : 0;
    status[p] = st > 0 ? 1 : 0;
}
}  // namespace

extern "C" int conan_shard_summary(const uint16_t* d_levels, int64_t levels_stride, int64_t nf, int64_t P,
                                   int32_t nlevels, int32_t lo_cnt, int32_t hi_first, int32_t equal_cut,
                                   uint8_t* d_decided, uint8_t* d_status, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_decided || !d_status) return conan::fail(CONAN_ERR_ARG, "conan_shard_summary: null output pointer");
    int rc = check_common(d_levels, levels_stride, nf, P, nlevels, lo_cnt, &hi_first, equal_cut, "conan_shard_summary");
    if (rc != CONAN_OK) return rc;
    shard_summary_kernel<<<(unsigned)((P + 255) / 256), 256, 0, stream>>>(d_levels, levels_stride, nf, P, lo_cnt,
                                                                        hi_first, d_decided, d_status);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
