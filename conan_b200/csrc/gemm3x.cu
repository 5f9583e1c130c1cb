// Stage 3 on sm_100a: C = A * B^T from three split products on the 5th-generation tensor cores (tcgen05 / TMEM).
//
// Serves the three dense contractions of the reference: the PCA covariance np.cov(vectors)
// (conan.py:1983; A = B = centred pair series, K = frames), the frame-frame contact-map distances
// scipy pdist(vectors.T / norm) (conan.py:2063; A = B = centred, scaled frame vectors, K = pairs,
// epilogue sqrt(|a|^2 + |b|^2 - 2ab)), and pair-pair correlation blocks.
//
// Operands arrive split into a "hi" and a "lo" part (conan_build_operand): x = hi + lo, each part a
// binary16 number -- the same 11-bit significands a TF32 split would give, in half the bytes -- so
// hi*hi + hi*lo + lo*hi recovers float32-level accuracy from three tensor-core products accumulated in
// float32 in TMEM ("3xTF32" with the parts stored as binary16: the kernel is bound by how fast operand
// tiles reach the SM, and the f16 MMA path also runs at twice the TF32 rate).
//
// One CTA computes one 128 x 128 output tile: warp 0 = TMA producer (four 128-row x 64-element boxes
// per k-block, 128-byte swizzle), warp 1 = TMEM allocator and single-thread tcgen05.mma issuer
// (UMMA 128x128x16, twelve per k-block), warps 2-5 = epilogue (tcgen05.ld 32 lanes x 32 columns,
// scaling / distance epilogue, 128-bit stores).  A three-stage shared-memory ring is handed back and
// forth with mbarriers; tcgen05.commit releases a stage when the MMAs that read it have retired.
// The tensor core adds into its float32 accumulator with truncation, which would let the error grow
// with K (frames or pairs, up to 10^5); so the MMAs accumulate only 128 k-values at a time into one of
// two TMEM accumulators while the epilogue warps drain the other into round-to-nearest float32
// register sums -- the bias stays at the 10^-6 level whatever K is, and the drain hides behind the MMAs.
// Every mbarrier wait is bounded and traps instead of hanging the GPU.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "common.h"

namespace {

constexpr int kBM = 128, kBN = 128, kBK = 64;      // tile; kBK binary16 elements = 128 bytes per row
constexpr int kStages = 3;
constexpr int kOperandBytes = kBM * kBK * 2;        // 16 KB: 128 rows x 128 B
constexpr int kStageBytes = 4 * kOperandBytes;      // A_hi, A_lo, B_hi, B_lo
constexpr int kThreads = 192;
constexpr int kTmemCols = 256;                       // two 128-column accumulators, used alternately
constexpr int kChunkKb = 2;                          // k-blocks accumulated inside the tensor core before a drain
constexpr int kUmmaK = 16;                          // binary16: 32 bytes per MMA along K
constexpr int kBand = 8;                            // tile rows per band of the launch order

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int who) {
    for (uint32_t spin = 0; spin < (1u << 26); spin++)
        if (mbar_try_wait(bar, parity)) return;
    printf("conan gemm3x: barrier wait timed out (role %d, block %d,%d)\n", who, (int)blockIdx.x, (int)blockIdx.y);
    __trap();
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}

// K-major operand tile in shared memory, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffff) >> 4);          // start address, 16-byte units
    d |= (uint64_t)1 << 16;                           // leading byte offset (ignored for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                 // stride byte offset between 8-row groups
    d |= (uint64_t)1 << 46;                           // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                           // SWIZZLE_128B
    return d;
}
// kind::f16: D = F32 (bits 4-5 = 1), A and B binary16 (format fields 0), both K-major, M = 128, N = 128.
constexpr uint32_t kInstrDesc = (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kInstrDesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct GemmParams {
    int M, N, K;                // K a multiple of 64 (operands are zero padded)
    float* C;
    int64_t ldc;
    int epilogue;               // 0: alpha * acc   1: alpha * sqrt(max(0, rn[m] + cn[n] - 2 acc))
    float alpha;
    const double* row_norm2;    // epilogue 1
    const double* col_norm2;
    int symmetric;              // only tiles with tn >= tm are launched; the mirror image is written as well
};

__global__ void __launch_bounds__(kThreads, 1)
gemm3x_kernel(const __grid_constant__ CUtensorMap mapAh, const __grid_constant__ CUtensorMap mapAl,
              const __grid_constant__ CUtensorMap mapBh, const __grid_constant__ CUtensorMap mapBl,
              const GemmParams prm) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // the 128-byte swizzle needs 1024-byte aligned tiles; the launch reserves the slack
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages], tmem_full_bar[2], tmem_empty_bar[2];
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Tiles are walked in bands of kBand tile rows, column by column inside a band, so the CTAs that
    // are resident together share a few A row strips and B column strips in L2.
    const int ntm = (prm.M + kBM - 1) / kBM, ntn = (prm.N + kBN - 1) / kBN;
    const int band = (int)blockIdx.x / (kBand * ntn);
    const int rows_in_band = min(kBand, ntm - band * kBand);
    const int within = (int)blockIdx.x - band * kBand * ntn;
    const int tn = within / rows_in_band;
    const int tm = band * kBand + within % rows_in_band;
    if (prm.symmetric && tn < tm) return;        // lower triangle: written by the mirror store of (tn, tm)
    const int m0 = tm * kBM, n0 = tn * kBN;
    const int nkb = prm.K / kBK;
    const int nchunks = (nkb + kChunkKb - 1) / kChunkKb;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; b++) { mbar_init(&tmem_full_bar[b], 1); mbar_init(&tmem_empty_bar[b], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                     "n"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ---- TMA producer ----
        if (lane == 0) {
            for (int kb = 0; kb < nkb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (uint32_t)(kb / kStages) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u, 0);
                unsigned char* st = smem + (size_t)s * kStageBytes;
                mbar_expect_tx(&full_bar[s], kStageBytes);
                tma_load_2d(st, &mapAh, &full_bar[s], kb * kBK, m0);
                tma_load_2d(st + kOperandBytes, &mapAl, &full_bar[s], kb * kBK, m0);
                tma_load_2d(st + 2 * kOperandBytes, &mapBh, &full_bar[s], kb * kBK, n0);
                tma_load_2d(st + 3 * kOperandBytes, &mapBl, &full_bar[s], kb * kBK, n0);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---- MMA issuer (one thread) ----
        if (lane == 0) {
            for (int kb = 0; kb < nkb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (uint32_t)(kb / kStages) & 1u;
                const int chunk = kb / kChunkKb, buf = chunk & 1;
                const bool chunk_start = (kb % kChunkKb) == 0;
                if (chunk_start) {   // the accumulator must have been drained by the epilogue warps
                    mbar_wait(&tmem_empty_bar[buf], ((uint32_t)(chunk >> 1) & 1u) ^ 1u, 3);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                mbar_wait(&full_bar[s], ph, 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc = tmem_base + (uint32_t)(buf * kBN);
                const uint32_t base = smem_u32(smem + (size_t)s * kStageBytes);
                const uint64_t dAh = make_smem_desc(base), dAl = make_smem_desc(base + kOperandBytes);
                const uint64_t dBh = make_smem_desc(base + 2 * kOperandBytes), dBl = make_smem_desc(base + 3 * kOperandBytes);
#pragma unroll
                for (int k = 0; k < kBK / kUmmaK; k++) {
                    const uint64_t adv = (uint64_t)((k * kUmmaK * 2) >> 4);   // 32 bytes per step, 16-byte units
                    umma_f16(acc, dAh + adv, dBh + adv, !(chunk_start && k == 0));
                    umma_f16(acc, dAh + adv, dBl + adv, 1);
                    umma_f16(acc, dAl + adv, dBh + adv, 1);
                }
                umma_commit(&empty_bar[s]);          // the stage may be refilled once these MMAs retire
                if ((kb % kChunkKb) == kChunkKb - 1 || kb == nkb - 1)
                    umma_commit(&tmem_full_bar[buf]);   // this chunk's accumulator is complete
            }
        }
        __syncwarp();
    } else {
        // ---- epilogue: 4 warps, each owns the 32 TMEM lanes of its quarter ----
        const int q = warp & 3;
        float racc[kBN];
#pragma unroll
        for (int j = 0; j < kBN; j++) racc[j] = 0.f;
        for (int chunk = 0; chunk < nchunks; chunk++) {
            const int buf = chunk & 1;
            mbar_wait(&tmem_full_bar[buf], (uint32_t)(chunk >> 1) & 1u, 2);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int c = 0; c < kBN / 32; c++) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * kBN + c * 32), v);
#pragma unroll
                for (int j = 0; j < 32; j++) racc[c * 32 + j] += __uint_as_float(v[j]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0)
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&tmem_empty_bar[buf])) : "memory");
        }
        const int m = m0 + q * 32 + lane;
        const float rn = (prm.epilogue == 1 && m < prm.M) ? (float)prm.row_norm2[m] : 0.f;
#pragma unroll
        for (int c = 0; c < kBN / 32; c++) {
            const int nb = n0 + c * 32;
            float o[32];
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const float acc = racc[c * 32 + j];
                if (prm.epilogue == 1) {
                    const float cn = (nb + j < prm.N) ? (float)prm.col_norm2[nb + j] : 0.f;
                    o[j] = prm.alpha * sqrtf(fmaxf(rn + cn - 2.f * acc, 0.f));
                } else {
                    o[j] = prm.alpha * acc;
                }
            }
            if (m < prm.M) {
                float* dst = prm.C + (size_t)m * prm.ldc + nb;
                if (nb + 32 <= prm.N && (prm.ldc & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
                } else {
                    for (int j = 0; j < 32; j++) if (nb + j < prm.N) dst[j] = o[j];
                }
                if (prm.symmetric && tm != tn) {
                    for (int j = 0; j < 32; j++)
                        if (nb + j < prm.N) prm.C[(size_t)(nb + j) * prm.ldc + m] = o[j];
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTmemCols) : "memory");
    }
}

// ------------------------------------------------------------------------------------------
// CTA pairs.  The single-CTA kernel above is bound by the bytes it stages per MMA: 64 KB of operand tiles per
// k-block and SM.  Here two CTAs on neighbouring SMs (a cluster of two) compute one 256 x 256 tile with
// tcgen05.mma.cta_group::2: each CTA stages its own 128 rows of A and its own 128 of the tile's 256 B rows
// -- the same 64 KB per k-block -- and the pair's MMAs (M = 256, N = 256, issued by the even CTA, reading both
// CTAs' shared memory) do twice the work on them.  Each CTA's TMEM holds its 128 rows x 256 columns, twice
// (all 512 columns), for the same chunked hand-over: 128 k-values per accumulator, drained by sixteen
// epilogue warps (a quarter of the lanes x a quarter of the columns each) into round-to-nearest sums.
// All tensor maps cover this GPU's memory only, and no load is multicast: the pair shares operands through
// the MMA's own reads of the partner's shared memory.
// ------------------------------------------------------------------------------------------
constexpr int kPairThreads = 576;                    // TMA warp, MMA warp, sixteen epilogue warps
constexpr int kPairBN = 256;                         // columns of the pair's tile (and of each CTA's accumulator)
constexpr int kPairTmemCols = 512;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;       // a shared::cluster address with this bit cleared names the even CTA of the pair
// kind::f16, D = F32, A / B binary16 K-major, M = 256 (the pair), N = 256
constexpr uint32_t kInstrDescPair = (1u << 4) | ((uint32_t)(kPairBN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    // the bytes are counted on the even CTA's barrier, whichever CTA of the pair issues the load
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kInstrDescPair), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    // arrives on the barrier at this offset in both CTAs once the pair's MMAs issued so far have retired
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_even_cta(uint64_t* bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerBitMask) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
gemm3x_pair_kernel(const __grid_constant__ CUtensorMap mapAh, const __grid_constant__ CUtensorMap mapAl,
                   const __grid_constant__ CUtensorMap mapBh, const __grid_constant__ CUtensorMap mapBl,
                   const GemmParams prm) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full_bar[kStages], empty_bar[kStages], tmem_full_bar[2], tmem_empty_bar[2];
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();           // 0: the even CTA, which issues the pair's MMAs
    // tiles of 256 x 256, walked in bands of kBand / 2 tile rows (see the single-CTA kernel)
    constexpr int kBandP = kBand / 2;
    const int ntm = (prm.M + 255) / 256, ntn = (prm.N + kPairBN - 1) / kPairBN;
    const int pair = (int)(blockIdx.x >> 1);
    const int band = pair / (kBandP * ntn);
    const int rows_in_band = min(kBandP, ntm - band * kBandP);
    const int within = pair - band * kBandP * ntn;
    const int tn = within / rows_in_band;
    const int tm = band * kBandP + within % rows_in_band;
    if (prm.symmetric && tn < tm) return;              // both CTAs of the pair leave together
    const int m0 = tm * 256 + (int)rank * kBM;          // this CTA's rows of the tile
    const int n0 = tn * kPairBN;
    const int nkb = prm.K / kBK;
    const int nchunks = (nkb + kChunkKb - 1) / kChunkKb;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; s++) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        // an accumulator is free again when the sixteen epilogue warps of both CTAs have drained it
        for (int b = 0; b < 2; b++) { mbar_init(&tmem_full_bar[b], 1); mbar_init(&tmem_empty_bar[b], 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)),
                     "n"(kPairTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    cluster_sync_all();        // barriers of both CTAs initialised, TMEM of both allocated
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ---- TMA producer (one thread in each CTA) ----
        if (lane == 0) {
            for (int kb = 0; kb < nkb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (uint32_t)(kb / kStages) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u, 0);
                unsigned char* st = smem + (size_t)s * kStageBytes;
                if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * kStageBytes);     // both CTAs' bytes
                tma_load_2d_pair(st, &mapAh, &full_bar[s], kb * kBK, m0);
                tma_load_2d_pair(st + kOperandBytes, &mapAl, &full_bar[s], kb * kBK, m0);
                tma_load_2d_pair(st + 2 * kOperandBytes, &mapBh, &full_bar[s], kb * kBK, n0 + (int)rank * kBM);
                tma_load_2d_pair(st + 3 * kOperandBytes, &mapBl, &full_bar[s], kb * kBK, n0 + (int)rank * kBM);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---- MMA issuer: one thread of the even CTA ----
        if (lane == 0 && rank == 0) {
            for (int kb = 0; kb < nkb; kb++) {
                const int s = kb % kStages;
                const uint32_t ph = (uint32_t)(kb / kStages) & 1u;
                const int chunk = kb / kChunkKb, buf = chunk & 1;
                const bool chunk_start = (kb % kChunkKb) == 0;
                if (chunk_start) {
                    mbar_wait(&tmem_empty_bar[buf], ((uint32_t)(chunk >> 1) & 1u) ^ 1u, 3);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
                mbar_wait(&full_bar[s], ph, 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc = tmem_base + (uint32_t)(buf * kPairBN);
                const uint32_t base = smem_u32(smem + (size_t)s * kStageBytes);
                const uint64_t dAh = make_smem_desc(base), dAl = make_smem_desc(base + kOperandBytes);
                const uint64_t dBh = make_smem_desc(base + 2 * kOperandBytes), dBl = make_smem_desc(base + 3 * kOperandBytes);
#pragma unroll
                for (int k = 0; k < kBK / kUmmaK; k++) {
                    const uint64_t adv = (uint64_t)((k * kUmmaK * 2) >> 4);
                    umma_f16_pair(acc, dAh + adv, dBh + adv, !(chunk_start && k == 0));
                    umma_f16_pair(acc, dAh + adv, dBl + adv, 1);
                    umma_f16_pair(acc, dAl + adv, dBh + adv, 1);
                }
                umma_commit_pair(&empty_bar[s]);
                if ((kb % kChunkKb) == kChunkKb - 1 || kb == nkb - 1) umma_commit_pair(&tmem_full_bar[buf]);
            }
        }
        __syncwarp();
    } else {
        // ---- epilogue: sixteen warps; warp w owns TMEM lanes 32 (w & 3) .. and 64 of the 256 columns ----
        const int q = warp & 3, cq = (warp - 2) >> 2;
        constexpr int kCols = kPairBN / 4;
        float racc[kCols];
#pragma unroll
        for (int j = 0; j < kCols; j++) racc[j] = 0.f;
        for (int chunk = 0; chunk < nchunks; chunk++) {
            const int buf = chunk & 1;
            mbar_wait(&tmem_full_bar[buf], (uint32_t)(chunk >> 1) & 1u, 2);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int c = 0; c < kCols / 32; c++) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * kPairBN + cq * kCols + c * 32), v);
#pragma unroll
                for (int j = 0; j < 32; j++) racc[c * 32 + j] += __uint_as_float(v[j]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive_even_cta(&tmem_empty_bar[buf]);
        }
        const int m = m0 + q * 32 + lane;
        const float rn = (prm.epilogue == 1 && m < prm.M) ? (float)prm.row_norm2[m] : 0.f;
#pragma unroll
        for (int c = 0; c < kCols / 32; c++) {
            const int nb = n0 + cq * kCols + c * 32;
            float o[32];
#pragma unroll
            for (int j = 0; j < 32; j++) {
                const float acc = racc[c * 32 + j];
                if (prm.epilogue == 1) {
                    const float cn = (nb + j < prm.N) ? (float)prm.col_norm2[nb + j] : 0.f;
                    o[j] = prm.alpha * sqrtf(fmaxf(rn + cn - 2.f * acc, 0.f));
                } else {
                    o[j] = prm.alpha * acc;
                }
            }
            if (m < prm.M) {
                float* dst = prm.C + (size_t)m * prm.ldc + nb;
                if (nb + 32 <= prm.N && (prm.ldc & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
                } else {
                    for (int j = 0; j < 32; j++) if (nb + j < prm.N) dst[j] = o[j];
                }
                if (prm.symmetric && tm != tn) {
                    for (int j = 0; j < 32; j++)
                        if (nb + j < prm.N) prm.C[(size_t)(nb + j) * prm.ldc + m] = o[j];
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    // nobody leaves while the partner may still signal its barriers or read its shared memory
    __syncthreads();
    cluster_sync_all();
    if (warp == 1) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kPairTmemCols) : "memory");
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

int make_map(EncodeTiledFn enc, CUtensorMap* map, const uint16_t* base, int rows, int64_t ld) {
    cuuint64_t gdim[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)kBM};
    cuuint32_t estr[2] = {1, 1};
    CUresult cr = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<uint16_t*>(base), gdim, gstr, box, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return cr == CUDA_SUCCESS ? 0 : (int)cr;
}

}  // namespace

extern "C" int conan_gemm_3xtf32(const uint16_t* d_a_hi, const uint16_t* d_a_lo, int32_t M, const uint16_t* d_b_hi,
                                 const uint16_t* d_b_lo, int32_t N, int64_t K, int64_t ld, int32_t epilogue, float alpha,
                                 const double* d_row_norm2, const double* d_col_norm2, int32_t symmetric, float* d_c,
                                 int64_t ldc, void* stream_) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_a_hi || !d_a_lo || !d_b_hi || !d_b_lo || !d_c)
        return conan::fail(CONAN_ERR_ARG, "conan_gemm_3xtf32: null pointer argument");
    if (M <= 0 || N <= 0 || K <= 0 || ld < K || ld % 64 != 0 || ldc < N)
        return conan::fail(CONAN_ERR_ARG, "conan_gemm_3xtf32: bad sizes M=%d N=%d K=%lld ld=%lld ldc=%lld (ld must be a "
                           "multiple of 64 and >= K)", M, N, (long long)K, (long long)ld, (long long)ldc);
    if (epilogue == 1 && (!d_row_norm2 || !d_col_norm2))
        return conan::fail(CONAN_ERR_ARG, "conan_gemm_3xtf32: the distance epilogue needs both norm arrays");
    if (epilogue != 0 && epilogue != 1) return conan::fail(CONAN_ERR_ARG, "conan_gemm_3xtf32: unknown epilogue %d", epilogue);
    if (symmetric && M != N) return conan::fail(CONAN_ERR_ARG, "conan_gemm_3xtf32: symmetric needs M == N");
    for (const void* p : {(const void*)d_a_hi, (const void*)d_a_lo, (const void*)d_b_hi, (const void*)d_b_lo})
        if ((reinterpret_cast<uintptr_t>(p) & 15) != 0)
            return conan::fail(CONAN_ERR_ARG, "conan_gemm_3xtf32: operands must be 16-byte aligned");
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return conan::fail(CONAN_ERR_CUDA, "conan_gemm_3xtf32: cuTensorMapEncodeTiled unavailable");
    CUtensorMap mAh, mAl, mBh, mBl;
    int e = make_map(enc, &mAh, d_a_hi, M, ld);
    if (!e) e = make_map(enc, &mAl, d_a_lo, M, ld);
    if (!e) e = make_map(enc, &mBh, d_b_hi, N, ld);
    if (!e) e = make_map(enc, &mBl, d_b_lo, N, ld);
    if (e) return conan::fail(CONAN_ERR_CUDA, "conan_gemm_3xtf32: cuTensorMapEncodeTiled failed (%d)", e);
    GemmParams prm;
    prm.M = M; prm.N = N; prm.K = (int)((K + kBK - 1) / kBK * kBK);
    if (prm.K > ld) prm.K = (int)ld;
    prm.C = d_c; prm.ldc = ldc; prm.epilogue = epilogue; prm.alpha = alpha;
    prm.row_norm2 = d_row_norm2; prm.col_norm2 = d_col_norm2; prm.symmetric = symmetric ? 1 : 0;
    const size_t smem = (size_t)kStages * kStageBytes + 1024;
    // CTA pairs (256 x 256 tiles) unless the problem is a single tile wide or CONAN_GEMM_CFG=0 asks for the
    // single-CTA kernel (128 x 128 tiles)
    const char* cfg = getenv("CONAN_GEMM_CFG");
    const bool pairs = cfg ? atoi(cfg) != 0 : (M > kBM || N > kBN);
    if (pairs) {
        CONAN_CUDA_TRY(cudaFuncSetAttribute(gemm3x_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int tmn = (M + 255) / 256, tnn = (N + kPairBN - 1) / kPairBN;
        dim3 grid((unsigned)(2 * tmn * tnn), 1);
        gemm3x_pair_kernel<<<grid, kPairThreads, smem, stream>>>(mAh, mAl, mBh, mBl, prm);
    } else {
        CONAN_CUDA_TRY(cudaFuncSetAttribute(gemm3x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int tmn = (M + kBM - 1) / kBM, tnn = (N + kBN - 1) / kBN;
        dim3 grid((unsigned)(tmn * tnn), 1);
        gemm3x_kernel<<<grid, kThreads, smem, stream>>>(mAh, mAl, mBh, mBl, prm);
    }
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
