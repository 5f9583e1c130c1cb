// Stage 1 on sm_100a: per-frame residue x residue minimum-distance matrix, quantised in-kernel
// to the xpm level index gmx mdmat would have written.
//
// Replaces `gmx mdmat -frames` as invoked by the reference (conan.py:1437-1453, :1783-1784)
// together with the text parse of dmf.xpm (conan.py:455-567): the uint16 level stack written
// here is what read_frame would have looked up in dict_legend.
//
// Two kernels:
//   prep_slots_kernel   [nf][natoms][3] coordinates -> a TMA-friendly "slot" layout
//                       slots[f][c][k][res] (c = x/y/z, k = atom slot inside the residue, residues
//                       padded to a multiple of the tile; empty slots repeat the residue's first
//                       atom so a minimum over slots is unaffected) plus one bounding sphere per
//                       residue.
//   mdmat_tile_kernel   one CTA per (frame, 64x64 residue tile).  Thread 0 issues the TMA tile
//                       loads (cp.async.bulk.tensor.4d) for the two residue tiles; while they are
//                       in flight all threads prune residue pairs by bounding sphere and compact
//                       the survivors into a shared-memory list; survivors then get the exact
//                       float32 minimum over atom pairs (same operation order as the CPU
//                       restatement: no FMA contraction), sqrt, level = rint(r*invlevel); the
//                       64x64 uint16 tile is staged in shared memory and written out row-wise.
//
// The kernel is FP32-pipe bound (8 flops per atom pair that survives pruning); see DESIGN.md.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "common.h"

namespace {

constexpr int kTile = 64;        // residues per tile side
constexpr int kThreads = 256;    // threads per CTA of mdmat_tile_kernel
constexpr int kMaxKpad = 128;    // largest supported atoms-per-residue (after padding to 4)
constexpr float kPruneSlack = 1.00001f;

__host__ __device__ inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// ------------------------------------------------------------------------------------------
// prep: slots + bounding spheres
// ------------------------------------------------------------------------------------------
__global__ void prep_slots_kernel(const float* __restrict__ xyz, const int* __restrict__ res_start,
                                  int64_t nf, int natoms, int nres, int nres_pad, int kpad,
                                  float* __restrict__ slots, float4* __restrict__ sph,
                                  int* __restrict__ cnt) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= nf * nres_pad) return;
    int64_t f = gid / nres_pad;
    int r = (int)(gid - f * nres_pad);
    float* sl = slots + (size_t)f * 3 * kpad * nres_pad + r;
    const size_t kstride = (size_t)nres_pad;
    const size_t cstride = (size_t)kpad * nres_pad;
    if (r >= nres) {
        for (int k = 0; k < kpad; k++) {
            sl[k * kstride] = 0.f; sl[cstride + k * kstride] = 0.f; sl[2 * cstride + k * kstride] = 0.f;
        }
        sph[(size_t)f * nres_pad + r] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (f == 0) cnt[r] = 0;
        return;
    }
    const int a0 = res_start[r], a1 = res_start[r + 1];
    const int n = a1 - a0;
    const float* x = xyz + ((size_t)f * natoms + a0) * 3;
    float sx = 0.f, sy = 0.f, sz = 0.f;
    for (int k = 0; k < n; k++) { sx += x[3 * k]; sy += x[3 * k + 1]; sz += x[3 * k + 2]; }
    const float inv = 1.0f / (float)n;
    const float cx = sx * inv, cy = sy * inv, cz = sz * inv;
    float r2max = 0.f;
    const float x0 = x[0], y0 = x[1], z0 = x[2];
    for (int k = 0; k < kpad; k++) {
        float px = x0, py = y0, pz = z0;
        if (k < n) {
            px = x[3 * k]; py = x[3 * k + 1]; pz = x[3 * k + 2];
            float dx = px - cx, dy = py - cy, dz = pz - cz;
            r2max = fmaxf(r2max, dx * dx + dy * dy + dz * dz);
        }
        sl[k * kstride] = px; sl[cstride + k * kstride] = py; sl[2 * cstride + k * kstride] = pz;
    }
    sph[(size_t)f * nres_pad + r] = make_float4(cx, cy, cz, sqrtf(r2max));
    if (f == 0) cnt[r] = n;
}

// ------------------------------------------------------------------------------------------
// PTX helpers (mbarrier + TMA)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a TMA that never lands must not hang the GPU; trap instead.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spin = 0; spin < (1u << 24); spin++)
        if (mbar_try_wait(bar, parity)) return;
    if ((threadIdx.x & 31) == 0)
        printf("conan mdmat: TMA tile load never completed (block %d warp %d)\n", (int)blockIdx.x, (int)(threadIdx.x >> 5));
    __trap();
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0,
                                            int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}

// ------------------------------------------------------------------------------------------
// main kernel
// ------------------------------------------------------------------------------------------
struct MdmatParams {
    const float4* sph;       // [nf][nres_pad]
    const int* cnt;          // [nres_pad]
    const float* box;        // [nf][9] or nullptr
    uint16_t* levels;        // [nf][stride]
    int64_t stride;          // elements between frames
    unsigned long long* counters;   // [0] near-boundary results [1] un-pruned residue pairs [2] atom pairs evaluated
    int64_t P;
    int64_t frame_base;      // first frame of this launch
    int nres, nres_pad, kpad;
    int symmetric;
    int x0, xres, y0, yres;  // block mode (symmetric: x0=y0=0, xres=yres=nres)
    int tx0, ty0;            // first 64-aligned tile of the x / y residue range
    int ntx, nty, ntiles;    // tile grid
    float trunc, trunc2, invlevel;
    int maxlevel;
    int pbc;
    int use_tma;             // 0: stage tiles with plain loads (debug aid, CONAN_MDMAT_NO_TMA=1)
    const float* slots;      // [nf][3][kpad][nres_pad], for the plain-load path
};

__device__ __forceinline__ float min_image(float d, float L) {
    // gmx pbc_dx, rectangular box: shift by whole box lengths into (-L/2, L/2]
    if (L > 0.f) {
        const float h = 0.5f * L;
        while (d > h) d -= L;
        while (d <= -h) d += L;
    }
    return d;
}

// r2 exactly as the CPU restatement computes it: every operation rounded, no contraction.
__device__ __forceinline__ float dist2_exact(float dx, float dy, float dz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

template <int ICH>
__device__ __forceinline__ float chunk_min(const float* __restrict__ aI, const float* __restrict__ aJ,
                                           int cstride, int a0, int cj, float m) {
    float xi[ICH], yi[ICH], zi[ICH];
#pragma unroll
    for (int u = 0; u < ICH; u++) {
        xi[u] = aI[(a0 + u) * kTile];
        yi[u] = aI[cstride + (a0 + u) * kTile];
        zi[u] = aI[2 * cstride + (a0 + u) * kTile];
    }
#pragma unroll 2
    for (int b = 0; b < cj; b++) {
        const float xj = aJ[b * kTile], yj = aJ[cstride + b * kTile], zj = aJ[2 * cstride + b * kTile];
#pragma unroll
        for (int u = 0; u < ICH; u++) m = fminf(m, dist2_exact(xi[u] - xj, yi[u] - yj, zi[u] - zj));
    }
    return m;
}

__global__ void __launch_bounds__(kThreads, 3)
mdmat_tile_kernel(const __grid_constant__ CUtensorMap tmap, const MdmatParams prm) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int kpad = prm.kpad;
    const int cstride = kpad * kTile;                 // floats between x / y / z planes of a tile
    float* atomsI = reinterpret_cast<float*>(smem_raw);
    float* atomsJ = atomsI + 3 * cstride;
    float4* sphI = reinterpret_cast<float4*>(atomsJ + 3 * cstride);
    float4* sphJ = sphI + kTile;
    int* cntI = reinterpret_cast<int*>(sphJ + kTile);
    int* cntJ = cntI + kTile;
    uint16_t* outT = reinterpret_cast<uint16_t*>(cntJ + kTile);   // [kTile][kTile + 2]
    constexpr int kOutStride = kTile + 2;
    uint16_t* list = outT + kTile * kOutStride;                   // [kTile*kTile]
    uint64_t* bar = reinterpret_cast<uint64_t*>(list + kTile * kTile);
    int* counters = reinterpret_cast<int*>(bar + 1);              // [0] survivors, [1] boundary hits, [2] atom pairs

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t blk = blockIdx.x;
    const int64_t f_local = blk / prm.ntiles;
    const int t = (int)(blk - f_local * prm.ntiles);
    const int64_t f = prm.frame_base + f_local;
    int ti, tj;
    if (prm.symmetric) {
        // t enumerates tiles (ti <= tj) row-major: row ti starts at ti*ntx - ti*(ti-1)/2
        const float nt = (float)prm.ntx;
        ti = (int)((2.f * nt + 1.f - sqrtf((2.f * nt + 1.f) * (2.f * nt + 1.f) - 8.f * (float)t)) * 0.5f);
        if (ti < 0) ti = 0;
        if (ti > prm.ntx - 1) ti = prm.ntx - 1;
        while (ti > 0 && ti * prm.ntx - ti * (ti - 1) / 2 > t) ti--;
        while ((ti + 1) * prm.ntx - (ti + 1) * ti / 2 <= t) ti++;
        tj = ti + (t - (ti * prm.ntx - ti * (ti - 1) / 2));
    } else {
        ti = t / prm.nty;
        tj = t - ti * prm.nty;
    }
    const bool diag = prm.symmetric && ti == tj;
    // Tiles are aligned to multiples of kTile in the index group's residue numbering, so TMA
    // coordinates are always tile-aligned and inside the padded tensor.
    const int gi0 = (prm.tx0 + ti) * kTile;   // first residue of the row tile
    const int gj0 = (prm.ty0 + tj) * kTile;
    const int i_beg = prm.x0, i_end = prm.x0 + prm.xres;   // residue bounds [beg, end)
    const int j_beg = prm.y0, j_end = prm.y0 + prm.yres;

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
        counters[0] = 0;
        counters[1] = 0;
        counters[2] = 0;
    }
    __syncthreads();
    if (prm.use_tma) {
        if (tid == 0) {
            const uint32_t tile_bytes = 3u * (uint32_t)cstride * 4u;
            mbar_expect_tx(bar, diag ? tile_bytes : 2u * tile_bytes);
            tma_load_4d(atomsI, &tmap, bar, gi0, 0, 0, (int)f);
            if (!diag) tma_load_4d(atomsJ, &tmap, bar, gj0, 0, 0, (int)f);
        }
    } else {
        const float* src = prm.slots + (size_t)f * 3 * kpad * prm.nres_pad;
        for (int idx = tid; idx < 3 * cstride; idx += kThreads) {
            const int r = idx & (kTile - 1), ck = idx >> 6;   // ck = c*kpad + k
            atomsI[idx] = gi0 + r < prm.nres_pad ? src[(size_t)ck * prm.nres_pad + gi0 + r] : 0.f;
            if (!diag) atomsJ[idx] = gj0 + r < prm.nres_pad ? src[(size_t)ck * prm.nres_pad + gj0 + r] : 0.f;
        }
    }
    if (tid < kTile) {
        const int g = gi0 + tid;
        sphI[tid] = g < prm.nres_pad ? prm.sph[(size_t)f * prm.nres_pad + g] : make_float4(0, 0, 0, 0);
        cntI[tid] = g < prm.nres_pad ? prm.cnt[g] : 0;
    } else if (tid < 2 * kTile) {
        const int l = tid - kTile, g = gj0 + l;
        sphJ[l] = g < prm.nres_pad ? prm.sph[(size_t)f * prm.nres_pad + g] : make_float4(0, 0, 0, 0);
        cntJ[l] = g < prm.nres_pad ? prm.cnt[g] : 0;
    }
    float Lx = 0.f, Ly = 0.f, Lz = 0.f;
    if (prm.pbc) {
        const float* b = prm.box + 9 * f;
        Lx = b[0]; Ly = b[4]; Lz = b[8];
    }
    __syncthreads();

    // ---- phase 1: bounding-sphere pruning, survivors compacted into `list` -----------------
    {
        float4 sj[2];
        sj[0] = sphJ[lane];
        sj[1] = sphJ[lane + 32];
        const int gj[2] = {gj0 + lane, gj0 + lane + 32};
#pragma unroll 1
        for (int ii = 0; ii < kTile / (kThreads / 32); ii++) {
            const int i = warp * (kTile / (kThreads / 32)) + ii;
            const int gi = gi0 + i;
            const float4 si = sphI[i];
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int j = lane + 32 * h;
                const bool valid = gi >= i_beg && gi < i_end && gj[h] >= j_beg && gj[h] < j_end &&
                                   (!prm.symmetric || gj[h] > gi);
                bool keep = false, general = false;
                if (valid) {
                    float dx = si.x - sj[h].x, dy = si.y - sj[h].y, dz = si.z - sj[h].z;
                    const float rr = si.w + sj[h].w;
                    if (prm.pbc) {
                        // a plain (unwrapped) inner loop is exact only if no atom pair can need a shift
                        general = !(fabsf(dx) + rr < 0.49999f * Lx && fabsf(dy) + rr < 0.49999f * Ly &&
                                    fabsf(dz) + rr < 0.49999f * Lz);
                        dx = min_image(dx, Lx); dy = min_image(dy, Ly); dz = min_image(dz, Lz);
                    }
                    const float d2 = dx * dx + dy * dy + dz * dz;
                    const float T = (prm.trunc + rr) * kPruneSlack;
                    keep = !(d2 > T * T);
                    if (!keep) outT[i * kOutStride + j] = (uint16_t)prm.maxlevel;
                }
                const unsigned ball = __ballot_sync(0xffffffffu, keep);
                if (ball) {
                    int base = 0;
                    if (lane == 0) base = atomicAdd(&counters[0], __popc(ball));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (keep)
                        list[base + __popc(ball & ((1u << lane) - 1))] =
                            (uint16_t)(i | (j << 6) | (general ? 0x1000 : 0));
                }
            }
        }
    }
    __syncthreads();
    if (prm.use_tma) mbar_wait(bar, 0);   // atom tiles have landed

    // ---- phase 2: exact minimum over atom pairs for the survivors --------------------------
    {
        const int nsurv = counters[0];
        const float* aJbase = diag ? atomsI : atomsJ;
        int near = 0, evals = 0;
        for (int s = tid; s < nsurv; s += kThreads) {
            const int e = list[s];
            const int i = e & 63, j = (e >> 6) & 63;
            const int ci = cntI[i], cj = diag ? cntI[j] : cntJ[j];
            const float* aI = atomsI + i;
            const float* aJ = aJbase + j;
            float m = prm.trunc2;
            evals += ci * cj;
            if (!(e & 0x1000)) {
                int a0 = 0;
                while (ci - a0 > 4) { m = chunk_min<8>(aI, aJ, cstride, a0, cj, m); a0 += 8; }
                if (ci - a0 > 0) m = chunk_min<4>(aI, aJ, cstride, a0, cj, m);
            } else {
                for (int a = 0; a < ci; a++) {
                    const float xi = aI[a * kTile], yi = aI[cstride + a * kTile], zi = aI[2 * cstride + a * kTile];
                    for (int b = 0; b < cj; b++) {
                        const float dx = min_image(xi - aJ[b * kTile], Lx);
                        const float dy = min_image(yi - aJ[cstride + b * kTile], Ly);
                        const float dz = min_image(zi - aJ[2 * cstride + b * kTile], Lz);
                        m = fminf(m, dist2_exact(dx, dy, dz));
                    }
                }
            }
            const float r = __fsqrt_rn(m);
            const float tq = __fmul_rn(r, prm.invlevel);
            int c = __float2int_rn(tq);
            c = max(0, min(c, prm.maxlevel));
            outT[i * kOutStride + j] = (uint16_t)c;
            // results within 2 ulp of a level boundary (x.5) may legitimately differ from gmx
            const float frac = fabsf(tq - rintf(tq));
            near += (fabsf(frac - 0.5f) <= tq * 2.4e-7f) ? 1 : 0;
        }
        if (prm.counters) {
            near = __reduce_add_sync(0xffffffffu, near);
            evals = __reduce_add_sync(0xffffffffu, evals);
            if (lane == 0 && near) atomicAdd(&counters[1], near);
            if (lane == 0 && evals) atomicAdd(&counters[2], evals);
        }
    }
    __syncthreads();

    // ---- phase 3: write the tile, one warp per row ------------------------------------------
    {
        uint16_t* frame_out = prm.levels + (size_t)f * prm.stride;
        for (int i = warp; i < kTile; i += kThreads / 32) {
            const int gi = gi0 + i;
            if (gi < i_beg) continue;
            if (gi >= i_end) break;
            // valid columns of this row: [jlo, jhi) in tile coordinates
            int jlo = max(0, j_beg - gj0), jhi = min(kTile, j_end - gj0);
            int64_t rowbase;   // element index of column gj0 in this row
            if (prm.symmetric) {
                if (diag) jlo = max(jlo, i + 1);
                rowbase = (int64_t)gi * prm.nres - (int64_t)gi * (gi + 1) / 2 + (gj0 - gi - 1);
            } else {
                rowbase = (int64_t)(gi - prm.x0) * prm.yres + (gj0 - prm.y0);
            }
            if (jlo >= jhi) continue;
            const uint16_t* src = outT + i * kOutStride;
            uint16_t* dst = frame_out + rowbase;
            // peel to 4-byte alignment, then 32-bit stores
            int j = jlo;
            if (((reinterpret_cast<uintptr_t>(dst + j)) & 3) != 0) {
                if (lane == 0) dst[j] = src[j];
                j++;
            }
            const int npair = (jhi - j) >> 1;
            for (int q = lane; q < npair; q += 32) {
                const int jj = j + 2 * q;
                const uint32_t v = (uint32_t)src[jj] | ((uint32_t)src[jj + 1] << 16);
                *reinterpret_cast<uint32_t*>(dst + jj) = v;
            }
            if (((jhi - j) & 1) && lane == 0) dst[jhi - 1] = src[jhi - 1];
        }
        if (tid == 0 && prm.counters) {
            if (counters[1]) atomicAdd(&prm.counters[0], (unsigned long long)counters[1]);
            if (counters[0]) atomicAdd(&prm.counters[1], (unsigned long long)counters[0]);
            if (counters[2]) atomicAdd(&prm.counters[2], (unsigned long long)counters[2]);
        }
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct Layout {
    int nres_pad, kpad;
    size_t off_res_start, off_cnt, off_sph, off_slots, total;
};

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

bool make_layout(int64_t nf, int nres, int max_res_atoms, Layout* L) {
    if (nf <= 0 || nres <= 0 || max_res_atoms <= 0) return false;
    L->nres_pad = round_up(nres, kTile);
    L->kpad = round_up(max_res_atoms, 4);
    size_t o = 0;
    L->off_res_start = o; o = align256(o + sizeof(int) * (size_t)(nres + 1));
    L->off_cnt = o;       o = align256(o + sizeof(int) * (size_t)L->nres_pad);
    L->off_sph = o;       o = align256(o + sizeof(float4) * (size_t)nf * L->nres_pad);
    L->off_slots = o;     o = align256(o + sizeof(float) * (size_t)nf * 3 * L->kpad * L->nres_pad);
    L->total = o;
    return true;
}

size_t mdmat_smem_bytes(int kpad) {
    size_t s = 0;
    s += 2 * 3 * (size_t)kpad * kTile * sizeof(float);   // atom tiles
    s += 2 * kTile * sizeof(float4);                      // spheres
    s += 2 * kTile * sizeof(int);                         // counts
    s += (size_t)kTile * (kTile + 2) * sizeof(uint16_t);  // output tile
    s += (size_t)kTile * kTile * sizeof(uint16_t);        // survivor list
    s += 8 + 4 * sizeof(int);                             // mbarrier + counters
    return (s + 15) & ~(size_t)15;
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

}  // namespace

extern "C" size_t conan_mdmat_workspace_bytes(int64_t nf, int32_t natoms, int32_t nres, int32_t max_res_atoms) {
    Layout L;
    if (natoms <= 0 || !make_layout(nf, nres, max_res_atoms, &L)) return 0;
    return L.total;
}

extern "C" int conan_mdmat_frames(const float* d_xyz, const float* d_box, int64_t nf, int32_t natoms,
                                  const int32_t* h_res_start, int32_t nres, int32_t symmetric, int32_t x0,
                                  int32_t xres, int32_t y0, int32_t yres, float trunc, int32_t nlevels,
                                  int32_t pbc, uint16_t* d_levels, int64_t levels_stride,
                                  unsigned long long* d_counters, void* d_work, size_t work_bytes,
                                  void* stream_) {
    return conan_mdmat_frames_ex(d_xyz, d_box, nf, natoms, h_res_start, nres, symmetric, x0, xres, y0, yres, trunc,
                                 nlevels, pbc, d_levels, levels_stride, d_counters, d_work, work_bytes, stream_, 3);
}

extern "C" int conan_mdmat_frames_ex(const float* d_xyz, const float* d_box, int64_t nf, int32_t natoms,
                                     const int32_t* h_res_start, int32_t nres, int32_t symmetric, int32_t x0,
                                     int32_t xres, int32_t y0, int32_t yres, float trunc, int32_t nlevels,
                                     int32_t pbc, uint16_t* d_levels, int64_t levels_stride,
                                     unsigned long long* d_counters, void* d_work, size_t work_bytes,
                                     void* stream_, int32_t phases) {
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    if (!d_xyz || !h_res_start || !d_levels || !d_work)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: null pointer argument");
    if (nf < 0 || natoms <= 0 || nres <= 0)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: nf=%lld natoms=%d nres=%d", (long long)nf, natoms, nres);
    if (nlevels < 2 || nlevels > 7744)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: nlevels=%d outside [2, 7744] (the xpm two-character limit)", nlevels);
    if (!(trunc > 0.f)) return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: trunc must be positive");
    if (pbc && !d_box) return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: pbc=1 needs d_box");
    if (nf == 0) return CONAN_OK;
    int max_atoms = 0;
    if (h_res_start[0] != 0 || h_res_start[nres] != natoms)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: res_start must run from 0 to natoms");
    for (int r = 0; r < nres; r++) {
        int n = h_res_start[r + 1] - h_res_start[r];
        if (n <= 0) return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: residue %d has no atoms", r);
        if (n > max_atoms) max_atoms = n;
    }
    if (symmetric) { x0 = 0; y0 = 0; xres = nres; yres = nres; }
    if (x0 < 0 || y0 < 0 || xres <= 0 || yres <= 0 || x0 + xres > nres || y0 + yres > nres)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: residue block [%d,+%d) x [%d,+%d) outside 0..%d",
                           x0, xres, y0, yres, nres);
    Layout L;
    make_layout(nf, nres, max_atoms, &L);
    if (L.kpad > kMaxKpad)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: a residue has %d atoms; the limit is %d", max_atoms, kMaxKpad);
    if (work_bytes < L.total)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: workspace %zu bytes, need %zu", work_bytes, L.total);
    if ((reinterpret_cast<uintptr_t>(d_work) & 255) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: workspace must be 256-byte aligned");

    char* w = static_cast<char*>(d_work);
    int* d_res_start = reinterpret_cast<int*>(w + L.off_res_start);
    int* d_cnt = reinterpret_cast<int*>(w + L.off_cnt);
    float4* d_sph = reinterpret_cast<float4*>(w + L.off_sph);
    float* d_slots = reinterpret_cast<float*>(w + L.off_slots);

    if (phases & 1) {
        CONAN_CUDA_TRY(cudaMemcpyAsync(d_res_start, h_res_start, sizeof(int) * (size_t)(nres + 1),
                                       cudaMemcpyHostToDevice, stream));
        const int64_t total = nf * L.nres_pad;
        const int threads = 128;
        prep_slots_kernel<<<(unsigned)((total + threads - 1) / threads), threads, 0, stream>>>(
            d_xyz, d_res_start, nf, natoms, nres, L.nres_pad, L.kpad, d_slots, d_sph, d_cnt);
        CONAN_CUDA_TRY(cudaGetLastError());
    }
    if (!(phases & 2)) return CONAN_OK;

    EncodeTiledFn encode = get_encode_fn();
    if (!encode) return conan::fail(CONAN_ERR_CUDA, "conan_mdmat_frames: cuTensorMapEncodeTiled unavailable");

    const size_t smem = mdmat_smem_bytes(L.kpad);
    CONAN_CUDA_TRY(cudaFuncSetAttribute(mdmat_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

    MdmatParams prm;
    prm.sph = d_sph; prm.cnt = d_cnt; prm.box = d_box; prm.levels = d_levels; prm.counters = d_counters;
    prm.P = symmetric ? (int64_t)nres * (nres - 1) / 2 : (int64_t)xres * yres;
    prm.nres = nres; prm.nres_pad = L.nres_pad; prm.kpad = L.kpad; prm.symmetric = symmetric ? 1 : 0;
    prm.x0 = x0; prm.xres = xres; prm.y0 = y0; prm.yres = yres;
    prm.tx0 = x0 / kTile;
    prm.ty0 = y0 / kTile;
    prm.ntx = (x0 + xres - 1) / kTile - prm.tx0 + 1;
    prm.nty = (y0 + yres - 1) / kTile - prm.ty0 + 1;
    prm.ntiles = symmetric ? prm.ntx * (prm.ntx + 1) / 2 : prm.ntx * prm.nty;
    prm.trunc = trunc; prm.trunc2 = trunc * trunc; prm.invlevel = (float)(nlevels - 1) / trunc;
    prm.maxlevel = nlevels - 1; prm.pbc = pbc ? 1 : 0;
    {
        const char* e = getenv("CONAN_MDMAT_NO_TMA");
        prm.use_tma = (e && e[0] == '1') ? 0 : 1;
    }
    prm.slots = d_slots;
    if (prm.P == 0) return CONAN_OK;
    if (levels_stride < prm.P)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: levels_stride=%lld < P=%lld", (long long)levels_stride, (long long)prm.P);
    prm.stride = levels_stride;

    // Frames per launch: the TMA frame coordinate is an int and the grid is 1-D.
    const int64_t max_blocks = 1ll << 30;
    const int64_t frames_per_launch = max_blocks / prm.ntiles > 0 ? max_blocks / prm.ntiles : 1;
    for (int64_t fb = 0; fb < nf; fb += frames_per_launch) {
        const int64_t nfl = (nf - fb < frames_per_launch) ? nf - fb : frames_per_launch;
        // The tensor map covers only this launch's frames, so coordinate 3 stays small.
        CUtensorMap tmap;
        cuuint64_t gdim[4] = {(cuuint64_t)L.nres_pad, (cuuint64_t)L.kpad, 3, (cuuint64_t)nfl};
        cuuint64_t gstr[3] = {(cuuint64_t)L.nres_pad * 4, (cuuint64_t)L.kpad * L.nres_pad * 4,
                              (cuuint64_t)3 * L.kpad * L.nres_pad * 4};
        cuuint32_t box[4] = {(cuuint32_t)kTile, (cuuint32_t)L.kpad, 3, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        float* base = d_slots + (size_t)fb * 3 * L.kpad * L.nres_pad;
        CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, gdim, gstr, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS)
            return conan::fail(CONAN_ERR_CUDA, "conan_mdmat_frames: cuTensorMapEncodeTiled failed (%d)", (int)cr);
        MdmatParams p2 = prm;
        // frame indices inside the kernel are relative to this launch's base pointers
        p2.sph = d_sph + (size_t)fb * L.nres_pad;
        p2.box = d_box ? d_box + 9 * fb : nullptr;
        p2.levels = d_levels + (size_t)fb * prm.stride;
        p2.frame_base = 0;
        p2.slots = d_slots + (size_t)fb * 3 * L.kpad * L.nres_pad;
        mdmat_tile_kernel<<<(unsigned)(nfl * prm.ntiles), kThreads, smem, stream>>>(tmap, p2);
        CONAN_CUDA_TRY(cudaGetLastError());
    }
    return CONAN_OK;
}

// ------------------------------------------------------------------------------------------
// Allocation-free variant of stage 1
// ------------------------------------------------------------------------------------------
// Some callers of this library are not in a position to hand over a workspace: an analysis hook
// that runs inside another program's time-step loop (contact maps computed on the fly while an
// MD engine is integrating) must not call cudaMalloc / cudaFree, because those synchronise the
// whole device and stall the engine's streams, and late in a run the engine's own pools may have
// left no memory to allocate from.  For them the scratch that conan_mdmat_frames normally takes
// from the caller lives in this fixed array, which every GPU that loads the library carries
// until the process exits.
static __device__ __align__(256) unsigned char g_mdmat_static_scratch[CONAN_MDMAT_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_static_scratch_ptr[kMaxDevices] = {nullptr};
std::mutex g_static_scratch_mutex;
}  // namespace

extern "C" int conan_mdmat_frames_static(const float* d_xyz, const float* d_box, int64_t nf, int32_t natoms,
                                         const int32_t* h_res_start, int32_t nres, int32_t symmetric,
                                         int32_t x0, int32_t xres, int32_t y0, int32_t yres, float trunc,
                                         int32_t nlevels, int32_t pbc, uint16_t* d_levels,
                                         int64_t levels_stride, unsigned long long* d_counters,
                                         void* stream_) {
    int dev = -1;
    CONAN_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames_static: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
    void* scratch = nullptr;
    {
        // Set-up, once per device and with that device current: in host code the array's name is
        // only a stand-in for the symbol, so its device address has to be asked for.
        std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
        if (!g_static_scratch_ptr[dev]) {
            void* p = nullptr;
            CONAN_CUDA_TRY(cudaGetSymbolAddress(&p, g_mdmat_static_scratch));
            g_static_scratch_ptr[dev] = p;
        }
        scratch = g_static_scratch_ptr[dev];
    }
    if (!h_res_start || nres <= 0 || natoms <= 0 || nf < 0)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames_static: bad argument");
    int max_atoms = 0;
    for (int r = 0; r < nres; r++) {
        const int n = h_res_start[r + 1] - h_res_start[r];
        if (n > max_atoms) max_atoms = n;
    }
    const size_t need = nf > 0 ? conan_mdmat_workspace_bytes(nf, natoms, nres, max_atoms > 0 ? max_atoms : 1) : 0;
    if (need > (size_t)CONAN_MDMAT_STATIC_SCRATCH_BYTES)
        return conan::fail(CONAN_ERR_SIZE,
                           "conan_mdmat_frames_static: %lld frames of %d residues need %zu bytes of scratch but the "
                           "built-in scratch holds %zu; pass fewer frames per call or use conan_mdmat_frames",
                           (long long)nf, nres, need, (size_t)CONAN_MDMAT_STATIC_SCRATCH_BYTES);
    // Nothing in the scratch has to start at zero: the first kernel of the call rewrites every
    // byte that the second one reads, so whatever an earlier call left behind is never seen.
    return conan_mdmat_frames_ex(d_xyz, d_box, nf, natoms, h_res_start, nres, symmetric, x0, xres, y0, yres, trunc,
                                 nlevels, pbc, d_levels, levels_stride, d_counters, scratch,
                                 (size_t)CONAN_MDMAT_STATIC_SCRATCH_BYTES, stream_, 3);
}
