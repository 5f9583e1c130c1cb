// Stage 1 on sm_100a: per-frame residue x residue minimum-distance matrix, quantised in-kernel
// to the xpm level index gmx mdmat would have written.
//
// Replaces `gmx mdmat -frames` as invoked by the reference (conan.py:1437-1453, :1783-1784)
// together with the text parse of dmf.xpm (conan.py:455-567): the uint16 level stack written
// here is what read_frame would have looked up in dict_legend.
//
// Three kernels:
//   block_perm_kernel   orders the residues of every block of 64 by their number of atoms.
//   prep_slots_kernel   [nf][natoms][3] coordinates -> a TMA-friendly "slot" layout: every residue's atoms
//                       relative to the centre of a near-minimal bounding sphere, four planes
//                       (x', y', z', |a'|^2) laid out [frame][plane][residue][k]; k = atom slot inside the
//                       residue (a residue's slots are contiguous, so both roles read them with 16-byte
//                       loads), residues padded to a multiple of the tile, empty slots repeat the residue's
//                       first atom (a minimum over slots is unaffected); plus the bounding sphere (centre,
//                       radius) per residue.
//   mdmat_tile_kernel   one CTA per (64x64 residue tile, group of consecutive frames).
//                       Thread 0 issues two TMA tile loads (cp.async.bulk.tensor.4d) covering the
//                       whole frame group; meanwhile the CTA fills its part of the output with the
//                       "beyond trunc" level and prunes residue pairs by bounding sphere (one sphere per
//                       residue for the whole group), compacting the survivors into one shared-memory list.
//                       Survivors are evaluated in the expanded form |b'|^2 - 2a".b' + |a"|^2 with the row
//                       atoms moved into the column residue's frame (a" = a' - D, D = centre_j - centre_i),
//                       as packed FFMA2 product sums and three-input minima, with a rigorous rounding bound E; if every
//                       value in [m-E, m+E] quantises to the same level that level is final, otherwise the
//                       pair is redone with the reference operation order (no FMA contraction) from
//                       the original coordinates, which is also the path for pairs whose atoms may need
//                       different periodic shifts.  Results are therefore bit-identical to the CPU restatement.
//
// The kernel is bound by FP32 / integer instruction issue; see DESIGN.md.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "common.h"

namespace {

constexpr int kTile = 64;          // residues per tile side
constexpr int kMaxFB = 8;          // frames per CTA at most (3 bits in a redo-list entry)
constexpr int kPlanes = 4;         // x' y' z' |a'|^2
constexpr int kPlanesJ = 4, kPlanesI = 3;   // planes staged per tile side: columns need |b'|^2, rows do not
constexpr int kListCap = 4096;     // survivor list entries (one per residue pair of the tile)
constexpr int kSmemLimit = 227 * 1024;   // opt-in shared memory per CTA on sm_100
constexpr float kPruneSlack = 1.00001f;
constexpr int kSphereIters = 16;          // steps of the bounding-sphere refinement in prep_slots_kernel
constexpr float kErrC = 1.25f * 5.9604645e-8f;   // safety factor times the unit roundoff, see fast_pair()

__host__ __device__ inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// ------------------------------------------------------------------------------------------
// prep: slots + bounding spheres
// ------------------------------------------------------------------------------------------
// Inside every block of 64 residues the tile kernel wants the residues ordered by their number of
// atoms: threads of a warp then work on residue pairs of similar size and run the same number of
// loop trips (residue sizes differ by a factor of three in a protein).  perm[position] = residue.
__global__ void block_perm_kernel(const int* __restrict__ res_start, int nres, int* __restrict__ perm) {
    __shared__ int s_cnt[kTile];
    const int l = threadIdx.x, r = blockIdx.x * kTile + l;
    const int n = r < nres ? res_start[r + 1] - res_start[r] : 0x7fffffff;     // padding sorts last
    s_cnt[l] = n;
    __syncthreads();
    int pos = 0;
    for (int k = 0; k < kTile; k++) pos += (s_cnt[k] < n || (s_cnt[k] == n && k < l)) ? 1 : 0;
    perm[blockIdx.x * kTile + pos] = r;
}

// KP > 0: residues have at most KP atoms and a thread keeps its residue in registers (one pass over global
// memory); KP = 0: any size, the atoms are re-read from global memory for every pass.
template <int KP>
__global__ void prep_slots_kernel(const float* __restrict__ xyz, const int* __restrict__ res_start,
                                  const int* __restrict__ perm, int64_t nf, int natoms, int nres, int nres_pad,
                                  int kpad, float* __restrict__ slots, float4* __restrict__ sph,
                                  int* __restrict__ cnt) {
    int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= nf * nres_pad) return;
    int64_t f = gid / nres_pad;
    const int pos = (int)(gid - f * nres_pad);         // position in the padded, per-block sorted order
    const int r = perm[pos];                            // the residue that sits there
    const size_t pstride = (size_t)kpad * nres_pad;     // floats per plane
    float* sl = slots + (size_t)f * kPlanes * pstride + (size_t)pos * kpad;   // [plane][residue][k]
    if (r >= nres) {
        for (int pl = 0; pl < kPlanes; pl++)
            for (int k = 0; k < kpad; k++) sl[(size_t)pl * pstride + k] = 0.f;
        sph[(size_t)f * nres_pad + pos] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (f == 0) cnt[pos] = 0;
        return;
    }
    const int a0 = res_start[r], a1 = res_start[r + 1];
    const int n = a1 - a0;
    const float* x = xyz + ((size_t)f * natoms + a0) * 3;
    constexpr int KR = KP > 0 ? KP : 1;
    float ax[KR], ay[KR], az[KR];
    if (KP > 0) {
#pragma unroll
        for (int k = 0; k < KR; k++) {
            const int kk = k < n ? k : 0;               // slots past the residue repeat its first atom
            ax[k] = x[3 * kk]; ay[k] = x[3 * kk + 1]; az[k] = x[3 * kk + 2];
        }
    }
    float sx = 0.f, sy = 0.f, sz = 0.f;
    if (KP > 0) {
#pragma unroll
        for (int k = 0; k < KR; k++)
            if (k < n) { sx += ax[k]; sy += ay[k]; sz += az[k]; }
    } else {
        for (int k = 0; k < n; k++) { sx += x[3 * k]; sy += x[3 * k + 1]; sz += x[3 * k + 2]; }
    }
    const float inv = 1.0f / (float)n;
    float cx = sx * inv, cy = sy * inv, cz = sz * inv;
    // A tighter bounding sphere than the one around the centroid (Badoiu-Clarkson: a few steps towards the
    // farthest atom with shrinking step length): every 1 % off the radii takes about 1 % of the residue pairs
    // off the survivor lists.  Any centre is valid; the slots are stored relative to it.
#pragma unroll 1
    for (int it = 1; it <= kSphereIters; it++) {
        float best = -1.f, fx = 0.f, fy = 0.f, fz = 0.f;
        if (KP > 0) {
#pragma unroll
            for (int k = 0; k < KR; k++) {              // (repeated slots cannot win against the atom they repeat)
                const float px = ax[k] - cx, py = ay[k] - cy, pz = az[k] - cz;
                const float d2 = fmaf(pz, pz, fmaf(py, py, px * px));
                if (d2 > best) { best = d2; fx = px; fy = py; fz = pz; }
            }
        } else {
            for (int k = 0; k < n; k++) {
                const float px = x[3 * k] - cx, py = x[3 * k + 1] - cy, pz = x[3 * k + 2] - cz;
                const float d2 = fmaf(pz, pz, fmaf(py, py, px * px));
                if (d2 > best) { best = d2; fx = px; fy = py; fz = pz; }
            }
        }
        const float w = 1.0f / (float)(it + 1);
        cx = fmaf(fx, w, cx); cy = fmaf(fy, w, cy); cz = fmaf(fz, w, cz);
    }
    float r2max = 0.f;
    if (KP > 0) {
        float rx[KR], ry[KR], rz[KR], rn[KR];
#pragma unroll
        for (int k = 0; k < KR; k++) {
            const float px = ax[k] - cx, py = ay[k] - cy, pz = az[k] - cz;
            const float n2 = fmaf(pz, pz, fmaf(py, py, px * px));
            r2max = fmaxf(r2max, n2);
            rx[k] = px; ry[k] = py; rz[k] = pz; rn[k] = n2;
        }
#pragma unroll
        for (int k = 0; k < KR; k += 4)
            if (k < kpad) {     // kpad is a multiple of 4 and a residue's record is 16-byte aligned
                *reinterpret_cast<float4*>(sl + k) = make_float4(rx[k], rx[k + 1], rx[k + 2], rx[k + 3]);
                *reinterpret_cast<float4*>(sl + pstride + k) = make_float4(ry[k], ry[k + 1], ry[k + 2], ry[k + 3]);
                *reinterpret_cast<float4*>(sl + 2 * pstride + k) = make_float4(rz[k], rz[k + 1], rz[k + 2], rz[k + 3]);
                *reinterpret_cast<float4*>(sl + 3 * pstride + k) = make_float4(rn[k], rn[k + 1], rn[k + 2], rn[k + 3]);
            }
    } else {
        for (int k = 0; k < kpad; k++) {
            const int kk = k < n ? k : 0;
            const float px = x[3 * kk] - cx, py = x[3 * kk + 1] - cy, pz = x[3 * kk + 2] - cz;
            const float n2 = fmaf(pz, pz, fmaf(py, py, px * px));
            r2max = fmaxf(r2max, n2);
            sl[k] = px; sl[pstride + k] = py; sl[2 * pstride + k] = pz; sl[3 * pstride + k] = n2;
        }
    }
    sph[(size_t)f * nres_pad + pos] = make_float4(cx, cy, cz, sqrtf(r2max) * 1.000001f);
    if (f == 0) cnt[pos] = n;
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
    const int* res_start;    // [nres + 1] (device copy)
    const int* perm;         // [nres_pad] residue at each position (sorted by size inside blocks of 64)
    const float* xyz;        // [nf][natoms][3] original coordinates (exact path)
    const float* box;        // [nf][9] or nullptr
    const float* slots;      // per frame [4][nres_pad][kpad], for the plain-load debug path
    uint16_t* levels;        // [nf][stride]
    int64_t stride;          // elements between frames
    unsigned long long* counters;   // [0] near-boundary results [1] un-pruned residue pairs [2] atom pairs evaluated
    int64_t nf;              // frames of this launch
    int natoms, nres, nres_pad, kpad;
    int fb;                  // frames per CTA
    int symmetric;
    int x0, xres, y0, yres;  // block mode (symmetric: x0=y0=0, xres=yres=nres)
    int tx0, ty0;            // first 64-aligned tile of the x / y residue range
    int ntx, nty, ntiles;    // tile grid
    float trunc, trunc2, invlevel;
    int maxlevel;
    int pbc;
    int use_tma;             // 0: stage tiles with plain loads (debug aid, CONAN_MDMAT_NO_TMA=1)
    int swz;                 // tiles land with a TMA swizzle: the 32-byte one for kpad == 8, the 64-byte one for kpad == 16
    int dbg;                 // timing experiments only (CONAN_MDMAT_DEBUG): 1 no exact path, 2 no fill, 4 no items, 8 no pruning
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

__device__ __forceinline__ float nearest_image(float d, float L) {
    return L > 0.f ? fmaf(-rintf(__fdividef(d, L)), L, d) : d;
}

// r2 exactly as the CPU restatement computes it: every operation rounded, no contraction.
__device__ __forceinline__ float dist2_exact(float dx, float dy, float dz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));   // FMNMX3
    return d;
}

// Packed FP32 pairs (FFMA2): one issue slot for two fused multiply-adds.  ptxas turns a pair built
// from the same scalar twice into the broadcast operand form, so only the row side lives in pairs.
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

// Minimum over one chunk of ICH row-atom slots of residue i against all cj atoms of residue j.
// aI, aJ: plane 0 of the residue's slots ([plane][residue][k], planes pstride floats apart; rows have
// x' y' z', columns x' y' z' |b'|^2).  The row atoms are first moved into the column residue's frame and
// pre-scaled, t = -2 (a' - D) with D = centre_j - centre_i (one FFMA2 per packed pair and coordinate), so
// that the column values are used as they were stored: |a - b|^2 = |b'|^2 + t.b' + |t|^2 / 4.
// Each product-sum is a chain of three fused multiply-adds rounded exactly as the scalar form.
// xi / xj: 0, or 4 where the 32-byte TMA swizzle has exchanged the residue's two 16-byte halves.
// K8: every residue has at most eight slots (kpad == 8), so the column loop is two trips, unrolled.
template <int ICH, bool K8>
__device__ __forceinline__ float chunk_min_fast(const float* __restrict__ aI, const float* __restrict__ aJ,
                                                int pstride, int a0, int xi, int xj, int cj, float Dx2, float Dy2,
                                                float Dz2, float m) {
    f32x2 tx[ICH / 2], ty[ICH / 2], tz[ICH / 2];
    float mu[ICH];
    {
        const f32x2 M2 = pack2(-2.f, -2.f);
        const f32x2 DX = pack2(Dx2, Dx2), DY = pack2(Dy2, Dy2), DZ = pack2(Dz2, Dz2);
        if (ICH == 2) {      // one packed pair: an 8-byte read per plane (the swizzle moves whole 16-byte quarters)
            const int o = a0 ^ xi;
            const float2 qx = *reinterpret_cast<const float2*>(aI + o);
            const float2 qy = *reinterpret_cast<const float2*>(aI + pstride + o);
            const float2 qz = *reinterpret_cast<const float2*>(aI + 2 * pstride + o);
            tx[0] = fma2(pack2(qx.x, qx.y), M2, DX);
            ty[0] = fma2(pack2(qy.x, qy.y), M2, DY);
            tz[0] = fma2(pack2(qz.x, qz.y), M2, DZ);
        }
#pragma unroll
        for (int v = 0; v < ICH / 4; v++) {
            const int o = (a0 + 4 * v) ^ xi;
            const float4 qx = *reinterpret_cast<const float4*>(aI + o);
            const float4 qy = *reinterpret_cast<const float4*>(aI + pstride + o);
            const float4 qz = *reinterpret_cast<const float4*>(aI + 2 * pstride + o);
            tx[2 * v] = fma2(pack2(qx.x, qx.y), M2, DX); tx[2 * v + 1] = fma2(pack2(qx.z, qx.w), M2, DX);
            ty[2 * v] = fma2(pack2(qy.x, qy.y), M2, DY); ty[2 * v + 1] = fma2(pack2(qy.z, qy.w), M2, DY);
            tz[2 * v] = fma2(pack2(qz.x, qz.y), M2, DZ); tz[2 * v + 1] = fma2(pack2(qz.z, qz.w), M2, DZ);
        }
    }
#pragma unroll
    for (int u = 0; u < ICH; u++) mu[u] = 3.0e38f;
    // four column atoms per trip (one 16-byte load per plane), two per three-input minimum; slots past cj
    // repeat atom 0 of the residue (prep_slots_kernel) and kpad is a multiple of 4, so a trip reads valid data
#pragma unroll 1
    for (int b = 0; b < (K8 ? 8 : cj); b += 4) {
        if (K8 && b >= cj) break;       // residues of up to four atoms: one trip
        const int o = b ^ xj;
        const float4 bx = *reinterpret_cast<const float4*>(aJ + o);
        const float4 by = *reinterpret_cast<const float4*>(aJ + pstride + o);
        const float4 bz = *reinterpret_cast<const float4*>(aJ + 2 * pstride + o);
        const float4 bn = *reinterpret_cast<const float4*>(aJ + 3 * pstride + o);
        {
            const f32x2 BX = pack2(bx.x, bx.x), BY = pack2(by.x, by.x), BZ = pack2(bz.x, bz.x), NB = pack2(bn.x, bn.x);
            const f32x2 CX = pack2(bx.y, bx.y), CY = pack2(by.y, by.y), CZ = pack2(bz.y, bz.y), NC = pack2(bn.y, bn.y);
#pragma unroll
            for (int u = 0; u < ICH / 2; u++) {
                const f32x2 vb = fma2(tz[u], BZ, fma2(ty[u], BY, fma2(tx[u], BX, NB)));
                const f32x2 vc = fma2(tz[u], CZ, fma2(ty[u], CY, fma2(tx[u], CX, NC)));
                float b0, b1, c0, c1;
                unpack2(vb, b0, b1);
                unpack2(vc, c0, c1);
                mu[2 * u] = fmin3(mu[2 * u], b0, c0);
                mu[2 * u + 1] = fmin3(mu[2 * u + 1], b1, c1);
            }
        }
        if (K8 || b + 2 < cj) {      // (generic sizes: the second half of the last trip may hold padding only)
            const f32x2 BX = pack2(bx.z, bx.z), BY = pack2(by.z, by.z), BZ = pack2(bz.z, bz.z), NB = pack2(bn.z, bn.z);
            const f32x2 CX = pack2(bx.w, bx.w), CY = pack2(by.w, by.w), CZ = pack2(bz.w, bz.w), NC = pack2(bn.w, bn.w);
#pragma unroll
            for (int u = 0; u < ICH / 2; u++) {
                const f32x2 vb = fma2(tz[u], BZ, fma2(ty[u], BY, fma2(tx[u], BX, NB)));
                const f32x2 vc = fma2(tz[u], CZ, fma2(ty[u], CY, fma2(tx[u], CX, NC)));
                float b0, b1, c0, c1;
                unpack2(vb, b0, b1);
                unpack2(vc, c0, c1);
                mu[2 * u] = fmin3(mu[2 * u], b0, c0);
                mu[2 * u + 1] = fmin3(mu[2 * u + 1], b1, c1);
            }
        }
    }
    // + |a' - D|^2 = |t|^2 / 4 (the scaling is exact), after the minimum: the rounded add is monotonic
#pragma unroll
    for (int u = 0; u < ICH / 2; u++) {
        const f32x2 n2 = fma2(tz[u], tz[u], fma2(ty[u], ty[u], mul2(tx[u], tx[u])));
        float n0, n1;
        unpack2(n2, n0, n1);
        m = fmin3(m, fmaf(n0, 0.25f, mu[2 * u]), fmaf(n1, 0.25f, mu[2 * u + 1]));
    }
    return m;
}

// The exact path: reference operation order on the original coordinates (with the minimum image
// when pbc is on), by a whole warp working on one residue pair: the atom pairs are dealt out to the lanes
// and the minimum is combined with shuffles (a minimum does not depend on the order it is taken in).
// Every lane returns the level; lane 0 counts the near-boundary result.
__device__ __forceinline__ int exact_pair_warp(const MdmatParams& prm, int64_t f, int gi, int gj, float Lx, float Ly,
                                               float Lz, int lane, int* near_counter) {
    const int a0 = prm.res_start[gi], na = prm.res_start[gi + 1] - a0;
    const int b0 = prm.res_start[gj], nb = prm.res_start[gj + 1] - b0;
    const float* x = prm.xyz + (size_t)f * prm.natoms * 3;
    float m = prm.trunc2;
    const float rnb = __frcp_rn((float)nb);
    for (int idx = lane; idx < na * nb; idx += 32) {
        // idx / nb without the integer division: exact for idx < 2^14 (na, nb <= 128), because
        // (idx + 0.5) / nb stays 0.5 / nb away from an integer and the float error is below that
        const int q = (int)(((float)idx + 0.5f) * rnb);
        const int a = a0 + q, b = b0 + idx - q * nb;
        float dx = x[3 * a] - x[3 * b], dy = x[3 * a + 1] - x[3 * b + 1], dz = x[3 * a + 2] - x[3 * b + 2];
        if (prm.pbc) { dx = min_image(dx, Lx); dy = min_image(dy, Ly); dz = min_image(dz, Lz); }
        m = fminf(m, dist2_exact(dx, dy, dz));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
    const float r = __fsqrt_rn(m);
    const float tq = __fmul_rn(r, prm.invlevel);
    int c = __float2int_rn(tq);
    c = max(0, min(c, prm.maxlevel));
    const float frac = fabsf(tq - rintf(tq));
    if (lane == 0 && fabsf(frac - 0.5f) <= tq * 2.4e-7f) atomicAdd(near_counter, 1);
    return c;
}

__device__ __forceinline__ int atom_add_shared(int* addr, int v) {
    // plain shared-memory atomic (inline PTX keeps the compiler from wrapping it in its own
    // warp-aggregation sequence; the callers already aggregate per warp)
    int old;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(smem_u32(addr)), "r"(v) : "memory");
    return old;
}

// One survivor item: residue pair (i, j) of the tile in frame g of the group.  Returns the level,
// or -1 if the expanded form cannot certify it (the pair then takes the exact path).
// aIp / aJp: plane 0 of residue i / j of the frame's row / column tile.
template <bool PBC, bool K8>
__device__ __forceinline__ int fast_pair(const MdmatParams& prm, const float* __restrict__ aIp,
                                         const float* __restrict__ aJp, int pstride, int xi, int xj, float4 si,
                                         float4 sj, int ci, int cj, const float* __restrict__ box) {
    float Dx = sj.x - si.x, Dy = sj.y - si.y, Dz = sj.z - si.z;
    // box = the three box lengths of the frame and their reciprocals.  Only pairs whose spheres stay inside half a box come here, so
    // the centre difference is well clear of +-L/2 and the nearest-image shift can be taken by rounding.
    bool shifted = false;
    if (PBC) {
        const float Rx = Dx, Ry = Dy, Rz = Dz;
        Dx = fmaf(-rintf(Dx * box[3]), box[0], Dx);     // box[3..5] = 1 / L
        Dy = fmaf(-rintf(Dy * box[4]), box[1], Dy);
        Dz = fmaf(-rintf(Dz * box[5]), box[2], Dz);
        shifted = Dx != Rx || Dy != Ry || Dz != Rz;
    }
    const float rr = si.w + sj.w;
    const float D2 = fmaf(Dz, Dz, fmaf(Dy, Dy, Dx * Dx));
    // Error of the expanded form against the reference's r2, term by term (u = 2^-24; Ri, Rj the sphere
    // radii, A = Ri + |D| >= |a' - D|, M = A + Rj >= every difference vector):
    //   positions: a', b', D and a' - D each carry u per component, the difference vector is off by
    //              <= 2 u M and its square by <= 4 u M sqrt(r2);
    //   the sum:   three rounded steps on partial sums <= Rj (Rj + 2 A), three on |b'|^2 <= Rj^2, three on
    //              |a' - D|^2 <= A^2, the last add on r2 itself: <= 3 u (M^2 + Rj^2) + u r2;
    //   reference: its own five roundings, <= 5 u r2.
    // E = 1.25 u (3 (M^2 + Rj^2) + 4 M sqrt(r2) + 6 r2), evaluated once the minimum is known.
    const float dn = D2 * rsqrtf(fmaxf(D2, 1.0e-20f));
    const float M = rr + dn;
    const float M2R = fmaf(M, M, sj.w * sj.w);
    // a pair that is brought together by a lattice shift: the reference forms a - b (of the size of the box)
    // before shifting, and so does the centre difference here: an extra u * L per component
    float Ebox = 0.f;
    if (PBC && shifted) Ebox = 2.5e-7f * fmaxf(box[0], fmaxf(box[1], box[2])) * (rr + dn * 1.001f);
    const float Dx2 = Dx + Dx, Dy2 = Dy + Dy, Dz2 = Dz + Dz;
    float m = 3.0e38f;
    if (K8) {
        if (ci > 4) m = chunk_min_fast<8, true>(aIp, aJp, pstride, 0, xi, xj, cj, Dx2, Dy2, Dz2, m);
        else m = chunk_min_fast<4, true>(aIp, aJp, pstride, 0, xi, xj, cj, Dx2, Dy2, Dz2, m);
    } else {
        // row atoms in chunks of 8, then what is left in a chunk of 8 (7-8 atoms), 4 (+ 2) or 2
        int a0 = 0;
        while (ci - a0 > 6) { m = chunk_min_fast<8, false>(aIp, aJp, pstride, a0, xi, xj, cj, Dx2, Dy2, Dz2, m); a0 += 8; }
        if (ci - a0 > 2) { m = chunk_min_fast<4, false>(aIp, aJp, pstride, a0, xi, xj, cj, Dx2, Dy2, Dz2, m); a0 += 4; }
        if (ci - a0 > 0) m = chunk_min_fast<2, false>(aIp, aJp, pstride, a0, xi, xj, cj, Dx2, Dy2, Dz2, m);
    }
    const float mm = fmaxf(m, 0.f);
    // sqrt through the reciprocal square root (2^-22.9 relative) instead of the IEEE sequence; the
    // certification margin below carries the difference
    float rs;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(fmaxf(mm, 1.0e-7f)));
    float sq = mm * rs;
    sq = fmaf(fmaf(-sq, sq, mm), 0.5f * rs, sq);    // one Newton step: within an ulp of the square root
    const float E = fmaf(kErrC, fmaf(4.f * M, sq, fmaf(3.f, M2R, 6.f * mm)), Ebox);
    if (m - E >= prm.trunc2 * 1.000001f) return prm.maxlevel;   // certainly beyond trunc
    if (mm < 1.0e-7f) return -1;
    const float tq = sq * prm.invlevel;
    const float cb = tq + 12582912.f;               // 1.5 * 2^23: rounds tq to the nearest integer (ties to even)
    const float c = cb - 12582912.f;
    // |sqrt(e) - sqrt(m)| = |e - m| / (sqrt(e) + sqrt(m)) <= 0.502 E / sqrt(m) for every e in [m-E, m+E]
    // while E < m / 100 (E / sqrt(m) otherwise); the level is certain when tq stays clear of the rounding
    // boundary (x.5) by that much plus the rounding of sqrt and mul here (1.3e-7 relative) and in the
    // reference (1.2e-7)
    const float half = E < 0.01f * mm ? 0.51f : 1.01f;
    const float dt = fmaf(E * prm.invlevel * half, rs, fmaf(tq, 3.0e-7f, 1.0e-6f));
    if (fabsf(fabsf(tq - c) - 0.5f) > dt) return min(__float_as_int(cb) - 0x4B400000, prm.maxlevel);
    return -1;
}

// NT threads per CTA, CPS CTAs per SM (see cta_shape).
template <bool PBC, int NT, int CPS>
__global__ void __launch_bounds__(NT, CPS)
mdmat_tile_kernel(const __grid_constant__ CUtensorMap tmapI, const __grid_constant__ CUtensorMap tmapJ,
                  const MdmatParams prm) {
    constexpr int kThreads = NT, kWarps = NT / 32;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int kpad = prm.kpad, fbmax = prm.fb;
    const int pstride = kpad * kTile;                 // floats between planes of one frame's tile
    float* aI = reinterpret_cast<float*>(smem_raw);                       // [fb][3][64][kpad]
    float* aJ = aI + (size_t)fbmax * kPlanesI * pstride;                  // [fb][4][64][kpad]
    float4* sphI = reinterpret_cast<float4*>(aJ + (size_t)fbmax * kPlanesJ * pstride);   // [fb][64]
    float4* sphJ = sphI + fbmax * kTile;
    float4* grpI = sphJ + fbmax * kTile;              // [64] sphere around the residue over the whole group
    float4* grpJ = grpI + kTile;
    int* cntI = reinterpret_cast<int*>(grpJ + kTile);
    int* cntJ = cntI + kTile;
    uint16_t* list = reinterpret_cast<uint16_t*>(cntJ + kTile);          // [kListCap]
    uint64_t* bar = reinterpret_cast<uint64_t*>(list + kListCap);
    int* counters = reinterpret_cast<int*>(bar + 1);   // [0] list [2] near [3] evals
    int* rowbase = counters + 6;                        // [64] element index of pair (row residue, residue 0) in a frame row
    int* origI = rowbase + kTile;                       // [64] residue at row position i / column position j of the tile
    int* origJ = origI + kTile;
    float* boxs = reinterpret_cast<float*>(origJ + kTile);   // [kMaxFB][6] box lengths of the group's frames and their reciprocals
    uint16_t** rowptr = reinterpret_cast<uint16_t**>(boxs + kMaxFB * 6);   // [kMaxFB] the group's frames in the level stack

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = (int)blockIdx.x;                                   // tile
    const int64_t f0 = (int64_t)blockIdx.y * fbmax;                  // first frame of the group
    const int nfb = (int)min((int64_t)fbmax, prm.nf - f0);
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
    // Tiles are aligned to multiples of kTile in the index group's residue numbering, so TMA
    // coordinates are always tile-aligned and inside the padded tensor.
    const bool diag = prm.symmetric && ti == tj;
    const int gi0 = (prm.tx0 + ti) * kTile, gj0 = (prm.ty0 + tj) * kTile;
    const int i_beg = prm.x0, i_end = prm.x0 + prm.xres;
    const int j_beg = prm.y0, j_end = prm.y0 + prm.yres;
    const int nres = prm.nres, yres = prm.yres, x0 = prm.x0, y0 = prm.y0, maxlevel = prm.maxlevel;
    const bool symmetric = prm.symmetric != 0;
    // element index of pair (gi, gj) inside a frame row (P < 2^31 for every supported size)
    auto pair_idx = [&](int gi, int gj) -> int {
        return symmetric ? gi * nres - ((gi * (gi + 1)) >> 1) + (gj - gi - 1) : (gi - x0) * yres + (gj - y0);
    };

    if (tid == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
        for (int c = 0; c < 5; c++) counters[c] = 0;
    }
    if (tid < kMaxFB) rowptr[tid] = prm.levels + (size_t)(f0 + tid) * prm.stride;
    __syncthreads();
    if (prm.use_tma) {
        if (tid == 0) {
            const uint32_t bytes = (uint32_t)fbmax * (kPlanesI + kPlanesJ) * (uint32_t)pstride * 4u;
            mbar_expect_tx(bar, bytes);
            tma_load_4d(aI, &tmapI, bar, 0, gi0, 0, (int)f0);
            tma_load_4d(aJ, &tmapJ, bar, 0, gj0, 0, (int)f0);
        }
    } else {
        for (int g = 0; g < nfb; g++) {
            const float* src = prm.slots + (size_t)(f0 + g) * kPlanes * kpad * prm.nres_pad;
            for (int idx = tid; idx < kPlanesI * pstride; idx += kThreads) {
                const int pl = idx / pstride, rk = idx - pl * pstride;   // rk = residue * kpad + k
                aI[(size_t)g * kPlanesI * pstride + idx] = src[(size_t)pl * kpad * prm.nres_pad + (size_t)gi0 * kpad + rk];
            }
            for (int idx = tid; idx < kPlanesJ * pstride; idx += kThreads) {
                const int pl = idx / pstride, rk = idx - pl * pstride;
                aJ[(size_t)g * kPlanesJ * pstride + idx] = src[(size_t)pl * kpad * prm.nres_pad + (size_t)gj0 * kpad + rk];
            }
        }
    }
    // per-frame spheres, and one sphere per residue that holds it in every frame of the group, so the
    // pruning runs once per group instead of once per frame.  In a periodic run the residue's centre in
    // frame g is its centre in the group's first frame plus the minimum image of the displacement, modulo
    // that frame's box, which is all the torus triangle inequality of the pruning test needs.
    if (PBC && tid < nfb * 3) {
        const int q = tid, g = q / 3, c = q % 3;
        const float L = prm.box[9 * (f0 + g) + 4 * c];
        boxs[6 * g + c] = L;
        boxs[6 * g + 3 + c] = L > 0.f ? 1.0f / L : 0.f;
    }
    if (tid < 2 * kTile) {
        const bool isI = tid < kTile;
        const int l = isI ? tid : tid - kTile;
        const int gr = (isI ? gi0 : gj0) + l;
        float4* dst = isI ? sphI : sphJ;
        const float4* src = prm.sph + (size_t)f0 * prm.nres_pad + gr;
        const float4 s0 = src[0];
        dst[l] = s0;
        // displacements of the centre since the group's first frame, and the group sphere around their mean
        float ux[kMaxFB], uy[kMaxFB], uz[kMaxFB], rw[kMaxFB];
        float mx = 0.f, my = 0.f, mz = 0.f;
        ux[0] = uy[0] = uz[0] = 0.f; rw[0] = s0.w;
#pragma unroll
        for (int g = 1; g < kMaxFB; g++) {
            ux[g] = uy[g] = uz[g] = 0.f; rw[g] = 0.f;
            if (g < nfb) {
                const float4 sg = src[(size_t)g * prm.nres_pad];
                dst[g * kTile + l] = sg;
                float dx = sg.x - s0.x, dy = sg.y - s0.y, dz = sg.z - s0.z;
                if (PBC) {
                    const float* b = prm.box + 9 * (f0 + g);
                    dx = nearest_image(dx, b[0]); dy = nearest_image(dy, b[4]); dz = nearest_image(dz, b[8]);
                }
                ux[g] = dx; uy[g] = dy; uz[g] = dz; rw[g] = sg.w;
                mx += dx; my += dy; mz += dz;
            }
        }
        const float invn = 1.0f / (float)nfb;
        mx *= invn; my *= invn; mz *= invn;
        float R = 0.f;
#pragma unroll
        for (int g = 0; g < kMaxFB; g++)
            if (g < nfb) {
                const float dx = ux[g] - mx, dy = uy[g] - my, dz = uz[g] - mz;
                R = fmaxf(R, sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx))) * 1.000001f + 1.0e-6f + rw[g]);
            }
        (isI ? grpI : grpJ)[l] = make_float4(s0.x + mx, s0.y + my, s0.z + mz, R);
        (isI ? cntI : cntJ)[l] = prm.cnt[gr];
        const int res = prm.perm[gr];
        (isI ? origI : origJ)[l] = res;
        if (isI) rowbase[l] = pair_idx(res, 0);          // + column residue = the pair's element index
    }

    // ---- fill: every pair of the tile starts at the "beyond trunc" level -------------------
    if (!(prm.dbg & 2)) {
        // a quarter of a warp per row: 8 lanes x 8 levels (one 16-byte store) cover the 64 columns
        const int sub = lane >> 3, l8 = lane & 7;
        const uint32_t fillv = (uint32_t)maxlevel * 0x00010001u;
        const int jlo0 = max(0, j_beg - gj0), jhi = min(kTile, j_end - gj0);
        const uint16_t ml = (uint16_t)maxlevel;
        for (int i = warp * 4 + sub; i < kTile; i += 4 * kWarps) {
            const int gi = gi0 + i;
            if (gi < i_beg || gi >= i_end) continue;
            const int jlo = diag ? max(jlo0, i + 1) : jlo0;
            const int n = jhi - jlo;
            if (n <= 0) continue;
            uint16_t* p = prm.levels + (size_t)f0 * prm.stride + pair_idx(gi, gj0) + jlo;
            // 16-byte aligned body, scalar head and tail (up to seven elements each); with a row pitch that is a
            // multiple of eight elements the split is the same in every frame of the group
            int mis = (int)((reinterpret_cast<uintptr_t>(p) >> 1) & 7);   // elements past alignment
            int head = min((8 - mis) & 7, n);
            int nq = (n - head) >> 3;
            int tail0 = head + 8 * nq;
            const bool same_split = (prm.stride & 7) == 0;
            for (int g = 0; g < nfb; g++, p += prm.stride) {
                if (!same_split && g > 0) {
                    mis = (int)((reinterpret_cast<uintptr_t>(p) >> 1) & 7);
                    head = min((8 - mis) & 7, n);
                    nq = (n - head) >> 3;
                    tail0 = head + 8 * nq;
                }
                if (l8 < head) p[l8] = ml;
                if (l8 < nq) *reinterpret_cast<uint4*>(p + head + 8 * l8) = make_uint4(fillv, fillv, fillv, fillv);
                if (l8 < n - tail0) p[tail0 + l8] = ml;
            }
        }
    }
    // spheres, counts and the fill are in place; the same barrier tells whether every residue of both tile sides has
    // exactly eight atoms (padding positions have none and never enter the list): their items then run a copy of the
    // atom loop with the counts as constants
    int mycnt = 8;                 // (threads 0..127 wrote the counts of their own positions above)
    if (tid < kTile) mycnt = cntI[tid];
    else if (tid < 2 * kTile) mycnt = cntJ[tid - kTile];
    const bool full8 = __syncthreads_and(mycnt == 8 || mycnt == 0) != 0;

    bool tiles_ready = !prm.use_tma;

    // ---- survivors: fast expanded form, exact redo where the level is not certain ----------
    // The list holds residue pairs; an entry stands for the pair in every frame of the group
    // (items = entries x frames).
    auto process_list = [&]() {
        if (!tiles_ready) { mbar_wait(bar, 0); tiles_ready = true; }
        const int nent = counters[0];
        const int nitems = (prm.dbg & 4) ? 0 : nent * nfb;
        const int swz = prm.swz ? 4 : 0;     // 32-byte swizzle: residues 4..7 of every eight have their halves exchanged
        const int swz16 = (prm.swz && kpad == 16) ? 12 : 0;
        const bool k8 = kpad == 8;
        int gq = 0, qq = tid;      // frame-major walk over (frame, entry): a warp's lanes share the frame,
                                   // so their tile reads fall into different banks
        // whole warps walk the items together (the exact path below is taken by all 32 lanes)
        for (int s0 = warp * 32; s0 < nitems; s0 += kThreads, qq += kThreads) {
            const bool have = s0 + lane < nitems;
            while (qq >= nent && gq < nfb - 1) { qq -= nent; gq++; }
            const int g = gq;
            const int e = have ? list[qq] : 0;
            const int i = e & 63, j = (e >> 6) & 63;
            int level = maxlevel;
            if (have) {
                const int ci = cntI[i], cj = cntJ[j];
                level = -1;
                if (!PBC || !(e & 0x1000)) {
                    const float* ri = aI + g * (kPlanesI * pstride) + i * kpad;
                    const float* rj = aJ + g * (kPlanesJ * pstride) + j * kpad;
                    if (k8 && full8)
                        level = fast_pair<PBC, true>(prm, ri, rj, 8 * kTile, swz & i, swz & j, sphI[g * kTile + i],
                                                     sphJ[g * kTile + j], 8, 8, PBC ? boxs + 6 * g : nullptr);
                    else if (k8)
                        level = fast_pair<PBC, true>(prm, ri, rj, 8 * kTile, swz & i, swz & j, sphI[g * kTile + i],
                                                     sphJ[g * kTile + j], ci, cj, PBC ? boxs + 6 * g : nullptr);
                    else   // 64-byte swizzle (kpad == 16): the four 16-byte quarters of a residue's row are permuted by (row / 2) % 4
                        level = fast_pair<PBC, false>(prm, ri, rj, pstride, swz16 & (i << 1), swz16 & (j << 1), sphI[g * kTile + i],
                                                      sphJ[g * kTile + j], ci, cj, PBC ? boxs + 6 * g : nullptr);
                }
            }
            // The pairs the expanded form could not certify are redone right here, one after the other, each
            // by the whole warp with the reference's operation order (no list, no block barrier).
            unsigned need = __ballot_sync(0xffffffffu, level < 0);
            if (prm.dbg & 1) need = 0;
            while (need) {
                const int src = __ffs(need) - 1;
                need &= need - 1;
                const int es = __shfl_sync(0xffffffffu, e, src), gs = __shfl_sync(0xffffffffu, g, src);
                const int is = es & 63, js = (es >> 6) & 63;
                float Lx = 0.f, Ly = 0.f, Lz = 0.f;
                if (PBC) {
                    // A pair whose spheres stay inside half a box and whose centres need no lattice shift has every
                    // atom difference inside (-L/2, L/2): the minimum image leaves it alone, so it is not applied.
                    const float4 a = sphI[gs * kTile + is], b = sphJ[gs * kTile + js];
                    const float* bl = boxs + 6 * gs;
                    const float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
                    const bool image = (es & 0x1000) || nearest_image(dx, bl[0]) != dx || nearest_image(dy, bl[1]) != dy ||
                                       nearest_image(dz, bl[2]) != dz;
                    if (image) { Lx = bl[0]; Ly = bl[1]; Lz = bl[2]; }
                }
                const int lv = exact_pair_warp(prm, f0 + gs, origI[is], origJ[js], Lx, Ly, Lz, lane, &counters[2]);
                if (lane == src) level = lv;
            }
            if (level < maxlevel)
                rowptr[g][rowbase[i] + origJ[j]] = (uint16_t)level;
        }
        __syncthreads();
    };

    // ---- bounding-sphere pruning, survivors compacted into the list -------------------------
    const float half_trunc = 0.5f * prm.trunc;
    {
        const float4* sI = grpI;
        const float4* sJ = grpJ;
        // periodic runs: the smallest box of the group (for the half-box test) and whether all boxes are equal
        float Lx = 0.f, Ly = 0.f, Lz = 0.f;
        bool same_box = true;
        if (PBC) {
            Lx = boxs[0]; Ly = boxs[1]; Lz = boxs[2];
            for (int q = 1; q < nfb; q++) {
                same_box = same_box && boxs[6 * q] == boxs[0] && boxs[6 * q + 1] == boxs[1] && boxs[6 * q + 2] == boxs[2];
                Lx = fminf(Lx, boxs[6 * q]); Ly = fminf(Ly, boxs[6 * q + 1]); Lz = fminf(Lz, boxs[6 * q + 2]);
            }
        }
        const float iLx = Lx > 0.f ? 1.0f / Lx : 0.f, iLy = Ly > 0.f ? 1.0f / Ly : 0.f, iLz = Lz > 0.f ? 1.0f / Lz : 0.f;
        // The tile is taken in 16 x 16 sub-blocks, two per warp.  Rows and columns are sorted by residue
        // size (block_perm_kernel), so the pairs of a sub-block have similar numbers of atoms on both
        // sides; a sub-block's survivors go into the list as one contiguous run (one reservation), and the
        // 32 lanes that later take 32 consecutive entries run the same number of loop trips.
#pragma unroll 1
        for (int sb = warp; sb < ((prm.dbg & 8) ? 0 : 16); sb += kWarps) {
            const int rb = (sb >> 2) * 16, cb = (sb & 3) * 16;
            const int j = cb + (lane & 15);
            const float4 sj = sJ[j];
            const float rjp = (sj.w + half_trunc) * kPruneSlack;   // the j half of the pruning distance
            const int jres = origJ[j];
            const bool jok = jres >= j_beg && jres < j_end;
            unsigned keepbits = 0, genbits = 0, total = 0;
            int evals = 0;               // atom pairs of the kept residue pairs (per frame)
            const int cjn = cntJ[j];
            unsigned before[8];          // survivors of the sub-block ahead of this lane's k-th pair
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int i = rb + (lane >> 4) + 2 * k;
                const float4 si = sI[i];
                const int ires = origI[i];
                const bool iok = ires >= i_beg && ires < i_end;
                float dx = si.x - sj.x, dy = si.y - sj.y, dz = si.z - sj.z;
                bool general = false;
                if (PBC) {
                    // the distance of the group centres on the torus: per component the smallest minimum
                    // image over the frames' boxes (one evaluation when the box does not change)
                    // (rounding to the nearest image instead of gmx's loop: the pruning distance carries a
                    // relative slack of 1e-5, far above the difference)
                    if (same_box) {
                        dx = fmaf(-rintf(dx * iLx), Lx, dx); dy = fmaf(-rintf(dy * iLy), Ly, dy); dz = fmaf(-rintf(dz * iLz), Lz, dz);
                    } else {
                        float mx = 3.0e38f, my = 3.0e38f, mz = 3.0e38f;
                        for (int q = 0; q < nfb; q++) {
                            mx = fminf(mx, fabsf(nearest_image(dx, boxs[6 * q])));
                            my = fminf(my, fabsf(nearest_image(dy, boxs[6 * q + 1])));
                            mz = fminf(mz, fabsf(nearest_image(dz, boxs[6 * q + 2])));
                        }
                        dx = mx; dy = my; dz = mz;
                    }
                    // one lattice shift serves every atom pair only if the spheres stay inside half a box
                    const float rr = si.w + sj.w;
                    general = !(fabsf(dx) + rr < 0.49999f * Lx && fabsf(dy) + rr < 0.49999f * Ly &&
                                fabsf(dz) + rr < 0.49999f * Lz);
                }
                const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                const float T = (si.w + half_trunc) * kPruneSlack + rjp;
                const bool keep = iok && jok && (!diag || jres > ires) && !(d2 > T * T);
                const unsigned ball = __ballot_sync(0xffffffffu, keep);
                before[k] = total + __popc(ball & ((1u << lane) - 1));
                total += __popc(ball);
                keepbits |= (keep ? 1u : 0u) << k;
                genbits |= (general ? 1u : 0u) << k;
                evals += keep ? cntI[i] * cjn : 0;
            }
            if (prm.counters) {
                evals = __reduce_add_sync(0xffffffffu, evals);
                if (lane == 0 && evals) atom_add_shared(&counters[3], evals);
            }
            if (total) {
                int base = 0;
                if (lane == 0) base = atom_add_shared(&counters[0], (int)total);
                base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if ((keepbits >> k) & 1u) {
                        const int i = rb + (lane >> 4) + 2 * k;
                        list[base + before[k]] = (uint16_t)(i | (j << 6) | (((genbits >> k) & 1u) ? 0x1000 : 0));
                    }
            }
        }
        __syncthreads();
    }
    process_list();

    if (tid == 0 && prm.counters) {
        if (counters[2]) atomicAdd(&prm.counters[0], (unsigned long long)counters[2]);
        if (counters[0]) atomicAdd(&prm.counters[1], (unsigned long long)(counters[0] * nfb));
        if (counters[3]) atomicAdd(&prm.counters[2], (unsigned long long)counters[3] * (unsigned long long)nfb);
    }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
struct Layout {
    int nres_pad, kpad;
    size_t off_res_start, off_cnt, off_perm, off_sph, off_slots, total;
};

inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

bool make_layout(int64_t nf, int nres, int max_res_atoms, Layout* L) {
    if (nf <= 0 || nres <= 0 || max_res_atoms <= 0) return false;
    L->nres_pad = round_up(nres, kTile);
    L->kpad = round_up(max_res_atoms, 4);
    size_t o = 0;
    L->off_res_start = o; o = align256(o + sizeof(int) * (size_t)(nres + 1));
    L->off_cnt = o;       o = align256(o + sizeof(int) * (size_t)L->nres_pad);
    L->off_perm = o;      o = align256(o + sizeof(int) * (size_t)L->nres_pad);
    L->off_sph = o;       o = align256(o + sizeof(float4) * (size_t)nf * L->nres_pad);
    L->off_slots = o;     o = align256(o + sizeof(float) * (size_t)nf * kPlanes * L->kpad * L->nres_pad);
    L->total = o;
    return true;
}

size_t mdmat_smem_bytes(int kpad, int fb) {
    size_t s = 0;
    s += (size_t)fb * (kPlanesI + kPlanesJ) * kpad * kTile * sizeof(float);   // atom tiles, rows and columns
    s += (size_t)(fb + 1) * 2 * kTile * sizeof(float4);         // per-frame and per-group spheres
    s += 2 * kTile * sizeof(int);                               // counts
    s += (size_t)kListCap * sizeof(uint16_t);                   // survivor list
    s += 8 + 6 * sizeof(int) + 3 * kTile * sizeof(int);         // mbarrier + counters + row bases + residue numbers
    s += kMaxFB * 6 * sizeof(float);                            // box lengths of the group's frames and reciprocals
    s += kMaxFB * sizeof(void*);                                // the group's rows of the level stack
    return (s + 15) & ~(size_t)15;
}

// CTA shapes.  0: 256 threads x 2 CTAs per SM, up to 128 registers per thread and 112 KB of shared memory
// (six frames per CTA at eight atoms per residue); 1: 256 threads x 3 CTAs per SM, 80 registers and 74 KB
// (three frames).  The first is faster (18.0 against 20.0 ms per 12 500 frames of 1000 residues): the tile
// kernel does not fit 80 registers without spilling in its item loop, and twice the frames per CTA halve
// the set-up and pruning work per frame.  More, thinner warps do not pay either: 320 threads x 2 (96
// registers) 19.2 ms, 384 x 2 (80 registers) 20.7 ms, 128 x 4 (128 registers, two frames) 20.8 ms.
struct CtaShape { int threads, per_sm; size_t budget; };
inline CtaShape cta_shape(int cfg) {
    if (cfg == 1) return CtaShape{256, 3, (size_t)74 * 1024};
    return CtaShape{256, 2, (size_t)112 * 1024};
}

int frames_per_cta(int kpad, size_t budget) {
    int fb = kMaxFB;
    while (fb > 1 && mdmat_smem_bytes(kpad, fb) > budget) fb--;
    return fb;
}

// Largest residue (in atom slots) whose single-frame tiles still fit the shared memory of one CTA.
int max_kpad() {
    int k = 4;
    while (mdmat_smem_bytes(k + 4, 1) <= (size_t)kSmemLimit) k += 4;
    return k;
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

constexpr int kDefaultCfg = 0;        // CTA shape used unless CONAN_MDMAT_CFG says otherwise (see cta_shape)
constexpr bool kDefaultSwz = true;

// The two tensor maps of one launch.  Encoding costs a few microseconds of host time; a caller that
// runs the same shapes on the same workspace call after call (a step loop) gets the previous pair back.
struct TileMaps {
    CUtensorMap I, J;
    const void* base = nullptr;
    int kpad = 0, nres_pad = 0, fb = 0, swz = 0;
    int64_t nfl = 0;
};

const TileMaps* tile_maps(EncodeTiledFn encode, float* base, int kpad, int nres_pad, int64_t nfl, int fb, int swz) {
    static thread_local TileMaps tm;
    if (tm.base == base && tm.kpad == kpad && tm.nres_pad == nres_pad && tm.nfl == nfl && tm.fb == fb && tm.swz == swz)
        return &tm;
    const cuuint64_t plane_bytes = (cuuint64_t)kpad * nres_pad * 4;
    cuuint64_t gdim[4] = {(cuuint64_t)kpad, (cuuint64_t)nres_pad, (cuuint64_t)kPlanes, (cuuint64_t)nfl};
    cuuint64_t gstr[3] = {(cuuint64_t)kpad * 4, plane_bytes, plane_bytes * kPlanes};
    cuuint32_t boxI[4] = {(cuuint32_t)kpad, (cuuint32_t)kTile, (cuuint32_t)kPlanesI, (cuuint32_t)fb};
    cuuint32_t boxJ[4] = {(cuuint32_t)kpad, (cuuint32_t)kTile, (cuuint32_t)kPlanesJ, (cuuint32_t)fb};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUtensorMapSwizzle sw = !swz ? CU_TENSOR_MAP_SWIZZLE_NONE : (kpad == 16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    tm.base = nullptr;
    CUresult cr = encode(&tm.I, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, gdim, gstr, boxI, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr == CUDA_SUCCESS)
        cr = encode(&tm.J, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, gdim, gstr, boxJ, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return nullptr;
    tm.base = base; tm.kpad = kpad; tm.nres_pad = nres_pad; tm.nfl = nfl; tm.fb = fb; tm.swz = swz;
    return &tm;
}

// Once per device: raise the dynamic shared-memory limit of every tile-kernel instantiation and make the
// runtime load all kernels of this file, so that later calls (the allocation-free entry point in
// particular) neither load code nor change function attributes inside a caller's step loop.
constexpr int kMaxDevices = 64;
bool g_device_ready[kMaxDevices] = {false};
std::mutex g_device_mutex;

int device_setup() {
    int dev = -1;
    CONAN_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_device_mutex);
    if (g_device_ready[dev]) return CONAN_OK;
    CONAN_CUDA_TRY(cudaFuncSetAttribute(mdmat_tile_kernel<false, 256, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    CONAN_CUDA_TRY(cudaFuncSetAttribute(mdmat_tile_kernel<true, 256, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    CONAN_CUDA_TRY(cudaFuncSetAttribute(mdmat_tile_kernel<false, 256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    CONAN_CUDA_TRY(cudaFuncSetAttribute(mdmat_tile_kernel<true, 256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit));
    cudaFuncAttributes fa;
    CONAN_CUDA_TRY(cudaFuncGetAttributes(&fa, block_perm_kernel));
    CONAN_CUDA_TRY(cudaFuncGetAttributes(&fa, prep_slots_kernel<8>));
    CONAN_CUDA_TRY(cudaFuncGetAttributes(&fa, prep_slots_kernel<16>));
    CONAN_CUDA_TRY(cudaFuncGetAttributes(&fa, prep_slots_kernel<0>));
    g_device_ready[dev] = true;
    return CONAN_OK;
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
    if (L.kpad > max_kpad())
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: a residue has %d atoms; the limit is %d (the tiles of "
                           "one frame must fit %d KB of shared memory)", max_atoms, max_kpad(), kSmemLimit / 1024);
    if (work_bytes < L.total)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: workspace %zu bytes, need %zu", work_bytes, L.total);
    if ((reinterpret_cast<uintptr_t>(d_work) & 255) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: workspace must be 256-byte aligned");

    char* w = static_cast<char*>(d_work);
    int* d_res_start = reinterpret_cast<int*>(w + L.off_res_start);
    int* d_cnt = reinterpret_cast<int*>(w + L.off_cnt);
    int* d_perm = reinterpret_cast<int*>(w + L.off_perm);
    float4* d_sph = reinterpret_cast<float4*>(w + L.off_sph);
    float* d_slots = reinterpret_cast<float*>(w + L.off_slots);

    if (phases & 1) {
        CONAN_CUDA_TRY(cudaMemcpyAsync(d_res_start, h_res_start, sizeof(int) * (size_t)(nres + 1),
                                       cudaMemcpyHostToDevice, stream));
        block_perm_kernel<<<(unsigned)(L.nres_pad / kTile), kTile, 0, stream>>>(d_res_start, nres, d_perm);
        const int64_t total = nf * L.nres_pad;
        const int threads = 128;
        const unsigned blocks = (unsigned)((total + threads - 1) / threads);
        if (L.kpad <= 8)
            prep_slots_kernel<8><<<blocks, threads, 0, stream>>>(d_xyz, d_res_start, d_perm, nf, natoms, nres, L.nres_pad,
                                                                 L.kpad, d_slots, d_sph, d_cnt);
        else if (L.kpad <= 16)
            prep_slots_kernel<16><<<blocks, threads, 0, stream>>>(d_xyz, d_res_start, d_perm, nf, natoms, nres, L.nres_pad,
                                                                  L.kpad, d_slots, d_sph, d_cnt);
        else
            prep_slots_kernel<0><<<blocks, threads, 0, stream>>>(d_xyz, d_res_start, d_perm, nf, natoms, nres, L.nres_pad,
                                                                 L.kpad, d_slots, d_sph, d_cnt);
        CONAN_CUDA_TRY(cudaGetLastError());
    }
    if (!(phases & 2)) return CONAN_OK;

    EncodeTiledFn encode = get_encode_fn();
    if (!encode) return conan::fail(CONAN_ERR_CUDA, "conan_mdmat_frames: cuTensorMapEncodeTiled unavailable");
    int rc = device_setup();
    if (rc != CONAN_OK) return rc;

    int cfg = kDefaultCfg;
    {
        const char* e = getenv("CONAN_MDMAT_CFG");
        if (e && (e[0] == '0' || e[0] == '1')) cfg = e[0] - '0';
    }
    CtaShape shape = cta_shape(cfg);
    if (mdmat_smem_bytes(L.kpad, 1) > shape.budget) shape.budget = (size_t)kSmemLimit;   // large residues: one CTA per SM
    const int fb = frames_per_cta(L.kpad, shape.budget);
    const size_t smem = mdmat_smem_bytes(L.kpad, fb);

    MdmatParams prm;
    prm.cnt = d_cnt; prm.res_start = d_res_start; prm.perm = d_perm; prm.counters = d_counters;
    prm.natoms = natoms; prm.nres = nres; prm.nres_pad = L.nres_pad; prm.kpad = L.kpad; prm.fb = fb;
    prm.symmetric = symmetric ? 1 : 0;
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
        // With eight slots a residue's record is 32 bytes per plane and eight consecutive residues would
        // meet in four of the eight 16-byte bank groups; the 32-byte swizzle spreads them over all eight.
        const char* d = getenv("CONAN_MDMAT_DEBUG");
        prm.dbg = d ? atoi(d) : 0;
        const char* z = getenv("CONAN_MDMAT_SWZ");
        const bool want = z ? z[0] == '1' : kDefaultSwz;
        prm.swz = (want && prm.use_tma && (L.kpad == 8 || L.kpad == 16)) ? 1 : 0;
    }
    const int64_t P = symmetric ? (int64_t)nres * (nres - 1) / 2 : (int64_t)xres * yres;
    if (P == 0) return CONAN_OK;
    if (nres > 46340 || P > 2147483647ll)
        return conan::fail(CONAN_ERR_SIZE, "conan_mdmat_frames: %d residues exceed the supported 46340 (pair index must fit 31 bits)", nres);
    if (levels_stride < P)
        return conan::fail(CONAN_ERR_ARG, "conan_mdmat_frames: levels_stride=%lld < P=%lld", (long long)levels_stride, (long long)P);
    prm.stride = levels_stride;

    // Frames per launch: the grid is (tiles, frame groups) and its second dimension ends at 65535.
    const int64_t frames_per_launch = (int64_t)65535 * fb;
    for (int64_t fbase = 0; fbase < nf; fbase += frames_per_launch) {
        const int64_t nfl = (nf - fbase < frames_per_launch) ? nf - fbase : frames_per_launch;
        // The tensor maps cover only this launch's frames, so coordinate 3 stays small.  Both describe the
        // same tensor [frame][4 planes][residue][kpad]; the row tile takes three planes, the column tile four.
        float* base = d_slots + (size_t)fbase * kPlanes * L.kpad * L.nres_pad;
        const TileMaps* maps = tile_maps(encode, base, L.kpad, L.nres_pad, nfl, fb, prm.swz);
        if (!maps) return conan::fail(CONAN_ERR_CUDA, "conan_mdmat_frames: cuTensorMapEncodeTiled failed");
        MdmatParams p2 = prm;
        // frame indices inside the kernel are relative to this launch's base pointers
        p2.sph = d_sph + (size_t)fbase * L.nres_pad;
        p2.xyz = d_xyz + (size_t)fbase * natoms * 3;
        p2.box = d_box ? d_box + 9 * fbase : nullptr;
        p2.levels = d_levels + (size_t)fbase * prm.stride;
        p2.slots = base;
        p2.nf = nfl;
        const int64_t ngroups = (nfl + fb - 1) / fb;
        const dim3 grid((unsigned)prm.ntiles, (unsigned)ngroups);
        if (shape.per_sm == 2) {
            if (prm.pbc) mdmat_tile_kernel<true, 256, 2><<<grid, 256, smem, stream>>>(maps->I, maps->J, p2);
            else mdmat_tile_kernel<false, 256, 2><<<grid, 256, smem, stream>>>(maps->I, maps->J, p2);
        } else {
            if (prm.pbc) mdmat_tile_kernel<true, 256, 3><<<grid, 256, smem, stream>>>(maps->I, maps->J, p2);
            else mdmat_tile_kernel<false, 256, 3><<<grid, 256, smem, stream>>>(maps->I, maps->J, p2);
        }
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
