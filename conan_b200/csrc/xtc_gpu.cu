// GPU-side .xtc decoding: the compressed trajectory bytes go to the device as they
// are on disk and every frame is decoded there, so the host-to-device copy carries ~3 bytes per atom
// instead of 12 and the host cores are out of the stage-1 feed (`gmx mdmat -f`, conan.py:1437-1453).
//
// A frame's bit stream (public xdrfile xdr3dfcoord scheme) is a chain of variable-width codes, but the
// widths depend only on the run flags between the codes.  Two decoders: in one pass a warp takes one
// frame, walks the flags together (sequential) and decodes 32 groups of atoms at a time, one per lane
// (xtc_decode_kernel); in two passes one thread per frame walks the flags and writes a descriptor per
// group, then one thread per group decodes (xtc_walk_kernel, xtc_groups_kernel; needs scratch memory).
// Streams are read as aligned big-endian 32-bit words (XDR pads every frame to 4 bytes); only the
// atoms of the index group are written.  The arithmetic is the integer arithmetic of the host decoder
// in xtc.cpp, so the coordinates are bit-identical.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdint>

#include "common.h"

namespace {

constexpr int kFirstIdx = 9;
constexpr int kNumMagic = 73;
constexpr int kDecodeThreads = 128;   // four frames (warps) per block
__constant__ int c_magic[kNumMagic] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 8, 10, 12, 16, 20, 25, 32, 40, 50, 64,
    80, 101, 128, 161, 203, 256, 322, 406, 512, 645, 812, 1024, 1290,
    1625, 2048, 2580, 3250, 4096, 5060, 6501, 8192, 10321, 13003,
    16384, 20642, 26007, 32768, 41285, 52015, 65536, 82570, 104031,
    131072, 165140, 208063, 262144, 330280, 416127, 524287, 660561,
    832255, 1048576, 1321122, 1664510, 2097152, 2642245, 3329021,
    4194304, 5284491, 6658042, 8388607, 10568983, 13316085, 16777216};

enum : int { kOk = 0, kBadHeader = 1, kBadSmallIdx = 2, kRunPastEnd = 3, kStreamPastEnd = 4 };

__device__ __forceinline__ uint32_t load_be32(const uint8_t* p) {
    return __byte_perm(__ldg(reinterpret_cast<const uint32_t*>(p)), 0, 0x0123);
}

// MSB-first bit stream over aligned 32-bit words, one word loaded ahead of its use so the load latency
// sits off the decode chain; reads past `end` give zeros (the walk reports a stream that is too short).
struct BitStream {
    const uint8_t* p;      // address of the word after `next`
    const uint8_t* end;
    uint64_t window;       // the next bits, left-aligned
    uint32_t next;         // the word that follows the window
    int have;
    // start at bit `pos` of the stream that begins at `begin`
    __device__ __forceinline__ void init_at(const uint8_t* begin, const uint8_t* e, int64_t pos) {
        end = e;
        p = begin + 4 * (pos >> 5);
        const uint32_t w0 = (p < end) ? load_be32(p) : 0u, w1 = (p + 4 < end) ? load_be32(p + 4) : 0u;
        next = (p + 8 < end) ? load_be32(p + 8) : 0u;
        p += 12;
        const int skip = (int)(pos & 31);
        window = (((uint64_t)w0 << 32) | w1) << skip;
        have = 64 - skip;
    }
    __device__ __forceinline__ uint32_t get(int nbits) {   // 0 <= nbits <= 32
        if (have < nbits) {
            window |= (uint64_t)next << (32 - have);
            have += 32;
            // a word past the end reads as zeros; whether its bits were used is checked when the frame is done
            next = (p < end) ? load_be32(p) : 0u;
            p += 4;
        }
        const uint32_t v = (uint32_t)((window >> 1) >> (63 - nbits));
        window <<= nbits;
        have -= nbits;
        return v;
    }
};

// The packed-integer layout of xdrfile's receiveints: full bytes least significant first, then the
// remaining high bits.  nbits <= 64.
__device__ __forceinline__ uint64_t get_packed64(BitStream& bs, int nbits) {
    if (nbits <= 0) return 0;
    const int full = (nbits - 1) >> 3, part = nbits - 8 * full;     // `full` whole bytes, then 1..8 bits
    uint64_t x;                                                     // the nbits as one big-endian number
    if (nbits <= 32) x = bs.get(nbits);
    else { const uint64_t hi = bs.get(nbits - 32); x = (hi << 32) | bs.get(32); }
    const uint64_t top = part < 64 ? x >> part : 0;                 // byte 0 is its most significant byte
    const uint64_t lsb = x & ((1ull << part) - 1);
    // reverse the `full` bytes of top
    const uint32_t tl = (uint32_t)top, th = (uint32_t)(top >> 32);
    const uint64_t rev = ((uint64_t)__byte_perm(tl, 0, 0x0123) << 32) | __byte_perm(th, 0, 0x0123);
    const uint64_t low = full ? rev >> (64 - 8 * full) : 0;
    return low | (lsb << (8 * full));
}

// v / s and v % s through a precomputed m = ceil(2^64 / s) (s >= 2): mulhi gives the quotient or one more
__device__ __forceinline__ uint64_t divmod64(uint64_t v, uint32_t s, uint64_t m, uint32_t& rem) {
    uint64_t q = __umul64hi(v, m);
    int64_t r = (int64_t)(v - q * s);
    if (r < 0) { q--; r += s; }
    rem = (uint32_t)r;
    return q;
}
__device__ __forceinline__ uint32_t divmod32(uint32_t v, uint32_t s, uint32_t m, uint32_t& rem) {
    uint32_t q = __umulhi(v, m);
    int32_t r = (int32_t)(v - q * s);
    if (r < 0) { q--; r += (int32_t)s; }
    rem = (uint32_t)r;
    return q;
}

struct Triple { int32_t c[3]; };

// up to 72 bits (three sizes below 2^24): long division over bytes, most significant first
__device__ __noinline__ Triple unpack3_wide(BitStream& bs, int nbits, uint32_t s1, uint32_t s2) {
    uint32_t bytes[12];
    Triple out;
    int nb = 0, left = nbits;
    while (left > 8) { bytes[nb++] = bs.get(8); left -= 8; }
    if (left > 0) bytes[nb++] = bs.get(left);
    for (int k = nb; k < 4; k++) bytes[k] = 0;
    const uint32_t div[2] = {s2, s1};
    for (int d = 0; d < 2; d++) {
        uint32_t num = 0;
        for (int k = nb - 1; k >= 0; k--) {
            // num < div <= 2^24, so num * 256 + byte fits 32 bits
            num = (num << 8) | bytes[k];
            const uint32_t qd = num / div[d];
            bytes[k] = qd;
            num -= qd * div[d];
        }
        out.c[2 - d] = (int32_t)num;
    }
    out.c[0] = (int32_t)(bytes[0] | (bytes[1] << 8) | (bytes[2] << 16) | (bytes[3] << 24));
    return out;
}

__device__ __forceinline__ int bits_for(uint32_t size) {
    uint32_t num = 1;
    int n = 0;
    while (size >= num && n < 32) { n++; num <<= 1; }
    return n;
}

struct DecodeParams {
    const uint8_t* data;
    int64_t nbytes;
    const int64_t* offsets;    // [nframes] byte offset of each frame to decode
    int64_t nframes;
    int natoms;
    const int32_t* slot;       // [natoms] output row of each atom or -1; null = all atoms
    int nsel;                  // rows per frame in the output
    float* xyz;                // [nframes][nsel][3]
    int32_t* status;           // [nframes]
};

// One warp per frame.  The width of every code is known once the run flags are read, so the warp first
// walks the frame's group structure -- a full-size code, a flag (1 or 6 bits), run/3 small codes --
// 32 groups at a time, all lanes in step on the same few bits (the chain is ~20 instructions and one
// cached load per group instead of the whole decode), and lane g keeps the description of group g;
// then the 32 groups are decoded in parallel, one per lane.
__global__ void __launch_bounds__(kDecodeThreads, 8) xtc_decode_kernel(const DecodeParams prm) {
    const int lane = threadIdx.x & 31;
    const int64_t f = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (f >= prm.nframes) return;
    const int64_t off = prm.offsets[f];
    const int natoms = prm.natoms;
    float* out = prm.xyz + (size_t)f * prm.nsel * 3;
    const int32_t* slot = prm.slot;
    auto finish = [&](int code) { if (lane == 0) prm.status[f] = code; };
    if (off < 0 || (off & 3) || off + 56 > prm.nbytes) return finish(kBadHeader);
    const uint8_t* p = prm.data + off;
    if ((int)load_be32(p) != 1995 || (int)load_be32(p + 4) != natoms) return finish(kBadHeader);
    const uint8_t* q = p + 56;
    if (natoms <= 9) {   // stored as plain floats
        if (off + 56 + (int64_t)natoms * 12 > prm.nbytes) return finish(kBadHeader);
        for (int a = lane; a < natoms; a += 32) {
            const int s = slot ? slot[a] : a;
            if (s >= 0 && s < prm.nsel)
                for (int d = 0; d < 3; d++) out[3 * (size_t)s + d] = __uint_as_float(load_be32(q + 12 * a + 4 * d));
        }
        return finish(kOk);
    }
    if (off + 92 > prm.nbytes) return finish(kBadHeader);
    const float precision = __uint_as_float(load_be32(q));
    int32_t minint[3];
    uint32_t sizeint[3];
    for (int d = 0; d < 3; d++) {
        minint[d] = (int32_t)load_be32(q + 4 + 4 * d);
        sizeint[d] = (uint32_t)((int32_t)load_be32(q + 16 + 4 * d) - minint[d] + 1);
    }
    int smallidx = (int32_t)load_be32(q + 28);
    const int32_t nb = (int32_t)load_be32(q + 32);
    if (smallidx < kFirstIdx || smallidx >= kNumMagic) return finish(kBadSmallIdx);
    if (nb < 0 || off + 92 + (((int64_t)nb + 3) & ~(int64_t)3) > prm.nbytes) return finish(kBadHeader);
    int bitsint[3] = {0, 0, 0}, bitsize = 0, bigbits;
    if ((sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffffu) {
        for (int d = 0; d < 3; d++) bitsint[d] = bits_for(sizeint[d]);
        bigbits = bitsint[0] + bitsint[1] + bitsint[2];
    } else {
        // bit length of the product of the three sizes (each below 2^24: at most 72 bits)
        const uint64_t p01 = (uint64_t)sizeint[0] * sizeint[1];
        const uint64_t lo = p01 * sizeint[2], hi = __umul64hi(p01, (uint64_t)sizeint[2]);
        bitsize = hi ? 128 - __clzll((long long)hi) : (lo ? 64 - __clzll((long long)lo) : 0);
        bigbits = bitsize;
    }
    // reciprocals for the two divisions of the full-size codes (sizes of 1 or more than 64 bits: slow path)
    const bool big_fast = bitsize > 0 && bitsize <= 64 && sizeint[1] >= 2 && sizeint[2] >= 2;
    uint64_t rcp1 = 0, rcp2 = 0;
    if (big_fast) { rcp1 = ~0ull / sizeint[1] + 1; rcp2 = ~0ull / sizeint[2] + 1; }

    const uint8_t* sbase = q + 36;
    const uint8_t* send = sbase + (((size_t)nb + 3) & ~(size_t)3);
    const int64_t stream_bits = (int64_t)(send - sbase) * 8;
    // the host decoder multiplies by the rounded reciprocal of the precision; so does this one
    const float inv = __fdiv_rn(1.0f, precision);
    auto emit = [&](int atom, const int32_t c[3]) {
        const int s = slot ? slot[atom] : atom;
        if (s >= 0 && s < prm.nsel) {   // (a slot outside the output is ignored rather than written)
            float* o = out + 3 * (size_t)s;
            o[0] = (float)c[0] * inv; o[1] = (float)c[1] * inv; o[2] = (float)c[2] * inv;
        }
    };
    auto word_at = [&](int64_t w) -> uint32_t {   // big-endian word w of the stream, zeros past its end
        const uint8_t* a = sbase + 4 * w;
        return a < send ? load_be32(a) : 0u;
    };

    auto six_at = [&](int64_t bit) -> uint32_t {   // the flag and the five run bits that may follow it
        const uint32_t w0 = word_at(bit >> 5), w1 = word_at((bit >> 5) + 1);
        return __funnelshift_l(w1, w0, (int)(bit & 31)) >> 26;
    };

    int64_t pos = 0;          // bit position of the next group (warp-uniform, like run / smallidx / i)
    int run = 0, i = 0, bad = kOk;
    while (i < natoms && bad == kOk) {
        // ---- where the next groups start: up to 32 of them, lane g keeps group g ----
        // As long as the flag bit after a full-size code is 0 the run length and the small-code width
        // stay as they are, so the groups sit at a fixed stride: lane k looks at the flag of the group
        // k strides ahead; everything before the first set flag is confirmed in one step, and the
        // group with the set flag (its 5 run bits give the new run length) is taken along.
        int64_t my_pos = 0;
        int my_run = 0, my_sidx = 0, my_first = -1;
        int nq = 0;           // groups collected so far
        while (nq < 32 && i < natoms && bad == kOk) {
            const int nsmall = run / 3, gatoms = 1 + nsmall;
            const int64_t stride = bigbits + 1 + (int64_t)nsmall * smallidx;
            const int first_k = i + lane * gatoms;
            const bool exists = first_k < natoms && lane < 32 - nq;
            const uint32_t six = exists ? six_at(pos + lane * stride + bigbits) : 0u;
            const unsigned flagged = __ballot_sync(0xffffffffu, exists && (six >> 5));
            const unsigned present = __ballot_sync(0xffffffffu, exists);
            const int nplain = min(flagged ? __ffs(flagged) - 1 : 32, __popc(present));   // groups with flag 0
            if (run > 0 && nplain > 0 && i + nplain * gatoms > natoms) { bad = kRunPastEnd; break; }
            const int j = lane - nq;          // this lane's group among the new ones
            if (j >= 0 && j < nplain) { my_pos = pos + j * stride; my_run = run; my_sidx = smallidx; my_first = i + j * gatoms; }
            pos += nplain * stride;
            i += nplain * gatoms;
            nq += nplain;
            if ((flagged >> nplain) & 1u) {
                // the group whose flag is set: new run length now, new code width from the next group on
                const int r5 = (int)(__shfl_sync(0xffffffffu, six, nplain) & 31u);
                int is_smaller = r5 % 3;
                run = r5 - is_smaller;
                is_smaller--;
                const int ns = run / 3;
                if (run > 0 && i + 1 + ns > natoms) { bad = kRunPastEnd; break; }
                if (lane == nq) { my_pos = pos; my_run = run; my_sidx = smallidx; my_first = i; }
                pos += bigbits + 6 + (int64_t)ns * smallidx;
                i += 1 + ns;
                nq++;
                smallidx += is_smaller;
                if (smallidx < kFirstIdx || smallidx >= kNumMagic) bad = kBadSmallIdx;
            }
        }
        if (pos > stream_bits && bad == kOk) bad = kStreamPastEnd;
        // ---- one group per lane (groups confirmed before a fault are still decoded; the reader pads
        //      with zeros past the end of the stream) ----
        if (my_first >= 0) {
            BitStream bs;
            bs.init_at(sbase, send, my_pos);
            int32_t cur[3], prev[3];
            if (bitsize == 0) {
                for (int d = 0; d < 3; d++) cur[d] = (int32_t)bs.get(bitsint[d]);
            } else if (big_fast) {
                // value = (c0 * s1 + c1) * s2 + c2
                const uint64_t v = get_packed64(bs, bitsize);
                uint32_t r2, r1;
                const uint64_t qq = divmod64(v, sizeint[2], rcp2, r2);
                const uint64_t q2 = divmod64(qq, sizeint[1], rcp1, r1);
                cur[2] = (int32_t)r2; cur[1] = (int32_t)r1; cur[0] = (int32_t)q2;
            } else {
                const Triple t = unpack3_wide(bs, bitsize, sizeint[1], sizeint[2]);
                for (int d = 0; d < 3; d++) cur[d] = t.c[d];
            }
            for (int d = 0; d < 3; d++) { cur[d] += minint[d]; prev[d] = cur[d]; }
            if (bs.get(1)) bs.get(5);
            if (my_run > 0) {
                const uint32_t sizesmall = (uint32_t)c_magic[my_sidx];
                const int smallnum = (int)(sizesmall >> 1);
                // ceil(2^32 / size): mulhi(v, m) is v / size or one more for every v < 2^32
                const uint32_t rcpsmall = 0xffffffffu / sizesmall + 1;
                int a = my_first + 1;
                for (int k = 0; k < my_run; k += 3) {
                    if (my_sidx <= 32) {
                        const uint32_t v = (uint32_t)get_packed64(bs, my_sidx);
                        uint32_t r2, r1;
                        const uint32_t qq = divmod32(v, sizesmall, rcpsmall, r2);
                        const uint32_t q2 = divmod32(qq, sizesmall, rcpsmall, r1);
                        cur[2] = (int32_t)r2; cur[1] = (int32_t)r1; cur[0] = (int32_t)q2;
                    } else {
                        const Triple t = unpack3_wide(bs, my_sidx, sizesmall, sizesmall);
                        for (int d = 0; d < 3; d++) cur[d] = t.c[d];
                    }
                    for (int d = 0; d < 3; d++) cur[d] += prev[d] - smallnum;
                    if (k == 0) {
                        // the first two atoms of a run are stored in swapped order
                        for (int d = 0; d < 3; d++) { const int32_t t = cur[d]; cur[d] = prev[d]; prev[d] = t; }
                        emit(my_first, prev);
                    } else {
                        for (int d = 0; d < 3; d++) prev[d] = cur[d];
                    }
                    emit(a++, cur);
                }
            } else {
                emit(my_first, cur);
            }
        }
    }
    finish(bad);
}


// ------------------------------------------------------------------------------------------
// Two passes (conan_xtc_decode_device_ws).  The walk over the run flags is a chain of a few integer
// operations per group; done by a whole warp for one frame it occupies an issue slot per step for one
// lane's worth of work.  Here pass 1 gives every frame one THREAD (32 frames share an issue slot), which
// writes one 8-byte descriptor per group -- bit position, first atom, number of small codes and their
// width -- and a record with the frame's header values; pass 2 decodes the groups, one per thread, with
// as many warps per frame as the frame has groups.  Same integer arithmetic, same coordinates.
// ------------------------------------------------------------------------------------------
constexpr int kMaxWsAtoms = 1 << 21;      // `first` has 21 bits in a descriptor
constexpr int kGroupThreads = 128;
constexpr int kGroupsPerCta = 512;
constexpr int kRunQueue = 64;           // parked runs per warp (at most 31 waiting + 32 new)

struct __align__(16) FrameRec {
    int64_t stream_off;      // byte offset of the code stream inside d_data
    uint64_t rcp1, rcp2;
    uint32_t stream_bytes;   // its length, padded to 4
    int32_t ngroups;         // 0: nothing for pass 2 (bad header, or a frame of plain floats written by pass 1)
    int32_t minint[3];
    uint32_t sizeint[3];
    int32_t bitsint[3];
    int32_t bitsize;
    int32_t big_fast;
    float inv;
    int32_t pad[2];
};
static_assert(sizeof(FrameRec) % 16 == 0, "FrameRec is read with aligned loads");

struct WalkParams {
    DecodeParams d;
    FrameRec* rec;           // [nframes]
    uint2* desc;             // [nframes][natoms]
};

constexpr int kWalkWindow = 256;        // words of every frame's stream held in shared memory by the walk
constexpr int kWalkWarps = 2;           // warps (of 32 frames) per CTA: 66 KB of windows, which fit beside the two resident CTAs
                                        // of the stage-2 kernel (a pipeline runs the walk of the next shard there); beside the
                                        // distance kernel, whose CTAs hold all registers, a walk CTA keeps one of them out
constexpr int kWalkSteps = 4;           // groups per round of the walk
// words a lane may advance in one round before its reads leave the window: a group is at most 96 + 6 + 10 x 72 bits
constexpr int kWalkMargin = ((kWalkSteps - 1) * (96 + 6 + 10 * 72) + 31) / 32 + 2;
constexpr size_t kWalkSmem = (size_t)kWalkWarps * 32 * (kWalkWindow + 1) * sizeof(uint32_t);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// c_rcp_magic[k] = ceil(2^32 / c_magic[k]): mulhi(v, m) is v / size or one more for every v < 2^32
__constant__ uint32_t c_rcp_magic[kNumMagic] = {
    0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u,
    0u, 536870912u, 429496730u, 357913942u, 268435456u, 214748365u, 171798692u, 134217728u,
    107374183u, 85899346u, 67108864u, 53687092u, 42524429u, 33554432u, 26676816u, 21157475u,
    16777216u, 13338408u, 10578738u, 8388608u, 6658865u, 5289369u, 4194304u, 3329433u,
    2643057u, 2097152u, 1664717u, 1321529u, 1048576u, 848808u, 660663u, 524288u,
    416139u, 330306u, 262144u, 208070u, 165147u, 131072u, 104033u, 82572u,
    65536u, 52017u, 41286u, 32768u, 26009u, 20643u, 16384u, 13005u,
    10322u, 8193u, 6502u, 5161u, 4096u, 3251u, 2581u, 2048u,
    1626u, 1291u, 1024u, 813u, 646u, 513u, 407u, 323u,
    256u};

__global__ void __launch_bounds__(32 * kWalkWarps) xtc_walk_kernel(const WalkParams wp) {
    // the next kWalkWindow words of each of the warp's 32 streams (already in host byte order, zeros past the
    // end), refilled for all lanes with coalesced loads whenever one lane reaches the end of its window:
    // the chain below then depends on shared-memory reads, not on 32 scattered global loads of which one
    // misses the L1 in nearly every step
    extern __shared__ uint32_t win_all[];      // [kWalkWarps][32][kWalkWindow + 1]; the warps work on their own
    uint32_t* win = win_all + (size_t)(threadIdx.x >> 5) * 32 * (kWalkWindow + 1);
    const DecodeParams& prm = wp.d;
    const int lane = threadIdx.x & 31;
    const int64_t f = (int64_t)blockIdx.x * (32 * kWalkWarps) + threadIdx.x;
    const int natoms = prm.natoms;
    FrameRec r;
    r.ngroups = 0;
    int bad = kOk, smallidx = 0, bigbits = 0;
    bool active = false;
    const uint32_t* words = nullptr;
    uint32_t nwords = 0;
    if (f < prm.nframes) {
        // (one pass through this do-block; `break` leaves it with `bad` set)
        do {
            const int64_t off = prm.offsets[f];
            bad = kBadHeader;
            if (off < 0 || (off & 3) || off + 56 > prm.nbytes) break;
            const uint8_t* p = prm.data + off;
            if ((int)load_be32(p) != 1995 || (int)load_be32(p + 4) != natoms) break;
            const uint8_t* q = p + 56;
            if (natoms <= 9) {   // stored as plain floats
                if (off + 56 + (int64_t)natoms * 12 > prm.nbytes) break;
                float* out = prm.xyz + (size_t)f * prm.nsel * 3;
                for (int a = 0; a < natoms; a++) {
                    const int sl = prm.slot ? prm.slot[a] : a;
                    if (sl >= 0 && sl < prm.nsel)
                        for (int d = 0; d < 3; d++) out[3 * (size_t)sl + d] = __uint_as_float(load_be32(q + 12 * a + 4 * d));
                }
                bad = kOk;
                break;
            }
            if (off + 92 > prm.nbytes) break;
            const float precision = __uint_as_float(load_be32(q));
            for (int d = 0; d < 3; d++) {
                r.minint[d] = (int32_t)load_be32(q + 4 + 4 * d);
                r.sizeint[d] = (uint32_t)((int32_t)load_be32(q + 16 + 4 * d) - r.minint[d] + 1);
            }
            smallidx = (int32_t)load_be32(q + 28);
            const int32_t nb = (int32_t)load_be32(q + 32);
            if (smallidx < kFirstIdx || smallidx >= kNumMagic) { bad = kBadSmallIdx; break; }
            if (nb < 0 || off + 92 + (((int64_t)nb + 3) & ~(int64_t)3) > prm.nbytes) break;
            r.bitsint[0] = r.bitsint[1] = r.bitsint[2] = 0;
            r.bitsize = 0;
            if ((r.sizeint[0] | r.sizeint[1] | r.sizeint[2]) > 0xffffffu) {
                for (int d = 0; d < 3; d++) r.bitsint[d] = bits_for(r.sizeint[d]);
                bigbits = r.bitsint[0] + r.bitsint[1] + r.bitsint[2];
            } else {
                const uint64_t p01 = (uint64_t)r.sizeint[0] * r.sizeint[1];
                const uint64_t lo = p01 * r.sizeint[2], hi = __umul64hi(p01, (uint64_t)r.sizeint[2]);
                r.bitsize = hi ? 128 - __clzll((long long)hi) : (lo ? 64 - __clzll((long long)lo) : 0);
                bigbits = r.bitsize;
            }
            r.big_fast = r.bitsize > 0 && r.bitsize <= 64 && r.sizeint[1] >= 2 && r.sizeint[2] >= 2;
            r.rcp1 = r.rcp2 = 0;
            if (r.big_fast) { r.rcp1 = ~0ull / r.sizeint[1] + 1; r.rcp2 = ~0ull / r.sizeint[2] + 1; }
            const uint8_t* sbase = q + 36;
            const uint32_t sbytes = (uint32_t)(((size_t)nb + 3) & ~(size_t)3);
            r.stream_off = (int64_t)(sbase - prm.data);
            r.stream_bytes = sbytes;
            r.inv = __fdiv_rn(1.0f, precision);
            r.pad[0] = r.pad[1] = 0;
            words = reinterpret_cast<const uint32_t*>(sbase);
            nwords = sbytes >> 2;
            bad = kOk;
            active = true;
        } while (false);
    }

    // the chain: position -> six bits behind the full-size code -> run length / code width -> next position.
    // Written without branches (a lane that is done idles under its predicate) and in rounds of kWalkSteps
    // groups, so the warp votes once per round: is a refill due, is anyone still walking.
    uint2* desc = wp.desc + (size_t)(f < prm.nframes ? f : 0) * natoms;
    uint32_t* mywin = win + (size_t)lane * (kWalkWindow + 1);
    uint32_t pos = 0, base = 0;
    bool filled = false;
    int ns = 0, i = 0, g = 0;
    const uint32_t nw_eff = active ? nwords : 0u;
    const uint32_t wlo = (uint32_t)reinterpret_cast<uintptr_t>(words), whi = (uint32_t)(reinterpret_cast<uintptr_t>(words) >> 32);
    while (__any_sync(0xffffffffu, active)) {
        const uint32_t w_now = (pos + (uint32_t)bigbits) >> 5;
        const bool need = active && (!filled || w_now + 1 + kWalkMargin - base >= (uint32_t)kWalkWindow);
        if (__any_sync(0xffffffffu, need)) {
            // all 32 windows restart at their lane's position: asynchronous 4-byte copies straight into shared
            // memory (zeros past the end of a stream), one wait for all of them
            if (active) base = w_now;
            __syncwarp();
#pragma unroll 2
            for (int k = 0; k < 32; k++) {
                const uint32_t nk = __shfl_sync(0xffffffffu, nw_eff, k), bk = __shfl_sync(0xffffffffu, base, k);
                const uint32_t lo = __shfl_sync(0xffffffffu, wlo, k), hi = __shfl_sync(0xffffffffu, whi, k);
                if (nk == 0) continue;
                const uint32_t* wk = reinterpret_cast<const uint32_t*>(((uintptr_t)hi << 32) | lo);
                const uint32_t dst = smem_u32(win + (size_t)k * (kWalkWindow + 1) + lane);
#pragma unroll
                for (int j = 0; j < kWalkWindow / 32; j++) {
                    const uint32_t idx = bk + 32 * j + lane;
                    const bool in = idx < nk;
                    cp_async4(dst + 128u * j, wk + (in ? idx : 0u), in ? 4u : 0u);
                }
            }
            cp_async_wait_all();
            __syncwarp();
            filled = true;
        }
#pragma unroll
        for (int u = 0; u < kWalkSteps; u++) {
            const uint32_t at = pos + (uint32_t)bigbits;
            const uint32_t rel = (at >> 5) - base;
            const uint32_t w0 = __byte_perm(mywin[rel], 0, 0x0123), w1 = __byte_perm(mywin[rel + 1], 0, 0x0123);
            const uint32_t six = __funnelshift_l(w1, w0, (int)(at & 31)) >> 26;
            const bool flag = (six >> 5) != 0;
            const int r5 = (int)(six & 31u);
            const int q3 = (r5 * 11) >> 5;                  // r5 / 3 for r5 < 32
            const int sm = r5 - 3 * q3;
            ns = flag ? q3 : ns;                            // number of small codes of this and the following groups
            const int d_idx = flag ? sm - 1 : 0;
            const bool fits = !(ns > 0 && i + 1 + ns > natoms);
            const bool ok = active && fits;
            if (ok) desc[g] = make_uint2(pos, (uint32_t)i | ((uint32_t)ns << 21) | ((uint32_t)smallidx << 25));
            if (active && !fits) bad = kRunPastEnd;
            g += ok ? 1 : 0;
            pos += ok ? (uint32_t)(bigbits + (flag ? 6 : 1) + ns * smallidx) : 0u;
            i += ok ? 1 + ns : 0;
            smallidx += ok ? d_idx : 0;
            const bool idx_bad = smallidx < kFirstIdx || smallidx >= kNumMagic;
            if (ok && idx_bad) bad = kBadSmallIdx;
            active = ok && !idx_bad && i < natoms;
        }
    }
    if (f < prm.nframes) {
        if (nwords && (uint64_t)pos > (uint64_t)nwords * 32 && bad == kOk) bad = kStreamPastEnd;
        r.ngroups = g;        // groups confirmed before a fault are still decoded (the reader pads with zeros)
        prm.status[f] = bad;
        wp.rec[f] = r;
    }
}

// 64 bits of the stream from bit `pos` on (big-endian bit order, zeros past the end)
__device__ __forceinline__ uint64_t peek64(const uint32_t* __restrict__ words, uint32_t nwords, uint32_t pos) {
    const uint32_t w = pos >> 5;
    const int sh = (int)(pos & 31);
    const uint32_t a = w < nwords ? __byte_perm(__ldg(words + w), 0, 0x0123) : 0u;
    const uint32_t b = w + 1 < nwords ? __byte_perm(__ldg(words + w + 1), 0, 0x0123) : 0u;
    const uint32_t c = w + 2 < nwords ? __byte_perm(__ldg(words + w + 2), 0, 0x0123) : 0u;
    return ((uint64_t)__funnelshift_l(b, a, sh) << 32) | __funnelshift_l(c, b, sh);
}

// the same from three words that are already here (a = the word that holds bit `pos`), and one bit of the 96
__device__ __forceinline__ uint64_t peek64_words(uint32_t a, uint32_t b, uint32_t c, uint32_t pos) {
    const int sh = (int)(pos & 31);
    return ((uint64_t)__funnelshift_l(b, a, sh) << 32) | __funnelshift_l(c, b, sh);
}
__device__ __forceinline__ uint32_t bit_of_words(uint32_t a, uint32_t b, uint32_t c, uint32_t idx) {   // idx < 96
    const uint32_t w = idx < 32 ? a : (idx < 64 ? b : c);
    return (w >> (31 - (idx & 31))) & 1u;
}

// xdrfile's receiveints layout: of the nbits (1..64, here the top bits of X) the first `full` whole bytes are the
// value's low bytes, least significant first, the last `part` (1..8) bits its top
__device__ __forceinline__ uint64_t packed_value(uint64_t X, int nbits, int full, int part) {
    const uint64_t x = X >> (64 - nbits);
    const uint64_t top = x >> part;
    const uint32_t lsb = (uint32_t)x & ((1u << part) - 1u);
    const uint64_t rev = ((uint64_t)__byte_perm((uint32_t)top, 0, 0x0123) << 32) | __byte_perm((uint32_t)(top >> 32), 0, 0x0123);
    const uint64_t low = full ? rev >> (64 - 8 * full) : 0ull;
    return low | ((uint64_t)lsb << (8 * full));
}

__global__ void __launch_bounds__(kGroupThreads) xtc_groups_kernel(const WalkParams wp) {
    const DecodeParams& prm = wp.d;
    const int64_t f = blockIdx.y;
    const FrameRec* rp = wp.rec + f;
    const int ngroups = rp->ngroups;
    const int g0 = (int)blockIdx.x * kGroupsPerCta;
    if (g0 >= ngroups) return;
    const int g1 = min(ngroups, g0 + kGroupsPerCta);
    const int natoms = prm.natoms;
    const int32_t minint[3] = {rp->minint[0], rp->minint[1], rp->minint[2]};
    const uint32_t sizeint[3] = {rp->sizeint[0], rp->sizeint[1], rp->sizeint[2]};
    const int bitsint[3] = {rp->bitsint[0], rp->bitsint[1], rp->bitsint[2]};
    const int bitsize = rp->bitsize;
    const bool big_fast = rp->big_fast != 0;
    const uint64_t rcp1 = rp->rcp1, rcp2 = rp->rcp2;
    const float inv = rp->inv;
    const uint8_t* sbase = prm.data + rp->stream_off;
    const uint8_t* send = sbase + rp->stream_bytes;
    const uint32_t* words = reinterpret_cast<const uint32_t*>(sbase);
    const uint32_t nwords = rp->stream_bytes >> 2;
    const int bigbits = bitsize ? bitsize : bitsint[0] + bitsint[1] + bitsint[2];
    const int bfull = bitsize ? (bitsize - 1) >> 3 : 0, bpart = bitsize - 8 * bfull;
    float* out = prm.xyz + (size_t)f * prm.nsel * 3;
    const int32_t* slot = prm.slot;
    const uint2* desc = wp.desc + (size_t)f * natoms;
    auto emit = [&](int atom, int32_t c0, int32_t c1, int32_t c2) {
        int s = atom;                  // without a slot map nsel == natoms and every atom has its row
        if (slot) {
            s = slot[atom];
            if (s < 0 || s >= prm.nsel) return;
        }
        float* o = out + 3u * (uint32_t)s;      // (at most 2^21 atoms per frame)
        o[0] = (float)c0 * inv; o[1] = (float)c1 * inv; o[2] = (float)c2 * inv;
    };
    // Groups with small codes behind the full-size one are rare and long (a dependent chain of up to ten
    // codes), so they are not decoded where they are met: the lane parks them in a queue of its warp, and
    // whenever 32 are waiting (and once more at the end) the warp decodes one run per lane.
    __shared__ uint32_t q_pos[kGroupThreads / 32][kRunQueue], q_meta[kGroupThreads / 32][kRunQueue];
    __shared__ int32_t q_prev[kGroupThreads / 32][3][kRunQueue];
    const int warp = (int)threadIdx.x >> 5, lane = (int)threadIdx.x & 31;
    uint32_t* qp = q_pos[warp];
    uint32_t* qm = q_meta[warp];
    int queued = 0;                 // warp-uniform

    auto decode_runs = [&](int from, int n) {         // n (<= 32) entries of the queue starting at `from`, one per lane
        if (lane < n) {
            const int e = from + lane;
            const uint32_t meta = qm[e];
            uint32_t sp = qp[e];
            const int my_first = (int)(meta & 0x1fffffu), my_ns = (int)((meta >> 21) & 15u), my_sidx = (int)(meta >> 25);
            int32_t prev[3] = {q_prev[warp][0][e], q_prev[warp][1][e], q_prev[warp][2][e]};
            int32_t cur[3];
            const uint32_t sizesmall = (uint32_t)c_magic[my_sidx];
            const int smallnum = (int)(sizesmall >> 1);
            int a = my_first + 1;
            if (my_sidx <= 32) {
                const uint32_t rcpsmall = c_rcp_magic[my_sidx];
                const int sfull = (my_sidx - 1) >> 3, spart = my_sidx - 8 * sfull;
                for (int k = 0; k < my_ns; k++, sp += (uint32_t)my_sidx) {
                    const uint32_t v = (uint32_t)packed_value(peek64(words, nwords, sp), my_sidx, sfull, spart);
                    uint32_t r2, r1;
                    const uint32_t qq = divmod32(v, sizesmall, rcpsmall, r2);
                    const uint32_t q2 = divmod32(qq, sizesmall, rcpsmall, r1);
                    cur[0] = (int32_t)q2 + prev[0] - smallnum;
                    cur[1] = (int32_t)r1 + prev[1] - smallnum;
                    cur[2] = (int32_t)r2 + prev[2] - smallnum;
                    if (k == 0) {
                        // the first two atoms of a run are stored in swapped order
                        emit(my_first, cur[0], cur[1], cur[2]);
                        emit(a++, prev[0], prev[1], prev[2]);
                    } else {
                        emit(a++, cur[0], cur[1], cur[2]);
                    }
                    for (int d = 0; d < 3; d++) prev[d] = cur[d];
                }
            } else {
                BitStream bs;
                bs.init_at(sbase, send, (int64_t)sp);
                for (int k = 0; k < my_ns; k++) {
                    const Triple t = unpack3_wide(bs, my_sidx, sizesmall, sizesmall);
                    for (int d = 0; d < 3; d++) cur[d] = t.c[d] + prev[d] - smallnum;
                    if (k == 0) {
                        emit(my_first, cur[0], cur[1], cur[2]);
                        emit(a++, prev[0], prev[1], prev[2]);
                    } else {
                        emit(a++, cur[0], cur[1], cur[2]);
                    }
                    for (int d = 0; d < 3; d++) prev[d] = cur[d];
                }
            }
        }
    };

    // The descriptors of the thread's groups and the three stream words behind each are fetched up front, all
    // loads in flight together (a descriptor and then its words are two dependent trips to memory; taken one group
    // at a time they were a quarter of the kernel's stall samples), and parked in shared memory for the loop below.
    constexpr int kIters = kGroupsPerCta / kGroupThreads;
    __shared__ uint32_t st_pos[kIters][kGroupThreads], st_meta[kIters][kGroupThreads];
    __shared__ uint32_t st_a[kIters][kGroupThreads], st_b[kIters][kGroupThreads], st_c[kIters][kGroupThreads];
    {
        uint2 d[kIters];
#pragma unroll
        for (int it = 0; it < kIters; it++) {
            const int g = g0 + it * kGroupThreads + (int)threadIdx.x;
            d[it] = g < g1 ? desc[g] : make_uint2(0u, 0u);
        }
        uint32_t wa[kIters], wb[kIters], wc[kIters];
#pragma unroll
        for (int it = 0; it < kIters; it++) {
            const uint32_t w = d[it].x >> 5;
            wa[it] = w < nwords ? __ldg(words + w) : 0u;
            wb[it] = w + 1 < nwords ? __ldg(words + w + 1) : 0u;
            wc[it] = w + 2 < nwords ? __ldg(words + w + 2) : 0u;
        }
#pragma unroll
        for (int it = 0; it < kIters; it++) {
            st_pos[it][threadIdx.x] = d[it].x; st_meta[it][threadIdx.x] = d[it].y;
            st_a[it][threadIdx.x] = __byte_perm(wa[it], 0, 0x0123);
            st_b[it][threadIdx.x] = __byte_perm(wb[it], 0, 0x0123);
            st_c[it][threadIdx.x] = __byte_perm(wc[it], 0, 0x0123);
        }
        // (every thread reads back only what it wrote itself: no barrier)
    }

    for (int gb = g0, it = 0; gb < g1; gb += kGroupThreads, it++) {      // whole warps stay in the loop: the queue is theirs
        const int g = gb + (int)threadIdx.x;
        bool park = false;
        uint32_t sp = 0, meta = 0;
        int32_t cur[3] = {0, 0, 0};
        if (g < g1) {
            const uint32_t pos = st_pos[it][threadIdx.x];
            meta = st_meta[it][threadIdx.x];
            const uint32_t wa = st_a[it][threadIdx.x], wb = st_b[it][threadIdx.x], wc = st_c[it][threadIdx.x];
            const int my_first = (int)(meta & 0x1fffffu), my_ns = (int)((meta >> 21) & 15u);
            if (big_fast) {
                // value = (c0 * s1 + c1) * s2 + c2
                const uint64_t v = packed_value(peek64_words(wa, wb, wc, pos), bitsize, bfull, bpart);
                uint32_t r2, r1;
                const uint64_t qq = divmod64(v, sizeint[2], rcp2, r2);
                const uint64_t q2 = divmod64(qq, sizeint[1], rcp1, r1);
                cur[2] = (int32_t)r2; cur[1] = (int32_t)r1; cur[0] = (int32_t)q2;
            } else {
                BitStream bs;
                bs.init_at(sbase, send, (int64_t)pos);
                if (bitsize == 0) {
                    for (int d = 0; d < 3; d++) cur[d] = (int32_t)bs.get(bitsint[d]);
                } else {
                    const Triple t = unpack3_wide(bs, bitsize, sizeint[1], sizeint[2]);
                    for (int d = 0; d < 3; d++) cur[d] = t.c[d];
                }
            }
            for (int d = 0; d < 3; d++) cur[d] += minint[d];
            if (my_ns == 0) {
                emit(my_first, cur[0], cur[1], cur[2]);
            } else {
                // the flag behind the full-size code, five run bits if it is set, then the small codes
                park = true;
                sp = pos + (uint32_t)bigbits;
                uint32_t flag;
                if ((pos & 31u) + (uint32_t)bigbits < 96u) {
                    flag = bit_of_words(wa, wb, wc, (pos & 31u) + (uint32_t)bigbits);
                } else {      // (full-size codes of more than 64 bits)
                    const uint32_t w = sp >> 5;
                    const uint32_t fw = w < nwords ? __byte_perm(__ldg(words + w), 0, 0x0123) : 0u;
                    flag = (fw >> (31 - (sp & 31))) & 1u;
                }
                sp += flag ? 6u : 1u;
            }
        }
        const unsigned parked = __ballot_sync(0xffffffffu, park);
        if (park) {
            const int at = queued + __popc(parked & ((1u << lane) - 1u));
            qp[at] = sp; qm[at] = meta;
            q_prev[warp][0][at] = cur[0]; q_prev[warp][1][at] = cur[1]; q_prev[warp][2][at] = cur[2];
        }
        queued += __popc(parked);
        __syncwarp();
        if (queued >= 32) {
            // the last 32 entries (the queue holds at most 63); what stays is its front
            decode_runs(queued - 32, 32);
            __syncwarp();
            queued -= 32;
        }
    }
    if (queued > 0) decode_runs(0, queued);
}

}  // namespace

extern "C" int conan_xtc_decode_device(const uint8_t* d_data, int64_t nbytes, const int64_t* d_offsets,
                                       int64_t nframes, int32_t natoms, const int32_t* d_slot, int32_t nsel,
                                       float* d_xyz, int32_t* d_status, void* stream) {
    if (!d_data || !d_offsets || !d_xyz || !d_status || nbytes < 0 || nframes < 0 || natoms < 0 || nsel < 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device: bad arguments");
    if ((reinterpret_cast<uintptr_t>(d_data) & 3) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device: d_data must be 4-byte aligned");
    if (!d_slot && nsel != natoms)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device: without a slot map nsel must equal natoms");
    if (nframes == 0) return CONAN_OK;
    DecodeParams prm{d_data, nbytes, d_offsets, nframes, natoms, d_slot, nsel, d_xyz, d_status};
    const int64_t per_block = kDecodeThreads / 32;
    const unsigned blocks = (unsigned)((nframes + per_block - 1) / per_block);
    xtc_decode_kernel<<<blocks, kDecodeThreads, 0, static_cast<cudaStream_t>(stream)>>>(prm);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

namespace {

int walk_setup() {
    // the walk's windows need more than the default 48 KB of shared memory: once per device
    static std::atomic<uint64_t> configured{0};
    int dev = 0;
    CONAN_CUDA_TRY(cudaGetDevice(&dev));
    const uint64_t bit = 1ull << (dev & 63);
    if (!(configured.load(std::memory_order_acquire) & bit)) {
        CONAN_CUDA_TRY(cudaFuncSetAttribute(xtc_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWalkSmem));
        configured.fetch_or(bit, std::memory_order_release);
    }
    return CONAN_OK;
}

// pass 1 over n frames whose records / descriptors start at `rec` / `desc`
int launch_walk(const DecodeParams& d, FrameRec* rec, uint2* desc, cudaStream_t st) {
    WalkParams wp;
    wp.d = d; wp.rec = rec; wp.desc = desc;
    const int per = 32 * kWalkWarps;
    xtc_walk_kernel<<<(unsigned)((d.nframes + per - 1) / per), per, kWalkSmem, st>>>(wp);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}

// pass 2 over the same n frames
int launch_groups(const DecodeParams& d, FrameRec* rec, uint2* desc, cudaStream_t st) {
    if (d.natoms <= 9) return CONAN_OK;        // frames of plain floats were written by pass 1
    for (int64_t f0 = 0; f0 < d.nframes; f0 += 65535) {
        const int64_t n = std::min<int64_t>(65535, d.nframes - f0);
        WalkParams wp;
        wp.d = d;
        wp.d.offsets = d.offsets ? d.offsets + f0 : nullptr;
        wp.d.nframes = n;
        wp.d.xyz = d.xyz + (size_t)f0 * d.nsel * 3;
        wp.d.status = d.status ? d.status + f0 : nullptr;
        wp.rec = rec + f0;
        wp.desc = desc + (size_t)f0 * d.natoms;
        const dim3 grid((unsigned)((d.natoms + kGroupsPerCta - 1) / kGroupsPerCta), (unsigned)n);
        xtc_groups_kernel<<<grid, kGroupThreads, 0, st>>>(wp);
        CONAN_CUDA_TRY(cudaGetLastError());
    }
    return CONAN_OK;
}

}  // namespace

extern "C" size_t conan_xtc_decode_workspace_bytes(int64_t nframes, int32_t natoms) {
    if (nframes <= 0 || natoms <= 0) return 0;
    return (size_t)nframes * (sizeof(FrameRec) + (size_t)natoms * sizeof(uint2));
}

extern "C" int conan_xtc_decode_device_ws(const uint8_t* d_data, int64_t nbytes, const int64_t* d_offsets,
                                          int64_t nframes, int32_t natoms, const int32_t* d_slot, int32_t nsel,
                                          float* d_xyz, int32_t* d_status, void* d_work, size_t work_bytes,
                                          void* stream) {
    if (!d_data || !d_offsets || !d_xyz || !d_status || nbytes < 0 || nframes < 0 || natoms < 0 || nsel < 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device_ws: bad arguments");
    if ((reinterpret_cast<uintptr_t>(d_data) & 3) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device_ws: d_data must be 4-byte aligned");
    if (!d_slot && nsel != natoms)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device_ws: without a slot map nsel must equal natoms");
    if (nframes == 0) return CONAN_OK;
    if (natoms > kMaxWsAtoms)     // the descriptors hold 21 bits of atom index: larger systems take the one-pass kernel
        return conan_xtc_decode_device(d_data, nbytes, d_offsets, nframes, natoms, d_slot, nsel, d_xyz, d_status, stream);
    const size_t per_frame = sizeof(FrameRec) + (size_t)natoms * sizeof(uint2);
    if (!d_work || (reinterpret_cast<uintptr_t>(d_work) & 15) != 0 || work_bytes < per_frame)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode_device_ws: the workspace must be 16-byte aligned and hold at "
                                          "least one frame (%zu bytes)", per_frame);
    // as many frames per pass as the workspace holds (records first, descriptors behind them)
    const int64_t chunk = (int64_t)std::min<size_t>((size_t)nframes, work_bytes / per_frame);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = walk_setup();
    if (rc != CONAN_OK) return rc;
    FrameRec* rec = static_cast<FrameRec*>(d_work);
    uint2* desc = reinterpret_cast<uint2*>(static_cast<unsigned char*>(d_work) + (size_t)chunk * sizeof(FrameRec));
    for (int64_t f0 = 0; f0 < nframes; f0 += chunk) {
        const int64_t n = std::min(chunk, nframes - f0);
        const DecodeParams d{d_data, nbytes, d_offsets + f0, n, natoms, d_slot, nsel, d_xyz + (size_t)f0 * nsel * 3, d_status + f0};
        if ((rc = launch_walk(d, rec, desc, st)) != CONAN_OK) return rc;
        if ((rc = launch_groups(d, rec, desc, st)) != CONAN_OK) return rc;
    }
    return CONAN_OK;
}

// The two passes as separate calls, for a pipeline that runs the walk of a batch early (it only needs the bytes, and
// it is a latency-bound kernel of few warps that fits beside other kernels) and the decode when the coordinate buffer
// is free.  The workspace must hold all nframes (conan_xtc_decode_workspace_bytes) and stay untouched in between.
extern "C" int conan_xtc_walk_device(const uint8_t* d_data, int64_t nbytes, const int64_t* d_offsets, int64_t nframes,
                                     int32_t natoms, const int32_t* d_slot, int32_t nsel, float* d_xyz,
                                     int32_t* d_status, void* d_work, size_t work_bytes, void* stream) {
    if (!d_data || !d_offsets || !d_xyz || !d_status || nbytes < 0 || nframes < 0 || natoms <= 0 || nsel < 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_walk_device: bad arguments");
    if ((reinterpret_cast<uintptr_t>(d_data) & 3) != 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_walk_device: d_data must be 4-byte aligned");
    if (!d_slot && nsel != natoms)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_walk_device: without a slot map nsel must equal natoms");
    if (natoms > kMaxWsAtoms)
        return conan::fail(CONAN_ERR_SIZE, "conan_xtc_walk_device: %d atoms per frame; the two-pass decoder takes up to %d "
                                           "(use conan_xtc_decode_device)", natoms, kMaxWsAtoms);
    if (nframes == 0) return CONAN_OK;
    const size_t per_frame = sizeof(FrameRec) + (size_t)natoms * sizeof(uint2);
    if (!d_work || (reinterpret_cast<uintptr_t>(d_work) & 15) != 0 || work_bytes / per_frame < (size_t)nframes)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_walk_device: the workspace must be 16-byte aligned and hold all %lld "
                                          "frames (%zu bytes each)", (long long)nframes, per_frame);
    int rc = walk_setup();
    if (rc != CONAN_OK) return rc;
    const DecodeParams d{d_data, nbytes, d_offsets, nframes, natoms, d_slot, nsel, d_xyz, d_status};
    return launch_walk(d, static_cast<FrameRec*>(d_work),
                       reinterpret_cast<uint2*>(static_cast<unsigned char*>(d_work) + (size_t)nframes * sizeof(FrameRec)),
                       static_cast<cudaStream_t>(stream));
}

extern "C" int conan_xtc_groups_device(const uint8_t* d_data, int64_t nbytes, int64_t nframes, int32_t natoms,
                                       const int32_t* d_slot, int32_t nsel, float* d_xyz, const void* d_work,
                                       size_t work_bytes, void* stream) {
    if (!d_data || !d_xyz || nbytes < 0 || nframes < 0 || natoms <= 0 || nsel < 0 || natoms > kMaxWsAtoms)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_groups_device: bad arguments");
    if (!d_slot && nsel != natoms)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_groups_device: without a slot map nsel must equal natoms");
    if (nframes == 0) return CONAN_OK;
    const size_t per_frame = sizeof(FrameRec) + (size_t)natoms * sizeof(uint2);
    if (!d_work || (reinterpret_cast<uintptr_t>(d_work) & 15) != 0 || work_bytes / per_frame < (size_t)nframes)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_groups_device: the workspace of the walk over the same %lld frames is "
                                          "expected", (long long)nframes);
    unsigned char* w = static_cast<unsigned char*>(const_cast<void*>(d_work));
    const DecodeParams d{d_data, nbytes, nullptr, nframes, natoms, d_slot, nsel, d_xyz, nullptr};
    return launch_groups(d, reinterpret_cast<FrameRec*>(w), reinterpret_cast<uint2*>(w + (size_t)nframes * sizeof(FrameRec)),
                         static_cast<cudaStream_t>(stream));
}
