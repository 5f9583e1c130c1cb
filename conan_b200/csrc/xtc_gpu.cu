// GPU-side .xtc decoding: the compressed trajectory bytes go to the device as they
// are on disk and every frame is decoded there, so the host-to-device copy carries ~3 bytes per atom
// instead of 12 and the host cores are out of the stage-1 feed (`gmx mdmat -f`, conan.py:1437-1453).
//
// A frame's bit stream (public xdrfile xdr3dfcoord scheme) is a chain of variable-width codes, but the
// widths depend only on the run flags between the codes, so a warp takes one frame: it walks the
// flags (cheap, sequential), then decodes 32 groups of atoms at a time in parallel, one per lane.
// Streams are read as aligned big-endian 32-bit words (XDR pads every frame to 4 bytes); only the
// atoms of the index group are written.  The arithmetic is the integer arithmetic of the host decoder
// in xtc.cpp, so the coordinates are bit-identical.
#include <cuda_runtime.h>

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
// sits off the decode chain; reads past `end` give zeros and are reported.
struct BitStream {
    const uint8_t* p;      // address of the word after `next`
    const uint8_t* end;
    uint64_t window;       // the next bits, left-aligned
    uint32_t next;         // the word that follows the window
    int have;
    __device__ __forceinline__ void init(const uint8_t* begin, const uint8_t* e) {
        p = begin; end = e; window = 0; have = 0;
        next = (p < end) ? load_be32(p) : 0u;
        p += 4;
    }
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
    // true if bits beyond the end of the stream were consumed
    __device__ __forceinline__ bool past_end() const {
        return p - ((have + 32) >> 3) > end;   // p minus the whole unread bytes (window + `next`)
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
