// GPU-side .xtc decoding: the compressed trajectory bytes go to the device as they
// are on disk and every frame is decoded there, so the host-to-device copy carries ~3 bytes per atom
// instead of 12 and the host cores are out of the stage-1 feed (`gmx mdmat -f`, conan.py:1437-1453).
//
// A frame's bit stream (public xdrfile xdr3dfcoord scheme) is one dependent chain -- the width of
// each code depends on the codes before it -- so the unit of parallelism is the frame: one thread
// walks one frame, a warp covers 32 frames.  Streams are read as aligned big-endian 32-bit words
// (XDR pads every frame to 4 bytes); only the atoms of the index group are written.  The arithmetic
// is the integer arithmetic of the host decoder in xtc.cpp, so the coordinates are bit-identical.
#include <cuda_runtime.h>

#include <cstdint>

#include "common.h"

namespace {

constexpr int kFirstIdx = 9;
constexpr int kNumMagic = 73;
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
    return __byte_perm(*reinterpret_cast<const uint32_t*>(p), 0, 0x0123);
}

// MSB-first bit stream over aligned 32-bit words; reads past `end` give zeros and are reported.
struct BitStream {
    const uint8_t* p;
    const uint8_t* end;
    uint64_t window;   // the next bits, left-aligned
    int have;
    bool overrun;
    __device__ __forceinline__ void init(const uint8_t* begin, const uint8_t* e) {
        p = begin; end = e; window = 0; have = 0; overrun = false;
    }
    __device__ __forceinline__ uint32_t get(int nbits) {   // 0 <= nbits <= 32
        if (have < nbits) {
            uint32_t w = 0;
            if (p < end) w = load_be32(p);
            else overrun = true;
            p += 4;
            window |= (uint64_t)w << (32 - have);
            have += 32;
        }
        const uint32_t v = nbits ? (uint32_t)(window >> (64 - nbits)) : 0u;
        window <<= nbits;
        have -= nbits;
        return v;
    }
};

// The packed-integer layout of xdrfile's receiveints: full bytes least significant first, then the
// remaining high bits.
__device__ __forceinline__ uint64_t get_packed64(BitStream& bs, int nbits) {   // nbits <= 64
    uint64_t v = 0;
    int shift = 0;
    while (nbits > 8) { v |= (uint64_t)bs.get(8) << shift; shift += 8; nbits -= 8; }
    if (nbits > 0) v |= (uint64_t)bs.get(nbits) << shift;
    return v;
}

// value = (c0 * s1 + c1) * s2 + c2
__device__ __forceinline__ void unpack3(BitStream& bs, int nbits, uint32_t s1, uint32_t s2, int32_t out[3]) {
    if (nbits <= 32) {
        const uint32_t v = (uint32_t)get_packed64(bs, nbits);
        const uint32_t q = v / s2;
        out[2] = (int32_t)(v - q * s2);
        const uint32_t q2 = q / s1;
        out[1] = (int32_t)(q - q2 * s1);
        out[0] = (int32_t)q2;
    } else if (nbits <= 64) {
        const uint64_t v = get_packed64(bs, nbits);
        const uint64_t q = v / s2;
        out[2] = (int32_t)(v - q * s2);
        const uint64_t q2 = q / s1;
        out[1] = (int32_t)(q - q2 * s1);
        out[0] = (int32_t)q2;
    } else {
        // up to 72 bits (three sizes below 2^24): long division over bytes, most significant first
        uint32_t bytes[12];
        int nb = 0, left = nbits;
        while (left > 8) { bytes[nb++] = bs.get(8); left -= 8; }
        if (left > 0) bytes[nb++] = bs.get(left);
        const uint32_t div[2] = {s2, s1};
#pragma unroll
        for (int d = 0; d < 2; d++) {
            uint32_t num = 0;
            for (int k = nb - 1; k >= 0; k--) {
                num = (num << 8) | bytes[k];
                const uint32_t qd = num / div[d];
                bytes[k] = qd;
                num -= qd * div[d];
            }
            out[2 - d] = (int32_t)num;
        }
        out[0] = (int32_t)(bytes[0] | (bytes[1] << 8) | (bytes[2] << 16) | (bytes[3] << 24));
    }
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

__global__ void __launch_bounds__(64) xtc_decode_kernel(const DecodeParams prm) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= prm.nframes) return;
    const int64_t off = prm.offsets[f];
    const int natoms = prm.natoms;
    float* out = prm.xyz + (size_t)f * prm.nsel * 3;
    auto fail_frame = [&](int code) { prm.status[f] = code; };
    if (off < 0 || (off & 3) || off + 56 > prm.nbytes) return fail_frame(kBadHeader);
    const uint8_t* p = prm.data + off;
    if ((int)load_be32(p) != 1995 || (int)load_be32(p + 4) != natoms) return fail_frame(kBadHeader);
    const uint8_t* q = p + 56;
    if (natoms <= 9) {   // stored as plain floats
        if (off + 56 + (int64_t)natoms * 12 > prm.nbytes) return fail_frame(kBadHeader);
        for (int a = 0; a < natoms; a++) {
            const int s = prm.slot ? prm.slot[a] : a;
            if (s >= 0)
                for (int d = 0; d < 3; d++) out[3 * (size_t)s + d] = __uint_as_float(load_be32(q + 12 * a + 4 * d));
        }
        prm.status[f] = kOk;
        return;
    }
    if (off + 92 > prm.nbytes) return fail_frame(kBadHeader);
    const float precision = __uint_as_float(load_be32(q));
    int32_t minint[3];
    uint32_t sizeint[3];
    for (int d = 0; d < 3; d++) {
        minint[d] = (int32_t)load_be32(q + 4 + 4 * d);
        sizeint[d] = (uint32_t)((int32_t)load_be32(q + 16 + 4 * d) - minint[d] + 1);
    }
    int smallidx = (int32_t)load_be32(q + 28);
    const int32_t nb = (int32_t)load_be32(q + 32);
    if (smallidx < kFirstIdx || smallidx >= kNumMagic) return fail_frame(kBadSmallIdx);
    if (nb < 0 || off + 92 + (((int64_t)nb + 3) & ~(int64_t)3) > prm.nbytes) return fail_frame(kBadHeader);
    int bitsint[3] = {0, 0, 0}, bitsize = 0;
    if ((sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffffu) {
        for (int d = 0; d < 3; d++) bitsint[d] = bits_for(sizeint[d]);
    } else {
        // bit length of the product of the three sizes (each below 2^24: at most 72 bits)
        const uint64_t p01 = (uint64_t)sizeint[0] * sizeint[1];
        const uint64_t lo = p01 * sizeint[2], hi = __umul64hi(p01, (uint64_t)sizeint[2]);
        bitsize = hi ? 128 - __clzll((long long)hi) : (lo ? 64 - __clzll((long long)lo) : 0);
    }
    int smaller = c_magic[max(kFirstIdx, smallidx - 1)] / 2;
    int smallnum = c_magic[smallidx] / 2;
    uint32_t sizesmall = (uint32_t)c_magic[smallidx];

    BitStream bs;
    bs.init(q + 36, q + 36 + (((size_t)nb + 3) & ~(size_t)3));
    // the host decoder multiplies by the rounded reciprocal of the precision; so does this one
    const float inv = 1.0f / precision;
    const int32_t* slot = prm.slot;
    auto emit = [&](int atom, const int32_t c[3]) {
        const int s = slot ? slot[atom] : atom;
        if (s >= 0) {
            float* o = out + 3 * (size_t)s;
            o[0] = (float)c[0] * inv; o[1] = (float)c[1] * inv; o[2] = (float)c[2] * inv;
        }
    };
    int run = 0, i = 0;
    int32_t cur[3], prev[3];
    while (i < natoms) {
        if (bitsize == 0) {
            for (int d = 0; d < 3; d++) cur[d] = (int32_t)bs.get(bitsint[d]);
        } else {
            unpack3(bs, bitsize, sizeint[1], sizeint[2], cur);
        }
        for (int d = 0; d < 3; d++) { cur[d] += minint[d]; prev[d] = cur[d]; }
        const int first_atom = i++;
        int is_smaller = 0;
        if (bs.get(1)) {
            run = (int)bs.get(5);
            is_smaller = run % 3;
            run -= is_smaller;
            is_smaller--;
        }
        if (run > 0) {
            if (i + run / 3 > natoms) return fail_frame(kRunPastEnd);
            for (int k = 0; k < run; k += 3) {
                unpack3(bs, smallidx, sizesmall, sizesmall, cur);
                for (int d = 0; d < 3; d++) cur[d] += prev[d] - smallnum;
                if (k == 0) {
                    // the first two atoms of a run are stored in swapped order
                    for (int d = 0; d < 3; d++) { const int32_t t = cur[d]; cur[d] = prev[d]; prev[d] = t; }
                    emit(first_atom, prev);
                } else {
                    for (int d = 0; d < 3; d++) prev[d] = cur[d];
                }
                emit(i++, cur);
            }
        } else {
            emit(first_atom, cur);
        }
        smallidx += is_smaller;
        if (smallidx < kFirstIdx || smallidx >= kNumMagic) return fail_frame(kBadSmallIdx);
        if (is_smaller < 0) {
            smallnum = smaller;
            smaller = smallidx > kFirstIdx ? c_magic[smallidx - 1] / 2 : 0;
        } else if (is_smaller > 0) {
            smaller = smallnum;
            smallnum = c_magic[smallidx] / 2;
        }
        sizesmall = (uint32_t)c_magic[smallidx];
    }
    prm.status[f] = bs.overrun ? kStreamPastEnd : kOk;
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
    const int threads = 64;
    const unsigned blocks = (unsigned)((nframes + threads - 1) / threads);
    xtc_decode_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(prm);
    CONAN_CUDA_TRY(cudaGetLastError());
    return CONAN_OK;
}
