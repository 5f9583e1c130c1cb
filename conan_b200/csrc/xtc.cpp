// Host-side .xtc reading for the stage-1 feed (frame index, threaded decode, simple encode).
//
// Takes over the trajectory reading `gmx mdmat -f` performs for the reference
// (conan.py:1437-1453).  The on-disk format is GROMACS' XDR .xtc: a big-endian header per
// frame followed by the xdr3dfcoord bit-packed integer coordinates (public xdrfile scheme).
// Frames are self-contained, so after one cheap header scan every frame can be decoded on
// its own thread straight into the (pinned) array that is copied to the GPU; only the atoms
// of the index group are written out.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <thread>
#include <vector>

#include "common.h"

namespace {

constexpr int kMagic = 1995;
constexpr int kFirstIdx = 9;
const int kMagicInts[] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 8, 10, 12, 16, 20, 25, 32, 40, 50, 64,
    80, 101, 128, 161, 203, 256, 322, 406, 512, 645, 812, 1024, 1290,
    1625, 2048, 2580, 3250, 4096, 5060, 6501, 8192, 10321, 13003,
    16384, 20642, 26007, 32768, 41285, 52015, 65536, 82570, 104031,
    131072, 165140, 208063, 262144, 330280, 416127, 524287, 660561,
    832255, 1048576, 1321122, 1664510, 2097152, 2642245, 3329021,
    4194304, 5284491, 6658042, 8388607, 10568983, 13316085, 16777216};
constexpr int kNumMagic = sizeof(kMagicInts) / sizeof(kMagicInts[0]);

inline int32_t be32(const uint8_t* p) {
    return (int32_t)(((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]);
}
inline float bef(const uint8_t* p) {
    int32_t v = be32(p);
    float f;
    std::memcpy(&f, &v, 4);
    return f;
}
inline void put32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v;
}
inline void putf(uint8_t* p, float f) {
    uint32_t v;
    std::memcpy(&v, &f, 4);
    put32(p, v);
}

// Bytes of the frame starting at p (0 if malformed / truncated).
size_t frame_size(const uint8_t* p, size_t avail, int32_t* natoms) {
    if (avail < 56 || be32(p) != kMagic) return 0;
    int32_t na = be32(p + 4);
    if (na < 0) return 0;
    *natoms = na;
    size_t s;
    if (na <= 9) {
        s = 56 + (size_t)na * 12;
    } else {
        if (avail < 92) return 0;
        int32_t nb = be32(p + 88);
        if (nb < 0) return 0;
        s = 92 + (((size_t)nb + 3) & ~(size_t)3);
    }
    return s <= avail ? s : 0;
}

// MSB-first bit stream over a byte range with a 64-bit window.
struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint64_t window = 0;  // valid bits are the top `have` bits
    int have = 0;
    inline void refill() {
        while (have <= 56) {
            uint64_t byte = (p < end) ? *p : 0;
            ++p;
            window |= byte << (56 - have);
            have += 8;
        }
    }
    inline uint32_t get(int nbits) {  // nbits <= 32
        if (nbits == 0) return 0;
        if (have < nbits) refill();
        uint32_t v = (uint32_t)(window >> (64 - nbits));
        window <<= nbits;
        have -= nbits;
        return v;
    }
};

// Packed triple: bytes are stored least-significant first, 8 bits each, the last chunk
// holding the remaining bits; value = (a*s1 + b)*s2 + c.
inline void unpack3(BitReader& br, int nbits, const uint32_t s[3], int32_t out[3]) {
    if (nbits <= 64) {
        uint64_t v = 0;
        int shift = 0, left = nbits;
        while (left > 8) { v |= (uint64_t)br.get(8) << shift; shift += 8; left -= 8; }
        if (left > 0) v |= (uint64_t)br.get(left) << shift;
        uint64_t q = v / s[2];
        out[2] = (int32_t)(v - q * s[2]);
        uint64_t q2 = q / s[1];
        out[1] = (int32_t)(q - q2 * s[1]);
        out[0] = (int32_t)q2;
    } else {
        unsigned __int128 v = 0;
        int shift = 0, left = nbits;
        while (left > 8) { v |= (unsigned __int128)br.get(8) << shift; shift += 8; left -= 8; }
        if (left > 0) v |= (unsigned __int128)br.get(left) << shift;
        unsigned __int128 q = v / s[2];
        out[2] = (int32_t)(uint64_t)(v - q * s[2]);
        unsigned __int128 q2 = q / s[1];
        out[1] = (int32_t)(uint64_t)(q - q2 * s[1]);
        out[0] = (int32_t)(uint64_t)q2;
    }
}

int bits_for(uint32_t size) {  // smallest n with 2^n > size... as xdrfile's sizeofint
    uint32_t num = 1;
    int n = 0;
    while (size >= num && n < 32) { n++; num <<= 1; }
    return n;
}

int bits_for_product(const uint32_t s[3]) {  // bits to hold s0*s1*s2 - 1 (xdrfile's sizeofints)
    unsigned __int128 prod = (unsigned __int128)s[0] * s[1] * s[2];
    // sizeofints counts bits of the product itself: number of bits n such that 2^n > prod's
    // top byte rule; equivalent to the bit length of prod.
    int n = 0;
    while (prod != 0) { n++; prod >>= 1; }
    return n;
}

// Decode one frame; writes only atoms listed in `slot` (slot[a] = output row or -1).
bool decode_frame(const uint8_t* p, size_t fsz, int32_t natoms, const int32_t* slot, float* out) {
    auto emit = [&](int atom, const int32_t c[3], float inv) {
        int s = slot ? slot[atom] : atom;
        if (s >= 0) {
            float* o = out + 3 * (size_t)s;
            o[0] = (float)c[0] * inv; o[1] = (float)c[1] * inv; o[2] = (float)c[2] * inv;
        }
    };
    const uint8_t* q = p + 56;
    if (natoms <= 9) {
        for (int a = 0; a < natoms; a++) {
            int s = slot ? slot[a] : a;
            if (s >= 0) for (int d = 0; d < 3; d++) out[3 * (size_t)s + d] = bef(q + 12 * a + 4 * d);
        }
        return true;
    }
    float precision = bef(q);
    int32_t minint[3], maxint[3];
    for (int d = 0; d < 3; d++) { minint[d] = be32(q + 4 + 4 * d); maxint[d] = be32(q + 16 + 4 * d); }
    int smallidx = be32(q + 28);
    if (smallidx < kFirstIdx || smallidx >= kNumMagic) return false;
    uint32_t sizeint[3], sizesmall[3];
    int bitsint[3] = {0, 0, 0}, bitsize;
    for (int d = 0; d < 3; d++) sizeint[d] = (uint32_t)(maxint[d] - minint[d] + 1);
    if ((sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffff) {
        for (int d = 0; d < 3; d++) bitsint[d] = bits_for(sizeint[d]);
        bitsize = 0;
    } else {
        bitsize = bits_for_product(sizeint);
    }
    int smaller = kMagicInts[std::max(kFirstIdx, smallidx - 1)] / 2;
    int smallnum = kMagicInts[smallidx] / 2;
    sizesmall[0] = sizesmall[1] = sizesmall[2] = (uint32_t)kMagicInts[smallidx];

    BitReader br{q + 36, p + fsz};
    const float inv = 1.0f / precision;
    int run = 0, i = 0;
    int32_t cur[3], prev[3];
    while (i < natoms) {
        if (bitsize == 0) {
            for (int d = 0; d < 3; d++) cur[d] = (int32_t)br.get(bitsint[d]);
        } else {
            unpack3(br, bitsize, sizeint, cur);
        }
        for (int d = 0; d < 3; d++) { cur[d] += minint[d]; prev[d] = cur[d]; }
        int first_atom = i++;
        int is_smaller = 0;
        if (br.get(1)) {
            run = (int)br.get(5);
            is_smaller = run % 3;
            run -= is_smaller;
            is_smaller--;
        }
        if (run > 0) {
            if (i + run / 3 > natoms) return false;
            for (int k = 0; k < run; k += 3) {
                unpack3(br, smallidx, sizesmall, cur);
                for (int d = 0; d < 3; d++) cur[d] += prev[d] - smallnum;
                if (k == 0) {
                    // the first two atoms of a run are stored in swapped order
                    for (int d = 0; d < 3; d++) std::swap(cur[d], prev[d]);
                    emit(first_atom, prev, inv);
                } else {
                    for (int d = 0; d < 3; d++) prev[d] = cur[d];
                }
                emit(i++, cur, inv);
            }
        } else {
            emit(first_atom, cur, inv);
        }
        smallidx += is_smaller;
        if (smallidx < kFirstIdx || smallidx >= kNumMagic) return false;
        if (is_smaller < 0) {
            smallnum = smaller;
            smaller = smallidx > kFirstIdx ? kMagicInts[smallidx - 1] / 2 : 0;
        } else if (is_smaller > 0) {
            smaller = smallnum;
            smallnum = kMagicInts[smallidx] / 2;
        }
        sizesmall[0] = sizesmall[1] = sizesmall[2] = (uint32_t)kMagicInts[smallidx];
    }
    return true;
}

// MSB-first bit writer.
struct BitWriter {
    std::vector<uint8_t>& buf;
    uint64_t acc = 0;
    int have = 0;
    void put(uint32_t v, int nbits) {
        if (nbits == 0) return;
        acc = (acc << nbits) | (v & (nbits == 32 ? 0xffffffffu : ((1u << nbits) - 1)));
        have += nbits;
        while (have >= 8) { buf.push_back((uint8_t)(acc >> (have - 8))); have -= 8; }
    }
    void flush() { if (have > 0) { buf.push_back((uint8_t)(acc << (8 - have))); have = 0; } }
};

}  // namespace

extern "C" int conan_xtc_index(const uint8_t* data, size_t nbytes, int64_t cap, int64_t* offsets,
                               int32_t* steps, float* times, float* boxes, int64_t* nframes,
                               int32_t* natoms) {
    if (!data || !nframes || !natoms) return conan::fail(CONAN_ERR_ARG, "conan_xtc_index: null argument");
    size_t off = 0;
    int64_t n = 0;
    int32_t na0 = 0;
    while (off < nbytes) {
        int32_t na;
        size_t s = frame_size(data + off, nbytes - off, &na);
        if (s == 0)
            return conan::fail(CONAN_ERR_FORMAT, "conan_xtc_index: malformed or truncated frame %lld at byte %zu",
                               (long long)n, off);
        if (n == 0) na0 = na;
        else if (na != na0)
            return conan::fail(CONAN_ERR_FORMAT, "conan_xtc_index: frame %lld has %d atoms, frame 0 has %d",
                               (long long)n, na, na0);
        if (offsets && n < cap) {
            offsets[n] = (int64_t)off;
            if (steps) steps[n] = be32(data + off + 8);
            if (times) times[n] = bef(data + off + 12);
            if (boxes) for (int i = 0; i < 9; i++) boxes[9 * n + i] = bef(data + off + 16 + 4 * i);
        }
        off += s;
        n++;
    }
    *nframes = n;
    *natoms = na0;
    return CONAN_OK;
}

extern "C" int conan_xtc_decode(const uint8_t* data, size_t nbytes, const int64_t* offsets,
                                const int64_t* frame_ids, int64_t nsel_frames, int32_t natoms,
                                const int32_t* atom_sel, int32_t nsel, float* xyz_out,
                                int32_t nthreads) {
    if (!data || !offsets || !xyz_out || nsel_frames < 0 || natoms <= 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode: bad argument");
    std::vector<int32_t> slot;
    int32_t nout = natoms;
    if (atom_sel) {
        if (nsel <= 0) return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode: empty atom selection");
        slot.assign((size_t)natoms, -1);
        for (int32_t s = 0; s < nsel; s++) {
            if (atom_sel[s] < 0 || atom_sel[s] >= natoms)
                return conan::fail(CONAN_ERR_ARG, "conan_xtc_decode: atom index %d out of range", atom_sel[s]);
            slot[(size_t)atom_sel[s]] = s;
        }
        nout = nsel;
    }
    int nt = std::max(1, std::min<int>(nthreads, (int)std::max<int64_t>(1, nsel_frames)));
    std::atomic<int64_t> next{0};
    std::atomic<int64_t> bad{-1};
    auto worker = [&]() {
        for (;;) {
            int64_t i = next.fetch_add(1);
            if (i >= nsel_frames || bad.load() >= 0) return;
            int64_t f = frame_ids ? frame_ids[i] : i;
            size_t off = (size_t)offsets[f];
            int32_t na = 0;
            size_t s = off < nbytes ? frame_size(data + off, nbytes - off, &na) : 0;
            if (s == 0 || na != natoms ||
                !decode_frame(data + off, s, natoms, atom_sel ? slot.data() : nullptr,
                              xyz_out + 3 * (size_t)nout * i)) {
                bad.store(f);
                return;
            }
        }
    };
    if (nt == 1) {
        worker();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; t++) pool.emplace_back(worker);
        for (auto& t : pool) t.join();
    }
    if (bad.load() >= 0)
        return conan::fail(CONAN_ERR_FORMAT, "conan_xtc_decode: frame %lld could not be decoded", (long long)bad.load());
    return CONAN_OK;
}

extern "C" int conan_xtc_encode(const float* xyz, const float* boxes, const int32_t* steps,
                                const float* times, int64_t nframes, int32_t natoms, float precision,
                                uint8_t* out, size_t cap, size_t* nbytes) {
    if (!xyz || !nbytes || nframes < 0 || natoms <= 0 || precision <= 0)
        return conan::fail(CONAN_ERR_ARG, "conan_xtc_encode: bad argument");
    size_t pos = 0;
    std::vector<uint8_t> bits;
    std::vector<int32_t> ints((size_t)natoms * 3);
    for (int64_t f = 0; f < nframes; f++) {
        const float* x = xyz + 3 * (size_t)natoms * f;
        uint8_t head[92];
        put32(head, kMagic);
        put32(head + 4, (uint32_t)natoms);
        put32(head + 8, (uint32_t)(steps ? steps[f] : (int32_t)f));
        putf(head + 12, times ? times[f] : (float)f);
        for (int i = 0; i < 9; i++) putf(head + 16 + 4 * i, boxes ? boxes[9 * f + i] : 0.0f);
        put32(head + 52, (uint32_t)natoms);
        size_t fsz;
        bits.clear();
        if (natoms <= 9) {
            fsz = 56 + (size_t)natoms * 12;
            if (out && pos + fsz <= cap) {
                std::memcpy(out + pos, head, 56);
                for (int i = 0; i < natoms * 3; i++) putf(out + pos + 56 + 4 * i, x[i]);
            }
        } else {
            int32_t mn[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, mx[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
            for (int a = 0; a < natoms; a++)
                for (int d = 0; d < 3; d++) {
                    float v = x[3 * (size_t)a + d] * precision;
                    if (std::fabs(v) > 2.0e9f)
                        return conan::fail(CONAN_ERR_SIZE, "conan_xtc_encode: coordinate too large for this precision");
                    int32_t iv = (int32_t)(v >= 0 ? v + 0.5f : v - 0.5f);
                    ints[3 * (size_t)a + d] = iv;
                    mn[d] = std::min(mn[d], iv);
                    mx[d] = std::max(mx[d], iv);
                }
            uint32_t sizeint[3];
            int bitsint[3] = {0, 0, 0}, bitsize;
            int smallidx, first_smallidx;
            for (int d = 0; d < 3; d++) sizeint[d] = (uint32_t)(mx[d] - mn[d] + 1);
            if ((sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffff) {
                for (int d = 0; d < 3; d++) bitsint[d] = bits_for(sizeint[d]);
                bitsize = 0;
            } else {
                bitsize = bits_for_product(sizeint);
            }
            BitWriter bw{bits};
            auto send3 = [&](int nb, const uint32_t sz[3], const uint32_t c[3]) {
                unsigned __int128 v = ((unsigned __int128)c[0] * sz[1] + c[1]) * sz[2] + c[2];
                int left = nb;
                while (left > 8) { bw.put((uint32_t)(v & 0xff), 8); v >>= 8; left -= 8; }
                if (left > 0) bw.put((uint32_t)(v & ((1u << left) - 1)), left);
            };
            // the xdr3dfcoord scheme: an atom in full, then a run of up to eight atoms as small
            // differences to their predecessor while they stay close (water molecules, side chains);
            // the width of the small codes adapts to how close successive atoms are
            int64_t mindiff = INT64_MAX;
            for (int a = 1; a < natoms; a++) {
                int64_t dsum = 0;
                for (int d = 0; d < 3; d++) dsum += std::llabs((int64_t)ints[3 * (size_t)a + d] - ints[3 * (size_t)(a - 1) + d]);
                mindiff = std::min(mindiff, dsum);
            }
            const int kLastIdx = kNumMagic - 1;
            smallidx = kFirstIdx;
            while (smallidx < kLastIdx && kMagicInts[smallidx] < mindiff) smallidx++;
            first_smallidx = smallidx;
            const int maxidx = std::min(kLastIdx, smallidx + 8), minidx = maxidx - 8;
            int smaller = kMagicInts[std::max(kFirstIdx, smallidx - 1)] / 2;
            int smallnum = kMagicInts[smallidx] / 2;
            const int larger = kMagicInts[maxidx] / 2;
            uint32_t sizesmall[3] = {(uint32_t)kMagicInts[smallidx], (uint32_t)kMagicInts[smallidx], (uint32_t)kMagicInts[smallidx]};
            auto within = [&](const int32_t* p, const int32_t* q, int lim) {
                for (int d = 0; d < 3; d++)
                    if (std::llabs((int64_t)p[d] - q[d]) >= lim) return false;
                return true;
            };
            int i = 0, prevrun = -1;
            int32_t prevcoord[3] = {0, 0, 0};
            uint32_t tmp[8 * 3];
            while (i < natoms) {
                int32_t* cur = &ints[3 * (size_t)i];
                int is_small = 0, is_smaller;
                if (smallidx < maxidx && i >= 1 && within(cur, prevcoord, larger)) is_smaller = 1;
                else if (smallidx > minidx) is_smaller = -1;
                else is_smaller = 0;
                if (i + 1 < natoms && within(cur, cur + 3, smallnum)) {
                    for (int d = 0; d < 3; d++) std::swap(cur[d], cur[3 + d]);   // the pair is stored swapped
                    is_small = 1;
                }
                uint32_t c[3];
                for (int d = 0; d < 3; d++) c[d] = (uint32_t)(cur[d] - mn[d]);
                if (bitsize == 0) {
                    for (int d = 0; d < 3; d++) bw.put(c[d], bitsint[d]);
                } else {
                    send3(bitsize, sizeint, c);
                }
                for (int d = 0; d < 3; d++) prevcoord[d] = cur[d];
                i++;
                cur += 3;
                int run = 0;
                if (is_small == 0 && is_smaller == -1) is_smaller = 0;
                while (is_small && run < 8 * 3) {
                    if (is_smaller == -1) {
                        int64_t d2 = 0;
                        for (int d = 0; d < 3; d++) { const int64_t dd = (int64_t)cur[d] - prevcoord[d]; d2 += dd * dd; }
                        if (d2 >= (int64_t)smaller * smaller) is_smaller = 0;
                    }
                    for (int d = 0; d < 3; d++) tmp[run++] = (uint32_t)(cur[d] - prevcoord[d] + smallnum);
                    for (int d = 0; d < 3; d++) prevcoord[d] = cur[d];
                    i++;
                    cur += 3;
                    is_small = (i < natoms && within(cur, prevcoord, smallnum)) ? 1 : 0;
                }
                if (run != prevrun || is_smaller != 0) {
                    prevrun = run;
                    bw.put(1, 1);
                    bw.put((uint32_t)(run + is_smaller + 1), 5);
                } else {
                    bw.put(0, 1);
                }
                for (int k = 0; k < run; k += 3) send3(smallidx, sizesmall, tmp + k);
                if (is_smaller != 0) {
                    smallidx += is_smaller;
                    if (is_smaller < 0) {
                        smallnum = smaller;
                        smaller = kMagicInts[std::max(kFirstIdx, smallidx - 1)] / 2;
                    } else {
                        smaller = smallnum;
                        smallnum = kMagicInts[smallidx] / 2;
                    }
                    sizesmall[0] = sizesmall[1] = sizesmall[2] = (uint32_t)kMagicInts[smallidx];
                }
            }
            bw.flush();
            putf(head + 56, precision);
            for (int d = 0; d < 3; d++) { put32(head + 60 + 4 * d, (uint32_t)mn[d]); put32(head + 72 + 4 * d, (uint32_t)mx[d]); }
            put32(head + 84, (uint32_t)first_smallidx);
            put32(head + 88, (uint32_t)bits.size());
            size_t padded = (bits.size() + 3) & ~(size_t)3;
            fsz = 92 + padded;
            if (out && pos + fsz <= cap) {
                std::memcpy(out + pos, head, 92);
                std::memcpy(out + pos + 92, bits.data(), bits.size());
                std::memset(out + pos + 92 + bits.size(), 0, padded - bits.size());
            }
        }
        pos += fsz;
    }
    *nbytes = pos;
    if (out && pos > cap) return conan::fail(CONAN_ERR_SIZE, "conan_xtc_encode: output needs %zu bytes, capacity %zu", pos, cap);
    return CONAN_OK;
}
