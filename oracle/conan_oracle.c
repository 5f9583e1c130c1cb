/*
 * conan_oracle.c -- CPU restatement of the CONAN trajectory contact path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in conan_b200/ may import, link or call this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker or the timed CPU baseline.
 *
 * What it restates, and where the behaviour is specified:
 *
 *  Stage 1 (distance matrices).  The reference shells out to `gmx mdmat`
 *  (/root/reference/conan.py:1437-1453, invoked :1783-1784) and consumes the
 *  resulting dmf.xpm (:455-567).  GROMACS is a third-party program that is NOT
 *  under /root/reference (un-vendored, un-pinned; README.md:7 asks for
 *  "version 5.0 or later"; the fixtures were written by 4.6.5 / 5.0.7), so
 *  this file restates its published algorithm:
 *    - .xtc frames: XDR big-endian header + the xdr3dfcoord bit-packed integer
 *      coordinate scheme of the public xdrfile library (magicints table,
 *      FIRSTIDX 9, run-length "small" deltas, first/second atom swap);
 *    - gmx mdmat: per frame, for every residue pair the minimum over atom pairs
 *      of the (minimum-image) Euclidean distance in `real` = float32, truncated
 *      at `trunc`; written through write_xpm as level
 *      c = rint(r * (nlevels-1)/trunc) clamped to [0, nlevels-1].
 *  Parity is PINNED by the reference's own golden files
 *  test/1_ReadWrite/correct_output/matrices/00000..00010.dat (11 x 28 x 28 levels,
 *  dx = 0.001) and test/8_Asymmetric/correct_output/matrices (tests/test_oracle.py).
 *  Minimum-image PBC and -b/-e/-dt frame selection are unpinned by any fixture.
 *
 *  Stage 2 (per-pair statistics): conan.py:24-45 (update_interaction),
 *  :599-611 (update_aggregate), :782-874 (frame loop), :1414-1432 (rmsd_frame).
 *  Restated here over the quantised level stack (the reference sees the same
 *  information after parsing the xpm text).  Pinned against the live reference
 *  (oracle/ref_runner.py) and the golden aggregate tables.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off; FMA contraction must stay
 * off so that r2 = dx*dx + dy*dy + dz*dz rounds after every operation).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ */
/* XTC reader (xdrfile's xdr3dfcoord, read direction)                  */
/* ------------------------------------------------------------------ */

static const int magicints[] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 8, 10, 12, 16, 20, 25, 32, 40, 50, 64,
    80, 101, 128, 161, 203, 256, 322, 406, 512, 645, 812, 1024, 1290,
    1625, 2048, 2580, 3250, 4096, 5060, 6501, 8192, 10321, 13003,
    16384, 20642, 26007, 32768, 41285, 52015, 65536, 82570, 104031,
    131072, 165140, 208063, 262144, 330280, 416127, 524287, 660561,
    832255, 1048576, 1321122, 1664510, 2097152, 2642245, 3329021,
    4194304, 5284491, 6658042, 8388607, 10568983, 13316085, 16777216};
#define FIRSTIDX 9

typedef struct {
    const unsigned char *p;
    size_t cnt;
    unsigned int lastbits;
    unsigned int lastbyte;
} bitreader;

static int sizeofint(int size) {
    unsigned int num = 1;
    int nbits = 0;
    while ((unsigned int)size >= num && nbits < 32) {
        nbits++;
        num <<= 1;
    }
    return nbits;
}

static int sizeofints(int n, const unsigned int sizes[]) {
    unsigned int bytes[32], nbytes = 1, nbits = 0, bytecnt, tmp;
    unsigned int num;
    bytes[0] = 1;
    for (int i = 0; i < n; i++) {
        tmp = 0;
        for (bytecnt = 0; bytecnt < nbytes; bytecnt++) {
            tmp = bytes[bytecnt] * sizes[i] + tmp;
            bytes[bytecnt] = tmp & 0xff;
            tmp >>= 8;
        }
        while (tmp != 0) {
            bytes[bytecnt++] = tmp & 0xff;
            tmp >>= 8;
        }
        nbytes = bytecnt;
    }
    num = 1;
    nbytes--;
    while (bytes[nbytes] >= num) {
        nbits++;
        num *= 2;
    }
    return (int)(nbits + nbytes * 8);
}

static int receivebits(bitreader *b, int nbits) {
    int mask = (1 << nbits) - 1;
    int num = 0;
    unsigned int lastbits = b->lastbits, lastbyte = b->lastbyte;
    while (nbits >= 8) {
        lastbyte = (lastbyte << 8) | b->p[b->cnt++];
        num |= (lastbyte >> lastbits) << (nbits - 8);
        nbits -= 8;
    }
    if (nbits > 0) {
        if (lastbits < (unsigned int)nbits) {
            lastbits += 8;
            lastbyte = (lastbyte << 8) | b->p[b->cnt++];
        }
        lastbits -= nbits;
        num |= (lastbyte >> lastbits) & ((1 << nbits) - 1);
    }
    num &= mask;
    b->lastbits = lastbits;
    b->lastbyte = lastbyte;
    return num;
}

static void receiveints(bitreader *b, int n, int nbits, const unsigned int sizes[], int nums[]) {
    int bytes[32];
    int nbytes = 0, i, j, p, num;
    bytes[1] = bytes[2] = bytes[3] = 0;
    while (nbits > 8) {
        bytes[nbytes++] = receivebits(b, 8);
        nbits -= 8;
    }
    if (nbits > 0) bytes[nbytes++] = receivebits(b, nbits);
    for (i = n - 1; i > 0; i--) {
        num = 0;
        for (j = nbytes - 1; j >= 0; j--) {
            /* unsigned as in xdrfile (sizes[] is unsigned there): sizes up to 2^24 make num << 8 reach 2^32 */
            num = (int)(((unsigned int)num << 8) | (unsigned int)bytes[j]);
            p = (int)((unsigned int)num / sizes[i]);
            bytes[j] = p;
            num = (int)((unsigned int)num - (unsigned int)p * sizes[i]);
        }
        nums[i] = num;
    }
    nums[0] = bytes[0] | (bytes[1] << 8) | (bytes[2] << 16) | (bytes[3] << 24);
}

static int32_t be32(const unsigned char *p) {
    return (int32_t)(((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]);
}
static float bef(const unsigned char *p) {
    int32_t v = be32(p);
    float f;
    memcpy(&f, &v, 4);
    return f;
}

/* Size in bytes of the frame starting at `p`, or 0 on a malformed / truncated frame. */
static size_t xtc_frame_size(const unsigned char *p, size_t avail, int *natoms_out) {
    if (avail < 56) return 0;
    if (be32(p) != 1995) return 0;
    int natoms = be32(p + 4);
    if (natoms_out) *natoms_out = natoms;
    /* magic, natoms, step, time, box[9], natoms = 14 words = 56 bytes */
    if (natoms <= 9) {
        size_t s = 56 + (size_t)natoms * 12;
        return s <= avail ? s : 0;
    }
    /* precision, minint[3], maxint[3], smallidx, nbytes = 9 words */
    if (avail < 56 + 36) return 0;
    int nbytes = be32(p + 56 + 32);
    size_t s = 56 + 36 + (((size_t)nbytes + 3) & ~(size_t)3);
    return s <= avail ? s : 0;
}

/* Count frames in an in-memory .xtc; returns the count (or -1), natoms of frame 0. */
int orc_xtc_count(const unsigned char *data, size_t n, int *natoms) {
    size_t off = 0;
    int count = 0, na = 0;
    while (off < n) {
        int a;
        size_t s = xtc_frame_size(data + off, n - off, &a);
        if (s == 0) return -1;
        if (count == 0) na = a;
        else if (a != na) return -1;
        off += s;
        count++;
    }
    if (natoms) *natoms = na;
    return count;
}

/* Decode one frame.  xyz: natoms*3 floats (nm).  Returns bytes consumed, 0 on error. */
static size_t xtc_decode_frame(const unsigned char *p, size_t avail, int natoms_expected,
                               int *step, float *time, float *box, float *xyz) {
    int natoms;
    size_t fsz = xtc_frame_size(p, avail, &natoms);
    if (fsz == 0 || natoms != natoms_expected) return 0;
    *step = be32(p + 8);
    *time = bef(p + 12);
    for (int i = 0; i < 9; i++) box[i] = bef(p + 16 + 4 * i);
    if (be32(p + 52) != natoms) return 0;
    const unsigned char *q = p + 56;
    if (natoms <= 9) {
        for (int i = 0; i < natoms * 3; i++) xyz[i] = bef(q + 4 * i);
        return fsz;
    }
    float precision = bef(q);
    int minint[3], maxint[3];
    unsigned int sizeint[3], sizesmall[3], bitsizeint[3] = {0, 0, 0};
    for (int i = 0; i < 3; i++) minint[i] = be32(q + 4 + 4 * i);
    for (int i = 0; i < 3; i++) maxint[i] = be32(q + 16 + 4 * i);
    int smallidx = be32(q + 28);
    int bitsize;
    for (int i = 0; i < 3; i++) sizeint[i] = (unsigned int)(maxint[i] - minint[i] + 1);
    if ((sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffff) {
        for (int i = 0; i < 3; i++) bitsizeint[i] = (unsigned int)sizeofint((int)sizeint[i]);
        bitsize = 0;
    } else {
        bitsize = sizeofints(3, sizeint);
    }
    int tmp = smallidx - 1;
    tmp = (FIRSTIDX > tmp) ? FIRSTIDX : tmp;
    int smaller = magicints[tmp] / 2;
    int smallnum = magicints[smallidx] / 2;
    sizesmall[0] = sizesmall[1] = sizesmall[2] = (unsigned int)magicints[smallidx];

    bitreader br = {q + 36, 0, 0, 0};
    float inv_precision = 1.0f / precision;
    int run = 0, i = 0;
    float *lfp = xyz;
    int thiscoord[3], prevcoord[3];
    while (i < natoms) {
        if (bitsize == 0) {
            thiscoord[0] = receivebits(&br, (int)bitsizeint[0]);
            thiscoord[1] = receivebits(&br, (int)bitsizeint[1]);
            thiscoord[2] = receivebits(&br, (int)bitsizeint[2]);
        } else {
            receiveints(&br, 3, bitsize, sizeint, thiscoord);
        }
        i++;
        thiscoord[0] += minint[0];
        thiscoord[1] += minint[1];
        thiscoord[2] += minint[2];
        prevcoord[0] = thiscoord[0];
        prevcoord[1] = thiscoord[1];
        prevcoord[2] = thiscoord[2];

        int flag = receivebits(&br, 1);
        int is_smaller = 0;
        if (flag == 1) {
            run = receivebits(&br, 5);
            is_smaller = run % 3;
            run -= is_smaller;
            is_smaller--;
        }
        if (run > 0) {
            for (int k = 0; k < run; k += 3) {
                receiveints(&br, 3, smallidx, sizesmall, thiscoord);
                i++;
                thiscoord[0] += prevcoord[0] - smallnum;
                thiscoord[1] += prevcoord[1] - smallnum;
                thiscoord[2] += prevcoord[2] - smallnum;
                if (k == 0) {
                    /* first and second atom of a run are stored swapped */
                    int t;
                    t = thiscoord[0]; thiscoord[0] = prevcoord[0]; prevcoord[0] = t;
                    t = thiscoord[1]; thiscoord[1] = prevcoord[1]; prevcoord[1] = t;
                    t = thiscoord[2]; thiscoord[2] = prevcoord[2]; prevcoord[2] = t;
                    *lfp++ = prevcoord[0] * inv_precision;
                    *lfp++ = prevcoord[1] * inv_precision;
                    *lfp++ = prevcoord[2] * inv_precision;
                } else {
                    prevcoord[0] = thiscoord[0];
                    prevcoord[1] = thiscoord[1];
                    prevcoord[2] = thiscoord[2];
                }
                *lfp++ = thiscoord[0] * inv_precision;
                *lfp++ = thiscoord[1] * inv_precision;
                *lfp++ = thiscoord[2] * inv_precision;
            }
        } else {
            *lfp++ = thiscoord[0] * inv_precision;
            *lfp++ = thiscoord[1] * inv_precision;
            *lfp++ = thiscoord[2] * inv_precision;
        }
        smallidx += is_smaller;
        if (is_smaller < 0) {
            smallnum = smaller;
            if (smallidx > FIRSTIDX) smaller = magicints[smallidx - 1] / 2;
            else smaller = 0;
        } else if (is_smaller > 0) {
            smaller = smallnum;
            smallnum = magicints[smallidx] / 2;
        }
        sizesmall[0] = sizesmall[1] = sizesmall[2] = (unsigned int)magicints[smallidx];
    }
    return fsz;
}

/* Decode every frame of an in-memory .xtc.  Arrays sized by orc_xtc_count. */
int orc_xtc_decode_all(const unsigned char *data, size_t n, int natoms, int nframes,
                       int *steps, float *times, float *boxes, float *xyz) {
    size_t off = 0;
    for (int f = 0; f < nframes; f++) {
        size_t s = xtc_decode_frame(data + off, n - off, natoms, &steps[f], &times[f],
                                    boxes + 9 * (size_t)f, xyz + 3 * (size_t)natoms * f);
        if (s == 0) return -1;
        off += s;
    }
    return 0;
}

/* ------------------------------------------------------------------ */
/* Stage 1: gmx mdmat restatement                                      */
/* ------------------------------------------------------------------ */

/* Rectangular minimum image, the way gmx pbc_dx treats epbcdxRECTANGULAR:
 * shift by whole box lengths until the component lies in (-L/2, L/2]. */
static inline float min_image(float d, float L) {
    if (L > 0.0f) {
        float h = 0.5f * L;
        while (d > h) d -= L;
        while (d <= -h) d += L;
    }
    return d;
}

/*
 * xyz    [nf][natoms][3] float32 (nm), box [nf][9] (row-major box vectors)
 * sel    [nsel] indices into the trajectory's atoms (the index group)
 * res_of [nsel] residue index 0..nres-1 of each selected atom
 * pbc    0 = plain distances; 1 = rectangular minimum image with box diagonal
 * out    [nf][nres][nres] uint16 level indices (symmetric, diagonal 0)
 * Arithmetic follows gmx mdmat: float32 throughout, no FMA contraction.
 */
int orc_mdmat_levels(const float *xyz, const float *box, int nf, int natoms,
                     const int *sel, const int *res_of, int nsel, int nres,
                     float trunc, int nlevels, int pbc, uint16_t *out) {
    if (nlevels < 2 || nlevels > 65535 || nres <= 0) return -1;
    float *r2min = (float *)malloc(sizeof(float) * (size_t)nres * nres);
    if (!r2min) return -2;
    const float trunc2 = trunc * trunc;
    const float invlevel = (float)(nlevels - 1) / trunc;
    for (int f = 0; f < nf; f++) {
        const float *x = xyz + 3 * (size_t)natoms * f;
        const float *b = box + 9 * (size_t)f;
        float Lx = pbc ? b[0] : 0.0f, Ly = pbc ? b[4] : 0.0f, Lz = pbc ? b[8] : 0.0f;
        for (size_t k = 0; k < (size_t)nres * nres; k++) r2min[k] = trunc2;
        for (int a = 0; a < nsel; a++) {
            const float *xa = x + 3 * (size_t)sel[a];
            int ra = res_of[a];
            for (int c = a + 1; c < nsel; c++) {
                int rc = res_of[c];
                if (rc == ra) continue;
                const float *xc = x + 3 * (size_t)sel[c];
                float dx = min_image(xa[0] - xc[0], Lx);
                float dy = min_image(xa[1] - xc[1], Ly);
                float dz = min_image(xa[2] - xc[2], Lz);
                float r2 = dx * dx + dy * dy + dz * dz;
                int lo = ra < rc ? ra : rc, hi = ra < rc ? rc : ra;
                float *m = &r2min[(size_t)lo * nres + hi];
                if (r2 < *m) *m = r2;
            }
        }
        uint16_t *o = out + (size_t)nres * nres * f;
        for (int i = 0; i < nres; i++) {
            o[(size_t)i * nres + i] = 0;
            for (int j = i + 1; j < nres; j++) {
                float r = sqrtf(r2min[(size_t)i * nres + j]);
                int c = (int)rintf(r * invlevel);
                if (c < 0) c = 0;
                if (c > nlevels - 1) c = nlevels - 1;
                o[(size_t)i * nres + j] = (uint16_t)c;
                o[(size_t)j * nres + i] = (uint16_t)c;
            }
        }
    }
    free(r2min);
    return 0;
}

/* ------------------------------------------------------------------ */
/* Stage 2: streaming per-pair statistics over a level stack           */
/* ------------------------------------------------------------------ */

/*
 * levels [nf][P] uint16, frame-major.  lo_cnt / hi_first are the level-space forms
 * of TRUNC_INTER / TRUNC_INTER_HIGH: r < trunc_inter  <=> k < lo_cnt,
 * r > trunc_inter_high <=> k >= hi_first (legend values are monotone in k).
 * equal_cut != 0 reproduces the `trunc_inter == trunc_inter_high` branch of
 * update_interaction (conan.py:29-30): status = (r < trunc_inter), history ignored.
 * Outputs (all caller-allocated):
 *   sum_k, sum_k2 [P] int64; n_int, n_enc, first_f, last_f [P] int32 (first_f=-1 if never on,
 *   last_f = -1 likewise); status [P] uint8 (final status); kmin,kmax [P] uint16;
 *   native, nonnative [nf] int32; ssd_prev, ssd_first [nf] int64 (sum over pairs of squared
 *   level differences to the previous / first frame; frame 0 gets 0).
 */
int orc_pair_stats_shard(const uint16_t *levels, int nf, int64_t P, int lo_cnt, int hi_first,
                         int equal_cut, int frame0, const uint16_t *first_row,
                         const uint16_t *prev_row, const uint8_t *status_in, int64_t *sum_k,
                         int64_t *sum_k2, int32_t *n_int, int32_t *n_enc, int32_t *first_f,
                         int32_t *last_f, uint8_t *status, uint16_t *kmin, uint16_t *kmax,
                         int32_t *native, int32_t *nonnative, int64_t *ssd_prev,
                         int64_t *ssd_first, uint8_t *decided);

int orc_pair_stats(const uint16_t *levels, int nf, int64_t P, int lo_cnt, int hi_first,
                   int equal_cut, int64_t *sum_k, int64_t *sum_k2, int32_t *n_int,
                   int32_t *n_enc, int32_t *first_f, int32_t *last_f, uint8_t *status,
                   uint16_t *kmin, uint16_t *kmax, int32_t *native, int32_t *nonnative,
                   int64_t *ssd_prev, int64_t *ssd_first) {
    return orc_pair_stats_shard(levels, nf, P, lo_cnt, hi_first, equal_cut, 0, NULL, NULL, NULL,
                                sum_k, sum_k2, n_int, n_enc, first_f, last_f, status, kmin, kmax,
                                native, nonnative, ssd_prev, ssd_first, NULL);
}

/*
 * The same loop for a shard of a longer trajectory: frame numbers start at frame0, the
 * interaction status entering the shard is status_in (NULL = off), "native" and the
 * rmsd-to-first reference come from first_row (NULL = this shard's first row), and prev_row
 * (NULL = none) is the frame before the shard.  decided[p] (optional) is set when some frame
 * of the shard was outside the hysteresis buffer.
 */
int orc_pair_stats_shard(const uint16_t *levels, int nf, int64_t P, int lo_cnt, int hi_first,
                         int equal_cut, int frame0, const uint16_t *first_row,
                         const uint16_t *prev_row, const uint8_t *status_in, int64_t *sum_k,
                         int64_t *sum_k2, int32_t *n_int, int32_t *n_enc, int32_t *first_f,
                         int32_t *last_f, uint8_t *status, uint16_t *kmin, uint16_t *kmax,
                         int32_t *native, int32_t *nonnative, int64_t *ssd_prev,
                         int64_t *ssd_first, uint8_t *decided) {
    uint8_t *nat = (uint8_t *)calloc((size_t)P, 1);
    if (!nat) return -2;
    if (!first_row) first_row = levels;
    if (equal_cut) hi_first = lo_cnt;
    for (int64_t p = 0; p < P; p++) {
        sum_k[p] = 0; sum_k2[p] = 0; n_int[p] = 0; n_enc[p] = 0;
        first_f[p] = -1; last_f[p] = -1; status[p] = status_in ? status_in[p] : 0;
        kmin[p] = 65535; kmax[p] = 0;
        nat[p] = (uint8_t)(first_row[p] < lo_cnt);
        if (decided) decided[p] = 0;
    }
    for (int f = 0; f < nf; f++) {
        const uint16_t *row = levels + (size_t)P * f;
        const uint16_t *prev = f ? row - P : (prev_row ? prev_row : row);
        int32_t nn = 0, nno = 0;
        int64_t sp = 0, s0 = 0;
        for (int64_t p = 0; p < P; p++) {
            int k = row[p];
            sum_k[p] += k;
            sum_k2[p] += (int64_t)k * k;
            if (k < kmin[p]) kmin[p] = (uint16_t)k;
            if (k > kmax[p]) kmax[p] = (uint16_t)k;
            int st = status[p], ns;
            if (equal_cut) ns = (k < lo_cnt);
            else if (k < lo_cnt) ns = 1;
            else if (k >= hi_first) ns = 0;
            else ns = st;
            if (decided && (k < lo_cnt || k >= hi_first)) decided[p] = 1;
            if (ns) {
                n_int[p]++;
                last_f[p] = frame0 + f;
                if (!st) {
                    if (first_f[p] < 0) first_f[p] = frame0 + f;
                    n_enc[p]++;
                }
            }
            status[p] = (uint8_t)ns;
            if (ns) { if (nat[p]) nn++; else nno++; }
            int64_t dp = (int64_t)k - prev[p], d0 = (int64_t)k - first_row[p];
            sp += dp * dp;
            s0 += d0 * d0;
        }
        native[f] = nn; nonnative[f] = nno; ssd_prev[f] = sp; ssd_first[f] = s0;
    }
    free(nat);
    return 0;
}
