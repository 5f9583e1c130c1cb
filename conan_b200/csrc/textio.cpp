// Host-side text writers of the file contract: matrices/NNNNN.dat
// (conan.py:570-597) and the multi-frame dmf.xpm that `gmx mdmat -frames` writes and
// read_index / read_next_time / read_frame parse (conan.py:455-567).  Both are formatted
// straight from the uint16 level stack on many host threads: every level's text is produced once
// ("%9.4f" of its legend value, or its xpm code) and lines are assembled by copying.
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdint>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.h"

namespace {

using conan::fail;

// Level of the pair (x-residue i, y-residue j) of one frame row; -1 on the diagonal of a symmetric run.
inline int level_at(const uint16_t* row, bool symmetric, int64_t nres, int yres, int i, int j) {
    if (!symmetric) return row[(int64_t)i * yres + j];
    if (i == j) return -1;
    const int64_t a = std::min(i, j), b = std::max(i, j);
    return row[a * nres - a * (a + 1) / 2 + (b - a - 1)];
}

bool write_all(int fd, const char* p, size_t n, off_t off, bool positioned) {
    while (n) {
        const ssize_t w = positioned ? pwrite(fd, p, n, off) : write(fd, p, n);
        if (w < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        p += w;
        n -= (size_t)w;
        off += w;
    }
    return true;
}

std::string fmt(const char* f, double v) {
    char b[64];
    snprintf(b, sizeof b, f, v);
    return b;
}
std::string fmt_int(const char* f, int v) {
    char b[32];
    snprintf(b, sizeof b, f, v);
    return b;
}

// Frames dealt out to nthreads workers.  The error text is thread-local (conan_last_error), so the first
// failing worker's message is carried back to the calling thread.
template <class Fn>
int run_frames(int64_t nf, int nthreads, Fn&& body) {
    const int nt = (int)std::max<int64_t>(1, std::min<int64_t>(nthreads > 0 ? nthreads : 1, nf));
    std::atomic<int64_t> next(0);
    std::atomic<int> status(CONAN_OK);
    std::mutex mu;
    std::string message;
    auto work = [&]() {
        for (;;) {
            const int64_t f = next.fetch_add(1);
            if (f >= nf || status.load() != CONAN_OK) return;
            const int rc = body(f);
            if (rc != CONAN_OK) {
                std::lock_guard<std::mutex> lock(mu);
                if (status.load() == CONAN_OK) {
                    message = conan::error_buffer();
                    status.store(rc);
                }
            }
        }
    };
    if (nt == 1) {
        work();
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; t++) pool.emplace_back(work);
        for (auto& th : pool) th.join();
    }
    if (status.load() != CONAN_OK) return fail(status.load(), "%s", message.c_str());
    return CONAN_OK;
}

}  // namespace

extern "C" int conan_write_matrices(const uint16_t* h_levels, int64_t levels_stride, int64_t nf, int64_t file_index0,
                                    int32_t symmetric, int32_t xterm, int32_t xres, int32_t yterm, int32_t yres,
                                    const double* values, int32_t nlevels, int32_t dr_mode,
                                    const uint16_t* h_first_row, const uint16_t* h_prev_row, const char* directory,
                                    int32_t nthreads) {
    if (!h_levels || !values || !directory || nf < 0 || xres <= 0 || yres <= 0 || nlevels < 2)
        return fail(CONAN_ERR_ARG, "conan_write_matrices: bad arguments");
    if (symmetric && xres != yres) return fail(CONAN_ERR_ARG, "conan_write_matrices: symmetric needs xres == yres");
    if (dr_mode && file_index0 > 0 && (!h_first_row || !h_prev_row))
        return fail(CONAN_ERR_ARG, "conan_write_matrices: dr files of a later batch need the first and previous rows");
    // text of every level's distance and of the diagonal zero, "%9.4f" as conan.py:595 prints it
    std::vector<std::string> vtext(nlevels);
    for (int l = 0; l < nlevels; l++) vtext[l] = fmt("%9.4f", values[l]) + "\n";
    const std::string zero = fmt("%9.4f", 0.0) + "\n";
    std::vector<std::string> xpre(xres), ypre(yres);
    for (int i = 0; i < xres; i++) xpre[i] = fmt_int("%4d ", xterm + i);
    for (int j = 0; j < yres; j++) ypre[j] = fmt_int("%4d ", yterm + j);
    const std::string head = "# i  j   r_ij\n# r_ij is the shortest distance between any atom of residues  i, j.\n";
    const std::string head_dr =
        "# i  j   dr_ij\n# dr_ij is the change (compared to the previous frame)\n"
        "# in shortest distance between any atom of residues  i, j.\n";
    const std::string dir(directory);
    const bool sym = symmetric != 0;

    auto dump = [&](const std::string& path, const std::string& text) -> int {
        const int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
        if (fd < 0) return fail(CONAN_ERR_IO, "cannot open %s: %s", path.c_str(), strerror(errno));
        const bool ok = write_all(fd, text.data(), text.size(), 0, false);
        close(fd);
        return ok ? CONAN_OK : fail(CONAN_ERR_IO, "short write to %s", path.c_str());
    };

    return run_frames(nf, nthreads, [&](int64_t f) -> int {
        const uint16_t* row = h_levels + f * levels_stride;
        char name[64];
        std::string text;
        text.reserve(head.size() + (size_t)xres * ((size_t)yres * 22 + 1));
        text = head;
        for (int i = 0; i < xres; i++) {
            for (int j = 0; j < yres; j++) {
                const int l = level_at(row, sym, xres, yres, i, j);
                text += xpre[i];
                text += ypre[j];
                text += l < 0 ? zero : vtext[std::min(l, nlevels - 1)];
            }
            text += '\n';
        }
        snprintf(name, sizeof name, "/%05lld.dat", (long long)(file_index0 + f));
        int rc = dump(dir + name, text);
        if (rc != CONAN_OK) return rc;
        if (!dr_mode || file_index0 + f == 0) return CONAN_OK;
        // change of the distance since the first / the previous frame (conan.py:829-838)
        for (int which = 1; which <= 2; which++) {
            if (!(dr_mode & which)) continue;
            const uint16_t* ref = which == 1 ? (file_index0 > 0 ? h_first_row : h_levels)
                                             : (f > 0 ? row - levels_stride : h_prev_row);
            text = head_dr;
            char b[48];
            for (int i = 0; i < xres; i++) {
                for (int j = 0; j < yres; j++) {
                    const int l = level_at(row, sym, xres, yres, i, j), l0 = level_at(ref, sym, xres, yres, i, j);
                    const double d = l < 0 ? 0.0 : values[std::min(l, nlevels - 1)] - values[std::min(l0, nlevels - 1)];
                    text += xpre[i];
                    text += ypre[j];
                    snprintf(b, sizeof b, "%9.4f\n", d);
                    text += b;
                }
                text += '\n';
            }
            snprintf(name, sizeof name, which == 1 ? "/%05lld_dr.init.dat" : "/%05lld_dr.prev.dat",
                     (long long)(file_index0 + f));
            rc = dump(dir + name, text);
            if (rc != CONAN_OK) return rc;
        }
        return CONAN_OK;
    });
}

extern "C" int conan_write_xpm(const uint16_t* h_levels, int64_t levels_stride, int64_t nf, int32_t symmetric,
                               int32_t xres, int32_t yres, const double* times, const char* legend_block,
                               const char* axes_block, const char* codes, int32_t nchar, int32_t nlevels,
                               const char* path, int32_t append, int32_t nthreads) {
    if (!h_levels || !times || !legend_block || !axes_block || !codes || !path || nf < 0 || xres <= 0 || yres <= 0 ||
        nchar < 1 || nchar > 2 || nlevels < 2)
        return fail(CONAN_ERR_ARG, "conan_write_xpm: bad arguments");
    if (symmetric && xres != yres) return fail(CONAN_ERR_ARG, "conan_write_xpm: symmetric needs xres == yres");
    const std::string legend(legend_block), axes(axes_block);
    const std::string head1 = "/* XPM */\n/* Generated by conan_b200 (gmx mdmat -frames layout) */\n";
    const std::string head2 =
        "/* legend:  \"Distance (nm)\" */\n/* x-label: \"Residue Index\" */\n"
        "/* y-label: \"Residue Index\" */\n/* type:    \"Continuous\" */\n";
    // frame sizes differ only in the title line; offsets first, then every thread writes its frames in place
    std::vector<std::string> titles(nf);
    std::vector<int64_t> offset(nf + 1, 0);
    const size_t body = (size_t)yres * ((size_t)xres * nchar + 4) - 1;   // rows "...",\n; the last one ends "\n
    for (int64_t f = 0; f < nf; f++) {
        titles[f] = "/* title:   \"t=" + fmt("%.0f", times[f]) + " ps\" */\n";
        offset[f + 1] = offset[f] + (int64_t)(head1.size() + titles[f].size() + head2.size() + legend.size() + axes.size() + body);
    }
    const int fd = open(path, O_WRONLY | O_CREAT | (append ? 0 : O_TRUNC), 0644);
    if (fd < 0) return fail(CONAN_ERR_IO, "cannot open %s: %s", path, strerror(errno));
    off_t base = 0;
    if (append) {
        struct stat st;
        if (fstat(fd, &st) != 0) { close(fd); return fail(CONAN_ERR_IO, "cannot stat %s", path); }
        base = st.st_size;
    }
    const bool sym = symmetric != 0;
    const int rc = run_frames(nf, nthreads, [&](int64_t f) -> int {
        const uint16_t* row = h_levels + f * levels_stride;
        std::string text;
        text.reserve((size_t)(offset[f + 1] - offset[f]));
        text = head1;
        text += titles[f];
        text += head2;
        text += legend;
        text += axes;
        // rows from the last y residue down to the first, x along the row (gmx write_xpm order)
        for (int j = yres - 1; j >= 0; j--) {
            text += '"';
            for (int i = 0; i < xres; i++) {
                int l = level_at(row, sym, xres, yres, i, j);
                if (l < 0) l = 0;
                l = std::min(l, nlevels - 1);
                text.append(codes + (size_t)l * nchar, (size_t)nchar);
            }
            text += j > 0 ? "\",\n" : "\"\n";
        }
        if ((int64_t)text.size() != offset[f + 1] - offset[f])
            return fail(CONAN_ERR_IO, "conan_write_xpm: frame %lld has %zu bytes, expected %lld", (long long)f, text.size(),
                        (long long)(offset[f + 1] - offset[f]));
        return write_all(fd, text.data(), text.size(), base + offset[f], true)
                   ? CONAN_OK : fail(CONAN_ERR_IO, "short write to %s", path);
    });
    close(fd);
    return rc;
}

// ---------------------------------------------------------------------------------------------
// residue-pair tables: one row per (i, j), a blank line after each i (the layout of
// aggregate/timeline.dat, mdmat_average_rmsf.dat, pearson_data.dat, pca/pc.N.dat, cluster maps ...)
// ---------------------------------------------------------------------------------------------
namespace {

struct Piece {
    std::string fmt;   // one printf conversion with the literal text that follows it
    bool integer;
};

// Split a printf row format into its conversions; integer conversions are widened to long long.
bool split_format(const char* format, std::vector<Piece>* out) {
    const std::string f(format);
    size_t pos = f.find('%');
    if (pos != 0) return false;
    while (pos != std::string::npos) {
        size_t next = f.find('%', pos + 1);
        std::string piece = f.substr(pos, next == std::string::npos ? std::string::npos : next - pos);
        // conversion character: first letter after the flags / width / precision
        size_t c = 1;
        while (c < piece.size() && std::strchr("-+ #0123456789.", piece[c])) c++;
        if (c >= piece.size()) return false;
        Piece p;
        const char conv = piece[c];
        if (conv == 'd' || conv == 'i') {
            p.integer = true;
            piece = piece.substr(0, c) + "lld" + piece.substr(c + 1);
        } else if (conv == 'f' || conv == 'e' || conv == 'g' || conv == 'E' || conv == 'G') {
            p.integer = false;
        } else {
            return false;
        }
        p.fmt = piece;
        out->push_back(p);
        pos = next;
    }
    return true;
}

}  // namespace

extern "C" int conan_write_pair_table(const char* path, const char* header, int32_t xterm, int32_t xres, int32_t yterm,
                                      int32_t yres, const double* const* columns, int32_t ncol, const char* row_format,
                                      int32_t nthreads) {
    if (!path || !row_format || xres <= 0 || yres <= 0 || ncol < 0 || (ncol > 0 && !columns))
        return fail(CONAN_ERR_ARG, "conan_write_pair_table: bad arguments");
    std::vector<Piece> pieces;
    if (!split_format(row_format, &pieces) || (int)pieces.size() != ncol + 2 || !pieces[0].integer || !pieces[1].integer)
        return fail(CONAN_ERR_ARG, "conan_write_pair_table: row format \"%s\" does not hold two integer and %d value conversions",
                    row_format, ncol);
    std::vector<std::string> ytext(yres);
    {
        char b[64];
        for (int j = 0; j < yres; j++) {
            snprintf(b, sizeof b, pieces[1].fmt.c_str(), (long long)(yterm + j));
            ytext[j] = b;
        }
    }
    std::vector<std::string> blocks(xres);
    const int rc = run_frames(xres, nthreads, [&](int64_t i) -> int {
        char b[96];
        snprintf(b, sizeof b, pieces[0].fmt.c_str(), (long long)(xterm + i));
        const std::string itext(b);
        std::string& text = blocks[i];
        text.reserve((size_t)yres * (itext.size() + 8 + 14 * (size_t)ncol) + 1);
        for (int j = 0; j < yres; j++) {
            text += itext;
            text += ytext[j];
            const int64_t e = i * yres + j;
            for (int c = 0; c < ncol; c++) {
                const double v = columns[c][e];
                if (pieces[c + 2].integer) snprintf(b, sizeof b, pieces[c + 2].fmt.c_str(), (long long)v);
                else snprintf(b, sizeof b, pieces[c + 2].fmt.c_str(), v);
                text += b;
            }
        }
        text += '\n';
        return CONAN_OK;
    });
    if (rc != CONAN_OK) return rc;
    const int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return fail(CONAN_ERR_IO, "cannot open %s: %s", path, strerror(errno));
    bool ok = true;
    if (header && *header) ok = write_all(fd, header, strlen(header), 0, false);
    for (int i = 0; ok && i < xres; i++) ok = write_all(fd, blocks[i].data(), blocks[i].size(), 0, false);
    close(fd);
    return ok ? CONAN_OK : fail(CONAN_ERR_IO, "short write to %s", path);
}
