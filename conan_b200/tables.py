"""Host-side finalisation of stage 2: integer statistics -> the reference's tables and files.

Follows the tail of read_process_frames (conan.py:890-959) and the per-frame lines written inside
its loop (:645, :854).  Inputs are the exact integer sums from the GPU (conan_b200.stage2) and the
legend values; everything here is O(nres^2) float64 arithmetic plus text formatting.
"""
import math

import numpy as np


class Layout:
    """Residue numbering of the maps: symmetric (nterm, nres) or asymmetric X x Y block."""

    def __init__(self, nres=None, nterm=1, xterm=None, xres=None, yterm=None, yres=None):
        self.asymm = xres is not None
        if self.asymm:
            self.xterm, self.xres, self.yterm, self.yres = int(xterm), int(xres), int(yterm), int(yres)
            self.nres = None
            self.P = self.xres * self.yres
        else:
            self.nres = int(nres)
            self.xterm = self.yterm = int(nterm)
            self.xres = self.yres = self.nres
            self.P = self.nres * (self.nres - 1) // 2

    def to_matrix(self, per_pair, diag=0):
        """[P] condensed -> [xres, yres] matrix (symmetric fill, `diag` on the diagonal)."""
        per_pair = np.asarray(per_pair)
        if self.asymm:
            return per_pair.reshape(self.xres, self.yres)
        m = np.full((self.nres, self.nres), diag, dtype=per_pair.dtype)
        iu = np.triu_indices(self.nres, 1)
        m[iu] = per_pair
        m[(iu[1], iu[0])] = per_pair
        return m

    def pair_labels(self):
        """(i, j) residue numbers (as used in pairs_list / the output files) of every pair p."""
        if self.asymm:
            pi, pj = np.divmod(np.arange(self.P), self.yres)
        else:
            pi, pj = np.triu_indices(self.nres, 1)
        return pi + self.xterm, pj + self.yterm


def write_pair_table(path, layout, columns, row_format, header=""):
    """One row per residue pair and a blank line after each x residue, formatted by the native threaded
    writer (conan_write_pair_table).  columns: [xres, yres] arrays, one per value conversion of row_format."""
    import ctypes
    import os
    from . import _lib
    cols = [np.ascontiguousarray(c, dtype=np.float64).reshape(layout.xres * layout.yres) for c in columns]
    ptrs = (ctypes.c_void_p * max(len(cols), 1))(*[c.ctypes.data for c in cols])
    lib = _lib.load()
    _lib.check(lib.conan_write_pair_table(os.fsencode(path), header.encode(), layout.xterm, layout.xres, layout.yterm,
                                          layout.yres, ptrs, len(cols), row_format.encode(),
                                          min(32, os.cpu_count() or 1)))


class Stage2Tables:
    """Everything read_process_frames derives after its frame loop."""

    def __init__(self, stats, values, time_vec, trunc, layout):
        self.layout = layout
        self.time_vec = [float(t) for t in time_vec]
        tn = len(self.time_vec)
        nl = len(values)
        dx = trunc / (nl - 1)
        t0, tmax = self.time_vec[0], self.time_vec[-1]
        self.t0, self.tmax, self.tn, self.trunc = t0, tmax, tn, trunc
        sum_k = stats["sum_k"].astype(np.int64)
        sum_k2 = stats["sum_k2"].astype(np.int64)
        n_int = stats["n_int"].astype(np.int64)
        times = np.asarray(self.time_vec)
        first_f, last_f = stats["first_f"], stats["last_f"]
        # conan.py:909-911: never on -> first = tmax; last stays 0 (empty_aggregate)
        self.first = np.where(first_f >= 0, times[np.clip(first_f, 0, tn - 1)], tmax)
        self.last = np.where(last_f >= 0, times[np.clip(last_f, 0, tn - 1)], 0.0)
        self.lifetime = n_int / tn                                    # :917
        self.n_enc = stats["n_enc"].astype(np.int64)
        self.n_int = n_int
        self.avg = dx * sum_k / tn                                    # :912
        # :913-916 with the cancellation done in exact integer arithmetic
        var_num = tn * sum_k2 - sum_k * sum_k
        self.std = dx * np.sqrt(np.maximum(var_num, 0).astype(np.float64)) / tn
        self.seen = sum_k < tn * (nl - 1)                             # pair ever below trunc (:536-537)
        self.varying = var_num != 0                                   # len(set(rvec)) > 1 (:1902)
        # per-frame lines
        self.native = stats["native"].astype(np.int64)
        self.nonnative = stats["nonnative"].astype(np.int64)
        if layout.asymm:                                              # :1429-1432
            self.rmsd_prev = np.sqrt(dx * dx * stats["ssd_prev"] / (layout.xres * layout.yres))
            self.rmsd_first = np.sqrt(dx * dx * stats["ssd_first"] / (layout.xres * layout.yres))
        else:
            self.rmsd_prev = np.sqrt(2.0 * dx * dx * stats["ssd_prev"]) / layout.nres
            self.rmsd_first = np.sqrt(2.0 * dx * dx * stats["ssd_first"]) / layout.nres
        self._interactions()

    def _interactions(self):
        lay = self.layout
        life = lay.to_matrix(self.lifetime, diag=0.0)
        if lay.asymm:                                                 # :937-947
            self.total_x = life.sum(axis=1)
            self.total_y = life.sum(axis=0)
        else:                                                         # :949-955, pairs with i < j-3 only (:817)
            n = lay.nres
            far = np.abs(np.subtract.outer(np.arange(n), np.arange(n))) > 3
            sel = life * far
            tot = sel.sum(axis=1)
            cnt = (sel > 0).sum(axis=1)
            self.local = np.where(tot > 0, tot / np.maximum(cnt, 1), tot)

    # ---- dictionaries of the `import conan` return tuple (conan.py:927-928, 941, 945, 955) -----
    def metric_dicts(self):
        lay = self.layout
        avg = lay.to_matrix(self.avg, diag=0.0)
        life = lay.to_matrix(self.lifetime, diag=1.0)
        avg_d, life_d = {}, {}
        for i in range(lay.xres):
            for j in range(lay.yres):
                if i == j and not lay.asymm:
                    continue
                avg_d[(i, j)] = float(avg[i, j])
                life_d[(i, j)] = float(1 - life[i, j])
        return avg_d, life_d

    def interact_dicts(self):
        lay = self.layout
        if lay.asymm:
            return ({lay.xterm + i: float(v) for i, v in enumerate(self.total_x)},
                    {lay.yterm + j: float(v) for j, v in enumerate(self.total_y)})
        return ({lay.xterm + i: float(v) for i, v in enumerate(self.local)},)

    # ---- files -------------------------------------------------------------------------------
    def write_aggregate(self, directory="aggregate"):
        lay = self.layout
        first = lay.to_matrix(self.first, diag=self.t0)
        last = lay.to_matrix(self.last, diag=self.tmax)
        life = lay.to_matrix(self.lifetime, diag=1.0)
        nenc = lay.to_matrix(self.n_enc, diag=1)
        avg = lay.to_matrix(self.avg, diag=0.0)
        std = lay.to_matrix(self.std, diag=0.0)
        span = self.tmax - self.t0
        write_pair_table(directory + "/timeline.dat", lay, [first, last, life, np.full_like(life, span), nenc],
                         "%4d %4d %12i %12i %12.6f %12.6f %5i\n",
                         "#  i  j   t_first t_last t_total (# encounters)\n"
                         "# first/last time residues i and j are within the interaction distance.\n"
                         "# total: fraction of frames in which residues i and j are within interaction distance (maximum = 1.0).\n")
        write_pair_table(directory + "/mdmat_average_rmsf.dat", lay, [avg, std], "%4d %4d %9.4f %9.4f\n",
                         "#  i  j  r_ij^avg sigma_(rij)\n"
                         "# average of the distance between residues i and j and the standard deviation (computed across frames.\n")
        with open(directory + "/native_contacts.dat", "w") as f:
            f.write("# time (ns)  #native #non-native\n")
            for t, a, b in zip(self.time_vec, self.native, self.nonnative):
                f.write("%9.4f %12.4f %12.4f\n" % (0.001 * t, a, b))
        with open(directory + "/time_mdmat_rmsd.dat", "w") as f:
            f.write("# time (ns)  RMSD_mdmat_previous RMSD_mdmat_total\n")
            for n in range(1, self.tn):
                f.write("%9.4f %9.4f %9.4f\n" % (0.001 * self.time_vec[n], self.rmsd_prev[n], self.rmsd_first[n]))
        if lay.asymm:
            with open(directory + "/total_interactions.x.dat", "w") as f:
                for i, v in enumerate(self.total_x):
                    f.write("%5i%12.6f\n" % (lay.xterm + i, v))
            with open(directory + "/total_interactions.y.dat", "w") as f:
                for j, v in enumerate(self.total_y):
                    f.write("%5i%12.6f\n" % (lay.yterm + j, v))
        else:
            with open(directory + "/local_interactions.dat", "w") as f:
                for i, v in enumerate(self.local):
                    f.write("%5i%12.6f\n" % (i + lay.xterm, v))
