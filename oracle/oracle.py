"""Python face of the CPU oracle: ctypes wrapper around conan_oracle.c plus numpy restatements.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs,
never by conan_b200/.  Each function names the reference code it follows
(/root/reference/conan.py:line).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libconan_oracle.so")
_lib = None


def build():
    """Compile conan_oracle.c with the committed Makefile (gcc, no FMA contraction)."""
    subprocess.run(["make", "-s", "-C", _HERE], check=True)
    return _LIB


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "conan_oracle.c")):
            build()
        _lib = ctypes.CDLL(_LIB)
        vp = ctypes.c_void_p
        _lib.orc_xtc_count.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int)]
        _lib.orc_xtc_decode_all.argtypes = [vp, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, vp, vp, vp, vp]
        _lib.orc_mdmat_levels.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_int, ctypes.c_int,
                                          ctypes.c_float, ctypes.c_int, ctypes.c_int, vp]
        _lib.orc_pair_stats.argtypes = [vp, ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int] + [vp] * 13
        _lib.orc_pair_stats_shard.argtypes = [vp, ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, vp, vp, vp] + [vp] * 14
    return _lib


def xtc_decode(data):
    """(times[nf], boxes[nf,3,3], xyz[nf,natoms,3]) of an in-memory .xtc."""
    lib = load()
    buf = np.frombuffer(data, dtype=np.uint8)
    na = ctypes.c_int()
    nf = lib.orc_xtc_count(buf.ctypes.data, buf.size, ctypes.byref(na))
    if nf < 0:
        raise ValueError("malformed xtc")
    steps = np.zeros(nf, np.int32)
    times = np.zeros(nf, np.float32)
    boxes = np.zeros((nf, 3, 3), np.float32)
    xyz = np.zeros((nf, na.value, 3), np.float32)
    if lib.orc_xtc_decode_all(buf.ctypes.data, buf.size, na.value, nf, steps.ctypes.data, times.ctypes.data,
                              boxes.ctypes.data, xyz.ctypes.data) != 0:
        raise ValueError("xtc decode failed")
    return times, boxes, xyz


def mdmat_levels(xyz, boxes, sel, res_start, trunc, nlevels, pbc=0):
    """gmx mdmat restatement: uint16 [nf, nres, nres] level matrices (conan.py:1437-1453 -> gmx)."""
    lib = load()
    xyz = np.ascontiguousarray(xyz, np.float32)
    nf, natoms = xyz.shape[0], xyz.shape[1]
    boxes = np.ascontiguousarray(boxes if boxes is not None else np.zeros((nf, 3, 3)), np.float32)
    sel = np.ascontiguousarray(sel, np.int32)
    res_start = np.asarray(res_start)
    nres = len(res_start) - 1
    res_of = np.ascontiguousarray(np.repeat(np.arange(nres), np.diff(res_start)), np.int32)
    out = np.zeros((nf, nres, nres), np.uint16)
    rc = lib.orc_mdmat_levels(xyz.ctypes.data, boxes.ctypes.data, nf, natoms, sel.ctypes.data, res_of.ctypes.data,
                              sel.size, nres, trunc, nlevels, pbc, out.ctypes.data)
    if rc != 0:
        raise ValueError("orc_mdmat_levels failed: %d" % rc)
    return out


def condensed(full):
    """[nf, nres, nres] -> [nf, P] in the C-ABI's symmetric pair order (i<j row-major)."""
    nres = full.shape[1]
    iu = np.triu_indices(nres, 1)
    return np.ascontiguousarray(full[:, iu[0], iu[1]])


def block(full, x0, xres, y0, yres):
    """[nf, nres, nres] -> [nf, xres*yres] block order (conan.py:541-551 picks the same entries)."""
    return np.ascontiguousarray(full[:, x0:x0 + xres, y0:y0 + yres].reshape(full.shape[0], -1))


def legend_values(trunc, nlevels):
    """read_index, conan.py:474-481: repeated addition, last level pinned to trunc."""
    vals = []
    dx = trunc / (nlevels - 1)
    value = 0.0
    for _ in range(nlevels - 1):
        vals.append(value)
        value = value + dx
    vals.append(trunc)
    return np.array(vals, dtype=np.float64)


def pair_stats(levels, lo_cnt, hi_first, equal_cut, frame0=0, first_row=None, prev_row=None, status_in=None):
    """C restatement of the frame loop (conan.py:782-874) over a [nf, P] level stack.

    The keyword arguments describe a shard of a longer trajectory (see orc_pair_stats_shard)."""
    lib = load()
    levels = np.ascontiguousarray(levels, np.uint16)
    nf, P = levels.shape
    keep = [None if a is None else np.ascontiguousarray(a, dt) for a, dt in
            ((first_row, np.uint16), (prev_row, np.uint16), (status_in, np.uint8))]
    ptr = [None if a is None else a.ctypes.data for a in keep]
    o = dict(sum_k=np.zeros(P, np.int64), sum_k2=np.zeros(P, np.int64), n_int=np.zeros(P, np.int32),
             n_enc=np.zeros(P, np.int32), first_f=np.zeros(P, np.int32), last_f=np.zeros(P, np.int32),
             status=np.zeros(P, np.uint8), kmin=np.zeros(P, np.uint16), kmax=np.zeros(P, np.uint16),
             native=np.zeros(nf, np.int32), nonnative=np.zeros(nf, np.int32),
             ssd_prev=np.zeros(nf, np.int64), ssd_first=np.zeros(nf, np.int64), decided=np.zeros(P, np.uint8))
    order = ["sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f", "status", "kmin", "kmax",
             "native", "nonnative", "ssd_prev", "ssd_first", "decided"]
    rc = lib.orc_pair_stats_shard(levels.ctypes.data, nf, P, lo_cnt, hi_first, equal_cut, frame0, *ptr,
                                  *[o[k].ctypes.data for k in order])
    if rc != 0:
        raise ValueError("orc_pair_stats failed: %d" % rc)
    return o
