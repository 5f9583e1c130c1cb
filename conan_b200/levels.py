"""Level-space bookkeeping shared by the stages: legend values, cut-offs, pair indexing.

The reference never sees distances, only the legend value of an xpm level:
read_index (conan.py:455-482) builds dict_legend by repeated addition of dx = trunc/(nlevels-1)
and pins the last level to trunc.  Everything downstream compares those float64 values with the
cut-offs, so cut-offs translate exactly into level indices.
"""
import numpy as np

MAX_NLEVELS = 7744  # 88^2 two-character xpm codes (docs/index.md:161)


def legend_values(trunc, nlevels):
    """float64 [nlevels]: value of each level exactly as read_index accumulates it (:474-481)."""
    vals = np.empty(nlevels, dtype=np.float64)
    dx = trunc / (nlevels - 1)
    v = 0.0
    for k in range(nlevels - 1):
        vals[k] = v
        v = v + dx
    vals[nlevels - 1] = trunc
    return vals


def adjust_truncs(trunc, trunc_dr, trunc_inter, trunc_inter_high, quiet=False):
    """Defaults of the interaction cut-offs, as conan.py:1715-1729 (same messages)."""
    say = (lambda *a: None) if quiet else print
    if not trunc_dr:
        trunc_dr = trunc
        say("No distance cutoff given for differential plot, will use maximum cutoff:", trunc)
    if not trunc_inter or trunc_inter == trunc:
        trunc_inter = 0.999999 * trunc
        trunc_inter_high = 0.999999 * trunc
        say("No distance cutoff given for defining interactions, will use maximum cutoff:", trunc)
    elif not trunc_inter_high or trunc_inter_high == trunc:
        say("No buffer value given for the interaction cutoff. We will use the same value for both.")
        trunc_inter_high = trunc_inter
    if trunc_inter_high < trunc_inter:
        say("It must hold that TRUNC_INTER_HIGH >= TRUNC_INTER. We will set them to be equal.")
        trunc_inter_high = trunc_inter
    return trunc_dr, trunc_inter, trunc_inter_high


def level_thresholds(values, trunc_inter, trunc_inter_high):
    """(lo_cnt, hi_first, equal_cut): r < trunc_inter <=> k < lo_cnt; r > trunc_inter_high <=> k >= hi_first.

    update_interaction (conan.py:24-45) takes its first branch when both cut-offs are equal."""
    lo_cnt = int(np.searchsorted(values, trunc_inter, side="left"))      # count of values < trunc_inter
    hi_first = int(np.searchsorted(values, trunc_inter_high, side="right"))  # first value > trunc_inter_high
    equal = trunc_inter == trunc_inter_high
    if equal:
        hi_first = lo_cnt
    return lo_cnt, hi_first, int(equal)


def n_pairs(nres=None, xres=None, yres=None):
    if xres is None:
        return nres * (nres - 1) // 2
    return xres * yres


def condensed_index(i, j, nres):
    """p of the symmetric pair i<j (0-based)."""
    return i * nres - i * (i + 1) // 2 + (j - i - 1)


def pair_table(nres=None, xres=None, yres=None):
    """(pi, pj) int32 arrays: 0-based residue indices of every pair p, in condensed order."""
    if xres is None:
        pi, pj = np.triu_indices(nres, 1)
    else:
        pi, pj = np.divmod(np.arange(xres * yres), yres)
    return pi.astype(np.int32), pj.astype(np.int32)


def padded_stride(P):
    """Row pitch (elements) of a level stack: a multiple of 64 so every frame row is 128-byte aligned."""
    return (P + 63) // 64 * 64
