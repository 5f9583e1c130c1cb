"""The CPU oracle against the reference's own golden vectors."""
import os

import numpy as np
import pytest

from conftest import GOLDEN


def _load_xtc(orc):
    return orc.xtc_decode(open(os.path.join(GOLDEN, "sample.xtc"), "rb").read())


def test_xtc_header_and_geometry(orc):
    times, boxes, xyz = _load_xtc(orc)
    assert xyz.shape == (11, 399, 3)
    assert np.array_equal(times, np.arange(11) * 100000.0)
    assert np.allclose(boxes[0].diagonal(), 6.49083, atol=1e-5)
    bond = np.linalg.norm(xyz[0, 1] - xyz[0, 0])
    assert 0.09 < bond < 0.12     # C-H bond of the ACE cap


@pytest.mark.parametrize("pbc", [0, 1])
def test_stage1_reproduces_all_golden_levels(orc, goldens, peptide, pbc):
    """11 x 28 x 28 levels of test/1_ReadWrite/correct_output/matrices: zero mismatches."""
    _, sel, res_start = peptide
    assert len(sel) == 193 and len(res_start) == 29
    times, boxes, xyz = _load_xtc(orc)
    lv = orc.mdmat_levels(xyz[:, sel], boxes, np.arange(len(sel)), res_start, 1.0, 1001, pbc)
    assert np.array_equal(lv, goldens["levels_sym"])
    assert np.array_equal(lv[:, :14, 14:], goldens["levels_asym"])   # test/8_Asymmetric block


def test_stage1_all_atoms_group_differs(orc, goldens):
    """The goldens pin the Protein-H group: selecting every atom changes hundreds of entries."""
    from conan_b200 import topology
    top = topology.read_gro(os.path.join(GOLDEN, "peptide.gro"))
    sel = top.default_groups()[0][1]
    times, boxes, xyz = _load_xtc(orc)
    lv = orc.mdmat_levels(xyz[:1], boxes[:1], sel, top.residue_runs(sel), 1.0, 1001, 0)
    assert (lv[0] != goldens["levels_sym"][0]).sum() > 100


def _stats_vs_python(levels, lo_cnt, hi_first, equal):
    """Straight Python restatement of update_interaction / update_aggregate (conan.py:24-45, :599-611)."""
    nf, P = levels.shape
    res = dict(n_int=np.zeros(P, int), n_enc=np.zeros(P, int), first=np.full(P, -1), last=np.full(P, -1),
               status=np.zeros(P, int))
    for p in range(P):
        st = 0
        for f in range(nf):
            k = levels[f, p]
            if equal:
                ns = int(k < lo_cnt)
            elif k < lo_cnt:
                ns = 1
            elif k >= hi_first:
                ns = 0
            else:
                ns = st
            if ns:
                res["n_int"][p] += 1
                res["last"][p] = f
                if not st:
                    if res["first"][p] == -1:
                        res["first"][p] = f
                    res["n_enc"][p] += 1
            st = ns
        res["status"][p] = st
    return res


@pytest.mark.parametrize("equal", [0, 1])
def test_stage2_c_matches_python_loops(orc, equal):
    rng = np.random.default_rng(1)
    levels = rng.integers(0, 101, size=(37, 23)).astype(np.uint16)
    lo, hi = 50, (50 if equal else 61)
    o = orc.pair_stats(levels, lo, hi, equal)
    r = _stats_vs_python(levels, lo, hi, equal)
    assert np.array_equal(o["n_int"], r["n_int"]) and np.array_equal(o["n_enc"], r["n_enc"])
    assert np.array_equal(o["first_f"], r["first"]) and np.array_equal(o["last_f"], r["last"])
    assert np.array_equal(o["status"], r["status"])
    assert np.array_equal(o["sum_k"], levels.astype(np.int64).sum(0))
    assert np.array_equal(o["sum_k2"], (levels.astype(np.int64) ** 2).sum(0))
    assert np.array_equal(o["kmin"], levels.min(0)) and np.array_equal(o["kmax"], levels.max(0))
    d = levels.astype(np.int64)
    assert np.array_equal(o["ssd_first"], ((d - d[0]) ** 2).sum(1))
    assert np.array_equal(o["ssd_prev"][1:], ((d[1:] - d[:-1]) ** 2).sum(1)) and o["ssd_prev"][0] == 0
