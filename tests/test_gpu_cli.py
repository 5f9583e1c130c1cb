"""The drop-in surface end to end on the GPU: `conan.py input.inp` with the reference's own test inputs.

Inputs are the reference's test trajectory (tests/golden/sample.xtc) and the topology extracted
from its md.tpr; outputs are compared with the reference's golden files (test/1_ReadWrite and
test/8_Asymmetric correct_output, parsed into tests/golden/ref_goldens.npz) with the reference's
own comparator tolerance (test/compare_all.py: atol = rtol = 1.1e-3) and, for the integer columns,
exactly.
"""
import io
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu


def _run_cli(workdir, text, stdin="2\n"):
    with open(os.path.join(workdir, "test.inp"), "w") as f:
        f.write(text)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "conan.py"), "test.inp"], cwd=workdir, input=stdin,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def _close(a, b):
    return a.shape == b.shape and np.isclose(a, b, atol=1.1e-3, rtol=1.1e-3).all()


def test_readwrite_matches_reference_goldens(lib, goldens, tmp_path):
    out = _run_cli(str(tmp_path), "TRAJ %s/sample.xtc\nCOORD %s/peptide.gro\nTRUNC 1.0\nNLEVEL 1001\n" % (GOLDEN, GOLDEN))
    assert "Protein-H" in out
    for name in ("timeline", "mdmat_average_rmsf", "native_contacts", "time_mdmat_rmsd", "local_interactions"):
        got = np.loadtxt(str(tmp_path / "aggregate" / (name + ".dat")))
        want = goldens["1_ReadWrite/aggregate/%s.dat" % name]
        assert _close(got, want), name
    tl = np.loadtxt(str(tmp_path / "aggregate" / "timeline.dat"))
    want = goldens["1_ReadWrite/aggregate/timeline.dat"]
    assert np.array_equal(tl[:, [0, 1, 2, 3, 6]], want[:, [0, 1, 2, 3, 6]])     # i, j, first, last, encounters
    # matrices/*.dat are the golden levels
    for f in (0, 5, 10):
        m = np.loadtxt(str(tmp_path / "matrices" / ("%05d.dat" % f)))
        assert np.array_equal(np.rint(m[:, 2] * 1000).astype(int).reshape(28, 28), goldens["levels_sym"][f])
    # the dmf.xpm written next to them parses back to the same stack
    from conan_b200 import xpm
    times, frames, trunc, nlevels = xpm.read_xpm(str(tmp_path / "dmf.xpm"))
    assert (trunc, nlevels) == (1.0, 1001) and np.array_equal(frames, goldens["levels_sym"])


def test_asymmetric_from_xpm_matches_reference_goldens(lib, goldens, tmp_path):
    """RUN_MDMAT no: the stack is read from a dmf.xpm, as in the reference's tests 2-8."""
    from conan_b200 import xpm
    with open(str(tmp_path / "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, goldens["levels_sym"], np.arange(11) * 100000.0, 1.0, 1001)
    _run_cli(str(tmp_path), "RUN_MDMAT no\nNRES_X 14\nNRES_Y 14\nNTERM_X 1\nNTERM_Y 15\nSTART_X 1\nSTART_Y 15\n")
    for name in ("timeline", "mdmat_average_rmsf", "native_contacts", "time_mdmat_rmsd", "total_interactions.x",
                 "total_interactions.y"):
        got = np.loadtxt(str(tmp_path / "aggregate" / (name + ".dat")))
        assert _close(got, goldens["8_Asymmetric/aggregate/%s.dat" % name]), name
    m = np.loadtxt(str(tmp_path / "matrices" / "00003.dat"))
    assert np.array_equal(np.rint(m[:, 2] * 1000).astype(int).reshape(14, 14), goldens["levels_asym"][3])


@pytest.mark.parametrize("case", ["hyst", "equalcut", "default", "asym", "fine", "dimer"])
def test_cases_generated_by_the_live_reference(lib, tmp_path, case):
    """Synthetic stacks with hysteresis etc.; expected files were produced by the unmodified reference."""
    from conan_b200 import xpm
    c = np.load(os.path.join(GOLDEN, "case_%s.npz" % case))
    with open(str(tmp_path / "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, c["in/levels_full"], c["in/times"], float(c["in/trunc"]), int(c["in/nlevels"]))
    keywords = str(c["in/keywords"]).replace("PEARSON_OBS obs.dat\n", "")
    _run_cli(str(tmp_path), keywords)
    checked = 0
    for key in c.files:
        if not key.startswith("file/aggregate/") or "pearson" in key:
            continue
        path = tmp_path / key[len("file/"):]
        assert path.exists(), key
        got, want = np.loadtxt(str(path)), c[key]
        tol = 1.0001e-4 if ("rmsf" in key or "rmsd" in key or "native" in key) else 1.0001e-6   # last printed digit
        assert got.shape == want.shape and np.abs(got - want).max() <= tol, key
        checked += 1
    assert checked >= 5
    for key in c.files:
        if key.startswith("file/matrices/"):
            got = np.loadtxt(str(tmp_path / key[len("file/"):]))
            assert np.allclose(got, c[key], atol=1.01e-4, rtol=0), key


def _xpm_from_goldens(goldens, workdir):
    from conan_b200 import xpm
    with open(os.path.join(workdir, "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, goldens["levels_sym"], np.arange(11) * 100000.0, 1.0, 1001)


def test_pearson_matches_reference_goldens(lib, goldens, tmp_path):
    """The reference's test 2: correlation with time, between residues and with an observable."""
    _xpm_from_goldens(goldens, str(tmp_path))
    with open(str(tmp_path / "observable.dat"), "w") as f:
        f.write("1\nRandom variable\n")
        for t, v in goldens["observable"]:
            f.write("%d\t%.3f\n" % (t, v))
    _run_cli(str(tmp_path), "RUN_MDMAT no\nPEARSON_TIME yes\nPEARSON_INTER yes\nPEARSON_OBS observable.dat\nMATRICES no\nMAKE_MOVIE no\n")
    for name in ("pearson_data", "pearson_data_obs_1", "pearson_map_density"):
        got = np.loadtxt(str(tmp_path / "aggregate" / (name + ".dat")))
        assert _close(got, goldens["2_Pearson/aggregate/%s.dat" % name]), name


def test_pca_matches_reference_goldens(lib, goldens, tmp_path):
    """The reference's test 6: ten principal components of the contact series."""
    _xpm_from_goldens(goldens, str(tmp_path))
    _run_cli(str(tmp_path), "RUN_MDMAT no\nPCA_MAKE 10\nMATRICES no\n")
    names = ["pca_output.txt", "pca_projections.dat"] + ["pc.%d.dat" % i for i in range(1, 10)]
    for name in names:
        got = np.loadtxt(str(tmp_path / "pca" / name))
        assert _close(got, goldens["6_PCA/pca/%s" % name]), name


def test_cluster_trajectory_matches_reference_goldens(lib, goldens, tmp_path):
    """The reference's test 3: frame-frame distances, Ward clustering into 1-3 clusters, medoids and maps."""
    _xpm_from_goldens(goldens, str(tmp_path))
    _run_cli(str(tmp_path), "RUN_MDMAT no\nK_TRAJ_CLUSTERS 1-3\nMATRICES no\n")
    checked = 0
    for key in goldens.files:
        if key.startswith("3_ClusterTraj/"):
            got = np.loadtxt(str(tmp_path / key[len("3_ClusterTraj/"):]))
            assert _close(got, goldens[key]), key
            checked += 1
    assert checked == 16


def test_stage3_files_of_the_live_reference_case(lib, tmp_path):
    """Hysteresis case with two observables, PCA and clustering; expected files from the unmodified reference."""
    from conan_b200 import xpm
    c = np.load(os.path.join(GOLDEN, "case_hyst.npz"))
    with open(str(tmp_path / "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, c["in/levels_full"], c["in/times"], float(c["in/trunc"]), int(c["in/nlevels"]))
    obs = c["in/obs"]
    with open(str(tmp_path / "obs.dat"), "w") as f:
        f.write("%d\n" % obs.shape[1])
        for k in range(obs.shape[1]):
            f.write("observable %d\n" % k)
        for t, row in zip(c["in/times"], obs):
            f.write("%d %s\n" % (t, " ".join("%.6f" % v for v in row)))
    _run_cli(str(tmp_path), str(c["in/keywords"]))
    checked = 0
    for key in c.files:
        if not key.startswith("file/") or key.startswith(("file/matrices", "file/aggregate/timeline", "file/aggregate/mdmat",
                                                          "file/aggregate/native", "file/aggregate/time_", "file/aggregate/local")):
            continue
        path = tmp_path / key[len("file/"):]
        assert path.exists(), key
        got, want = np.loadtxt(str(path)), c[key]
        assert got.shape == want.shape, key
        if "pc." in key or "pca_projections" in key:
            tol = 2e-5 * max(1.0, np.abs(want).max())      # eigenvectors of a float32 covariance
        elif "linkage" in key or "rmsd_interframe" in key or "assignment" in key:
            tol = 2e-5
        else:
            tol = 2.0e-6
        assert np.abs(got - want).max() <= tol, (key, np.abs(got - want).max())
        checked += 1
    assert checked >= 20


def test_reread_matches_reference_goldens(lib, goldens, tmp_path):
    """The reference's test 7: matrices of an earlier run are read back, differential maps are written."""
    from conan_b200 import levels as lv
    os.makedirs(str(tmp_path / "to_read"))
    values = lv.legend_values(1.0, 1001)
    for f in range(11):          # the reference's to_read/ folder is its own test-1 output
        with open(str(tmp_path / "to_read" / ("%05d.dat" % f)), "w") as out:
            out.write("# i  j   r_ij\n")
            m = values[goldens["levels_sym"][f]]
            for i in range(28):
                out.write("".join("%4d %4d %9.4f\n" % (i + 1, j + 1, m[i, j]) for j in range(28)) + "\n")
    with open(str(tmp_path / "to_read" / "index.dat"), "w") as out:
        out.write("NTERM 1\nNRES 28\nDT 100000\nTRUNC 1.0\n")
    _run_cli(str(tmp_path), "REREAD to_read/\nDR_MODE both\n")
    lv_sym = goldens["levels_sym"].astype(np.int64)
    for f in (0, 3, 10):
        m = np.loadtxt(str(tmp_path / "matrices" / ("%05d.dat" % f)))
        assert np.array_equal(np.rint(m[:, 2] * 1000).astype(int).reshape(28, 28), lv_sym[f])
    for f in (1, 7):
        di = np.loadtxt(str(tmp_path / "matrices" / ("%05d_dr.init.dat" % f)))
        dp = np.loadtxt(str(tmp_path / "matrices" / ("%05d_dr.prev.dat" % f)))
        assert np.array_equal(np.rint(di[:, 2] * 1000).astype(int).reshape(28, 28), lv_sym[f] - lv_sym[0])
        assert np.array_equal(np.rint(dp[:, 2] * 1000).astype(int).reshape(28, 28), lv_sym[f] - lv_sym[f - 1])
    tl = np.loadtxt(str(tmp_path / "aggregate" / "timeline.dat"))
    want = goldens["1_ReadWrite/aggregate/timeline.dat"]
    assert np.array_equal(tl[:, [0, 1, 4, 6]], want[:, [0, 1, 4, 6]])        # lifetimes and encounters as in test 1


def test_zoom_matches_live_reference(lib, tmp_path):
    """ZOOM_LIST: distance trace with hysteresis state and the pair's correlation with every other pair."""
    from conan_b200 import xpm
    c = np.load(os.path.join(GOLDEN, "case_zoom.npz"))
    with open(str(tmp_path / "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, c["in/levels_full"], c["in/times"], float(c["in/trunc"]), int(c["in/nlevels"]))
    with open(str(tmp_path / "zoom.txt"), "w") as f:
        f.write("# pairs\n2 7\n9 3\n1 12\n")
    _run_cli(str(tmp_path), str(c["in/keywords"]))
    n = 0
    for key in c.files:
        if key.startswith("file/zoom/"):
            got = np.loadtxt(str(tmp_path / key[len("file/"):]))
            # scipy's pearsonr keeps the float16 dtype of `vectors`, so the reference's own numbers carry
            # float16 rounding (~5e-4); the distance traces are exact
            tol = 3.0e-3 if "2d_correlate" in key else 2.0e-6
            assert got.shape == c[key].shape and np.abs(got - c[key]).max() <= tol, key
            n += 1
    assert n == 6
    # ... and against float64 Pearson coefficients of the reference's own `vectors`, to print precision
    vec = c["ret/vectors"].astype(np.float64)
    pairs = [tuple(int(v) for v in row) for row in c["ret/pairs_list"]]
    zp = [q for q in ((2, 7), (3, 9), (1, 12)) if q in pairs][0]
    zi = pairs.index(zp)
    got = np.loadtxt(str(tmp_path / "zoom" / ("2d_correlate_%d_%d.dat" % zp)))
    table = {(int(a), int(b)): v for a, b, v in got}
    for row, pair in enumerate(pairs):
        want = np.corrcoef(vec[zi], vec[row])[0, 1] if vec[row].std() > 0 else 0.0
        assert abs(table[pair] - want) < 2e-6 and abs(table[pair[::-1]] - want) < 2e-6, pair



def test_asymmetric_from_trajectory_computes_only_the_block(lib, goldens, tmp_path):
    """START_X / START_Y with a trajectory: stage 1 runs in block mode and the reference's test-8 goldens still match."""
    out = _run_cli(str(tmp_path), "TRAJ %s/sample.xtc\nCOORD %s/peptide.gro\nTRUNC 1.0\nNLEVEL 1001\nNRES_X 14\nNRES_Y 14\n"
                   "NTERM_X 1\nNTERM_Y 15\nSTART_X 1\nSTART_Y 15\n" % (GOLDEN, GOLDEN))
    assert "only the 14 x 14 block" in out and not (tmp_path / "dmf.xpm").exists()
    for name in ("timeline", "mdmat_average_rmsf", "native_contacts", "time_mdmat_rmsd", "total_interactions.x",
                 "total_interactions.y"):
        got = np.loadtxt(str(tmp_path / "aggregate" / (name + ".dat")))
        assert _close(got, goldens["8_Asymmetric/aggregate/%s.dat" % name]), name
    m = np.loadtxt(str(tmp_path / "matrices" / "00007.dat"))
    assert np.array_equal(np.rint(m[:, 2] * 1000).astype(int).reshape(14, 14), goldens["levels_asym"][7])


def test_begin_end_dt_select_frames(lib, orc, goldens, tmp_path):
    """BEGIN / END / DT: gmx's -b/-e/-dt filter and CONAN's own (time-begin) % dt test pick the same frames."""
    from conan_b200 import levels as lv
    _run_cli(str(tmp_path), "TRAJ %s/sample.xtc\nCOORD %s/peptide.gro\nTRUNC 1.0\nNLEVEL 1001\nBEGIN 200000\nEND 900000\n"
             "DT 200000\nTRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\n" % (GOLDEN, GOLDEN))
    frames = [2, 4, 6, 8]
    assert sorted(os.listdir(str(tmp_path / "matrices"))) == ["%05d.dat" % n for n in range(4)]
    for n, f in enumerate(frames):
        m = np.loadtxt(str(tmp_path / "matrices" / ("%05d.dat" % n)))
        assert np.array_equal(np.rint(m[:, 2] * 1000).astype(int).reshape(28, 28), goldens["levels_sym"][f])
    values = lv.legend_values(1.0, 1001)
    lo, hi, eq = lv.level_thresholds(values, 0.5, 0.6)
    want = orc.pair_stats(orc.condensed(goldens["levels_sym"][frames]), lo, hi, eq)
    nc = np.loadtxt(str(tmp_path / "aggregate" / "native_contacts.dat"))
    assert np.array_equal(nc[:, 0], np.array([200.0, 400.0, 600.0, 800.0]))
    assert np.array_equal(nc[:, 1], want["native"]) and np.array_equal(nc[:, 2], want["nonnative"])


@pytest.mark.parametrize("case", ["hyst", "asym"])
def test_import_api_return_tuple(lib, tmp_path, case):
    """`import conan` users: read_index + read_process_frames with the reference's signature and return
    tuple (conan.py:697, :1021-1023), `vectors` included (it is gathered from the GPU stack on first use)."""
    from conan_b200 import frontend, xpm
    c = np.load(os.path.join(GOLDEN, "case_%s.npz" % case))
    full, times = c["in/levels_full"], c["in/times"]
    trunc, nlevels = float(c["in/trunc"]), int(c["in/nlevels"])
    cwd = os.getcwd()
    os.chdir(str(tmp_path))
    try:
        with open("dmf.xpm", "w") as f:
            xpm.write_xpm(f, full, times, trunc, nlevels)
        nres = full.shape[1]
        asymm_data = None
        if case == "asym":
            xdict = {m: m for m in range(1, 6)}
            ydict = {m: m for m in range(6, 13)}
            asymm_data = (xdict, ydict, 5, 7, 1, 6)
        with open("dmf.xpm") as inputf:
            nres_r, nlevels_r, nchar, dict_legend, trunc_r = frontend.read_index(inputf)
            assert (nres_r, nlevels_r) == (nres, nlevels) and abs(trunc_r - trunc) < 1e-6
            ret = frontend.read_process_frames(inputf, nres, nlevels, nchar, dict_legend, asymm_data, trunc, 0.5, 0.6,
                                               None, None, None, False, None, 0, None, False, False, False, None,
                                               [], None, 0.5, matrices=False)
    finally:
        os.chdir(cwd)
    begin, dt, time_vec, pairs_list, pairs_legend, vectors = ret[:6]
    assert np.array_equal(time_vec, c["ret/time_vec"])
    want_pairs = [tuple(p) for p in c["ret/pairs_list"]]
    assert sorted(pairs_list) == sorted(want_pairs)
    assert vectors.shape == c["ret/vectors"].shape and vectors.dtype == np.float16
    got = np.asarray(vectors)
    order = [pairs_legend[p] for p in want_pairs]        # the reference lists pairs in order of first appearance
    assert np.array_equal(got[order], c["ret/vectors"])
    assert np.array_equal(vectors[order[0]], c["ret/vectors"][0]) and len(vectors) == len(want_pairs)


def test_pca_read_projects_on_components_of_an_earlier_run(lib, tmp_path):
    """PCA_READ (conan.py:1073-1095): the components come from the reference's own PCA_MAKE run on the same stack
    (case_hyst), the expected projections from the reference's PCA_READ run (case_pcaread)."""
    from conan_b200 import xpm
    c = np.load(os.path.join(GOLDEN, "case_pcaread.npz"))
    hyst = np.load(os.path.join(GOLDEN, "case_hyst.npz"))
    os.makedirs(str(tmp_path / "oldpca"))
    for k in (1, 2, 3):
        text, last = [], None
        for i, j, v in hyst["file/pca/pc.%d.dat" % k]:
            if last is not None and i != last:
                text.append("\n")
            text.append("%5i %5i %12.6f\n" % (i, j, v))
            last = i
        (tmp_path / "oldpca" / ("pc.%d.dat" % k)).write_text("".join(text) + "\n")
    with open(str(tmp_path / "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, c["in/levels_full"], c["in/times"], float(c["in/trunc"]), int(c["in/nlevels"]))
    out = _run_cli(str(tmp_path), str(c["in/keywords"]))
    assert "I found 3 principal components" in out and "Norm of PC # 1" in out
    got = np.loadtxt(str(tmp_path / "pca" / "pca_projections.dat"))
    want = c["file/pca/pca_projections.dat"]
    assert got.shape == want.shape and np.array_equal(got[:, 0], want[:, 0])
    assert np.abs(got[:, 1:] - want[:, 1:]).max() <= 2.0e-6
