"""`import conan`: every hot-path name of the reference is importable with the reference's signature and, run
on the host objects the reference's functions work on, reproduces what the UNMODIFIED reference wrote for the
same stacks (tests/golden/case_hyst.npz, case_asym.npz).  No GPU needed: these
helpers take frame dictionaries, aggregates rows, vectors and condensed distances from the caller."""
import inspect
import io
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

REFERENCE_SIGNATURES = {        # conan.py line -> argument names
    "read_next_time": ["inputf"],
    "read_frame": ["inputf", "nres", "nlevels", "nchar", "dict_legend", "nterm", "trunc", "asymm_data", "dimer"],
    "update_aggregate": ["r", "time", "trunc", "trunc_inter", "trunc_inter_high", "agg", "inter_mode"],
    "prepare_dimer_transpose": ["pairs_list", "pairs_legend"],
    "rmsd_frame": ["dict_a", "dict_b", "trunc", "xres", "yres", "asymm", "dimer"],
    "plot_frames": ["dict_frame", "first_frame", "prev_frame", "pairs_list", "dr_mode", "xterm", "xres", "yterm", "yres", "trunc",
                    "matrices", "clean_matrices", "n", "time", "rmsd_perframe_file", "asymm", "dimer"],
    "get_inter": ["residue_vectors", "nterm", "nres", "pearson_2d_density_file"],
    "plot_pc": ["xterm", "xres", "yterm", "yres", "ii", "vec", "pairs", "pairs_dict", "gnus_path", "domains"],
    "project_pc": ["v", "time_vec", "vectors", "pairs_list", "change_sign"],
    "read_pca": ["pca_dir", "xterm", "xres", "yterm", "yres", "time_vec", "asymm", "vectors_mean0", "pairs_list"],
    "get_best_center": ["cluster", "metric_cond", "n"],
    "call_hierarchical": ["dir", "vectors", "trunc", "trunc_inter", "condensed_dist", "n_elements", "ksi", "title", "time_vec", "type",
                          "xterm", "xres", "yterm", "yres", "asymm", "matrices"],
    "print_hierarchical_clusters": ["dir", "vectors", "trunc", "trunc_inter", "nframes", "clusters", "centers", "time_vec", "cost_list",
                                    "xterm", "xres", "yterm", "yres", "asymm", "type", "matrices"],
    "read_in_observables": ["inputf", "time_vec"],
    "read_in_observables_blind": ["inputf"],
    "read_index": ["inputf"],
    "read_matrix": ["inputf", "xterm", "xres", "yterm", "yres", "asymm", "trunc"],
    "write_matrix": ["outputf", "time", "xterm", "xres", "yterm", "yres", "trunc", "dict_frame", "dr", "asymm"],
    "update_interaction": ["r", "trunc_inter", "trunc_inter_high", "status", "inter_mode"],
    "weight_contacts": ["r", "trunc"],
    "cond_index": ["i", "j", "n"],
    "read_options": ["finput", "opts"],
    "adjust_truncs": ["trunc", "trunc_dr", "trunc_inter", "trunc_inter_high"],
    "read_index_folder": ["reread"],
    "parse_numbers": ["vals"],
    "read_stages": ["stagef"],
    "handle_backups": [],
}


@pytest.fixture(scope="module")
def conan():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    sys.modules.pop("conan", None)
    import conan as mod
    return mod


def test_reference_names_and_signatures(conan):
    for name, args in REFERENCE_SIGNATURES.items():
        assert hasattr(conan, name), name
        got = list(inspect.signature(getattr(conan, name)).parameters)
        assert got[:len(args)] == args, (name, got)
    # read_process_frames and mdmat_run keep the reference's leading arguments (extra keyword arguments follow)
    rpf = list(inspect.signature(conan.read_process_frames).parameters)
    assert rpf[:9] == ["inputf", "nres", "nlevels", "nchar", "dict_legend", "asymm_data", "trunc", "trunc_inter", "trunc_inter_high"]
    assert list(inspect.signature(conan.mdmat_run).parameters)[:10] == ["program", "traj", "coord", "nlevels", "trunc", "mean", "begin",
                                                                        "end", "dt", "indexf"]


def _frames_through_the_xpm_api(conan, c, workdir, asymm_data=None, nterm=1):
    """dmf.xpm of the fixture read frame by frame with read_index / read_next_time / read_frame."""
    from conan_b200 import xpm
    path = os.path.join(workdir, "dmf.xpm")
    with open(path, "w") as f:
        xpm.write_xpm(f, c["in/levels_full"], c["in/times"], float(c["in/trunc"]), int(c["in/nlevels"]))
    with open(path) as f:
        nres, nlevels, nchar, dict_legend, trunc = conan.read_index(f)
    frames, times = [], []
    with open(path) as f:
        t = conan.read_next_time(f)
        while t > float('-Inf'):
            frames.append(conan.read_frame(f, nres, nlevels, nchar, dict_legend, nterm, trunc, asymm_data, False))
            times.append(t)
            t = conan.read_next_time(f)
    return nres, trunc, frames, times


def test_frame_loop_reproduces_the_reference_tables(conan, tmp_path):
    """read_next_time / read_frame / update_aggregate / rmsd_frame driven the way read_process_frames drives them
    (conan.py:782-874) on the hysteresis case: timeline.dat, mdmat_average_rmsf.dat, time_mdmat_rmsd.dat."""
    c = np.load(os.path.join(GOLDEN, "case_hyst.npz"))
    nres, trunc, frames, times = _frames_through_the_xpm_api(conan, c, str(tmp_path))
    assert np.array_equal(times, c["in/times"]) and nres == c["in/levels_full"].shape[1]
    ti, th = float(c["ret/trunc_inter"]), float(c["ret/trunc_inter_high"])
    aggs = {}
    tn = len(times)
    rmsd = []
    conan.nres = nres                                   # the global rmsd_frame reads (conan.py:1432)
    for n, (t, fr) in enumerate(zip(times, frames)):
        for pair in set(aggs) | set(fr):
            if pair not in aggs:                        # a pair seen for the first time stood at trunc before (:856-863)
                aggs[pair] = np.array([-1, 0, 0, 0, 0, n * trunc, n * trunc * trunc])
            aggs[pair] = conan.update_aggregate(fr.get(pair, trunc), t, trunc, ti, th, aggs[pair])
        if n:
            rmsd.append((0.001 * t, conan.rmsd_frame(fr, frames[n - 1], trunc, nres, nres, False, False),
                         conan.rmsd_frame(fr, frames[0], trunc, nres, nres, False, False)))
    want_tl, want_avg = c["file/aggregate/timeline.dat"], c["file/aggregate/mdmat_average_rmsf.dat"]
    rows = {(int(r[0]), int(r[1])): r for r in want_tl}
    avgs = {(int(r[0]), int(r[1])): r for r in want_avg}
    for (i, j), agg in aggs.items():
        first, last, n_int, n_enc, status, sumr, sumr2 = agg
        w = rows[(i, j)]
        assert (w[2] == (first if first != -1 else times[-1])) and w[3] == last and w[6] == n_enc, (i, j)
        assert abs(w[4] - n_int / tn) < 1.0e-6
        assert abs(avgs[(i, j)][2] - sumr / tn) < 1.0001e-4
    assert np.abs(np.array(rmsd) - c["file/aggregate/time_mdmat_rmsd.dat"]).max() < 1.0001e-4
    # pairs_list of the reference = the pairs that ever came within trunc
    assert set(aggs) == {tuple(p) for p in c["ret/pairs_list"].tolist()}


def test_asymmetric_frames_and_dimer_orientation(conan, tmp_path):
    c = np.load(os.path.join(GOLDEN, "case_asym.npz"))
    xdict = {m: m for m in range(1, 6)}                 # START_X 1 NRES_X 5 NTERM_X 1
    ydict = {m: m for m in range(6, 13)}                # START_Y 6 NRES_Y 7 NTERM_Y 6
    nres, trunc, frames, times = _frames_through_the_xpm_api(conan, c, str(tmp_path), (xdict, ydict, 5, 7, 1, 6))
    seen = set().union(*[set(f) for f in frames])
    assert seen == {tuple(p) for p in c["ret/pairs_list"].tolist()}
    assert all(1 <= i <= 5 and 6 <= j <= 12 for i, j in seen)
    r = np.array([[0.001 * t, conan.rmsd_frame(fr, frames[n], trunc, 5, 7, True, False), conan.rmsd_frame(fr, frames[0], trunc, 5, 7, True, False)]
                  for n, (t, fr) in enumerate(zip(times[1:], frames[1:]))])
    assert np.abs(r - c["file/aggregate/time_mdmat_rmsd.dat"]).max() < 1.0001e-4
    # DIMER: a frame with more weighted contacts below the diagonal comes back transposed (conan.py:553-565)
    a = {(1, 2): 0.2, (3, 1): 0.9}
    assert conan.rmsd_frame(a, {(2, 1): 0.2, (1, 3): 0.9}, 1.0, 3, 3, True, True) == 0.0
    assert conan.rmsd_frame(a, {(2, 1): 0.2, (1, 3): 0.9}, 1.0, 3, 3, True, False) > 0.0
    npad, mapping = conan.prepare_dimer_transpose([(1, 2), (2, 1), (1, 3)], {(1, 2): 0, (2, 1): 1, (1, 3): 2})
    assert (npad, mapping) == (1, [1, 0, 3, 2])


def test_stage3_helpers_reproduce_the_reference_files(conan, tmp_path):
    """project_pc / plot_pc / read_pca / call_hierarchical / get_best_center / get_inter on the reference's own
    `vectors` of the hysteresis case against the files the reference wrote."""
    c = np.load(os.path.join(GOLDEN, "case_hyst.npz"))
    vectors = c["ret/vectors"]                                      # float16 [pairs, frames]
    pairs_list = [tuple(p) for p in c["ret/pairs_list"].tolist()]
    time_vec = list(c["ret/time_vec"])
    nres = c["in/levels_full"].shape[1]
    trunc = float(c["in/trunc"])
    legend = {}
    for row, p in enumerate(pairs_list):
        legend[p] = row
        legend[p[::-1]] = row
    conan.pairs_legend, conan.asymm, conan.nres, conan.gnus_path, conan.domains = legend, False, nres, "", 0
    old = os.getcwd()
    os.chdir(str(tmp_path))
    try:
        os.makedirs("pca")
        os.makedirs("cluster_trj")
        # PCA as conan.py:1973-2001 does it, with the helpers doing the projection and the maps
        mean0 = vectors - np.mean(vectors, axis=1)[:, None]
        w, v = np.linalg.eigh(np.cov(vectors))
        v = v[:, ::-1][:, :3]
        ordered = conan.project_pc(v, time_vec, mean0, pairs_list)
        for k in range(3):
            conan.plot_pc(1, nres, 1, nres, k, ordered[:, k], pairs_list, legend, "", 0)
        got = np.loadtxt("pca/pca_projections.dat")
        want = c["file/pca/pca_projections.dat"]
        assert got.shape == want.shape and np.abs(np.abs(got) - np.abs(want)).max() < 2e-3      # float16 mean removal (:1981)
        for k in (1, 2, 3):
            g, wv = np.loadtxt("pca/pc.%d.dat" % k), c["file/pca/pc.%d.dat" % k]
            assert np.array_equal(g[:, :2], wv[:, :2]) and np.abs(np.abs(g[:, 2]) - np.abs(wv[:, 2])).max() < 2e-3
        # read_pca projects on what was just written, without flipping signs
        os.rename("pca/pca_projections.dat", "pca/first.dat")
        conan.read_pca("pca", 1, nres, 1, nres, time_vec, False, mean0, pairs_list)
        assert np.abs(np.loadtxt("pca/pca_projections.dat") - np.loadtxt("pca/first.dat")).max() < 2e-5
        # trajectory clustering from the condensed frame-frame distances (conan.py:2034-2090)
        from scipy.spatial.distance import pdist
        cond = pdist(np.transpose(vectors) / (nres / np.sqrt(2.0)))
        conan.call_hierarchical("cluster_trj", vectors, trunc, float(c["ret/trunc_inter"]), cond, len(time_vec), [2, 3], "", time_vec,
                                "trj", 1, nres, 1, nres, False, False)
        assert np.abs(np.loadtxt("cluster_trj/linkage.txt") - c["file/cluster_trj/linkage.txt"]).max() < 2e-6
        for k in (2, 3):
            assert np.array_equal(np.loadtxt("cluster_trj/%d/assignment.dat" % k), c["file/cluster_trj/%d/assignment.dat" % k])
            for n in range(k):
                for name in ("average_map", "lifetime"):
                    key = "file/cluster_trj/%d/cluster%d/%s.dat" % (k, n, name)
                    assert np.abs(np.loadtxt("cluster_trj/%d/cluster%d/%s.dat" % (k, n, name)) - c[key]).max() < 1.0e-6, key
        ctr = conan.get_best_center([0, 1, 2, 3], cond, len(time_vec))
        sq = np.zeros((4, 4))
        assert 0 <= ctr <= 3
        # inter-residue correlation table
        rv = np.random.default_rng(1).normal(size=(4, 30))
        rv[2] = 1.0
        buf = io.StringIO()
        d = conan.get_inter(list(rv), 5, 4, buf)
        tab = np.loadtxt(io.StringIO(buf.getvalue()))
        assert tab.shape == (16, 3) and tab[0, 0] == 5
        assert abs(tab[1, 2] - np.corrcoef(rv[0], rv[1])[0, 1]) < 1e-6 and tab[10, 2] == 0.0 and d[(0, 2)] == 1.0
    finally:
        os.chdir(old)
