#!/usr/bin/env python3
"""Generates the committed fixtures in tests/golden/ from the reference tree.

Run HERE (where /root/reference exists):  python tests/golden/make_golden.py
Nothing under tests/ reads /root/reference at test time; the GPU box has no reference tree.

Fixtures written
  sample.xtc            copy of the reference's 11-frame test trajectory (input data, 17 KB)
  peptide.gro           topology of its 399 atoms, extracted from test/md.tpr, frame-0 coordinates
  ubi_12.xtc            first 12 frames of applications/1_ubi_unfold/pull_500ps.xtc (input data, 52 KB)
  ref_goldens.npz       the reference's own golden outputs (test/*/correct_output) parsed to arrays,
                        plus the 11 x 28 x 28 golden level matrices of test 1 and the 14 x 14 of test 8
  case_*.npz            inputs and outputs of the UNMODIFIED reference run here (oracle/ref_runner.py)
                        on synthetic level stacks that exercise hysteresis, observables, PCA,
                        frame clustering and the asymmetric mode
  ubi_pull_500ps.xtc    the reference's ubiquitin unfolding trajectory (applications/1_ubi_unfold, 472 frames
  ubi.pdb               of 1231 atoms, 2 MB) and its structure file: BASELINE.json configs[1]
  case_ubi.npz          outputs of the unmodified reference on that trajectory with the keywords of
                        applications/1_ubi_unfold/fine.inp plus a STAGES file; its dmf.xpm comes from the CPU
                        restatement of gmx mdmat (oracle/), since gmx itself is unavailable
  case_dimer.npz        DIMER yes on a small homodimer-like stack with a third of the frames transposed
  case_c5.npz           the same for a small two-chain interface system (50 x 100 residues, 200 frames) in the
                        asymmetric mode with K_TRAJ_CLUSTERS 2-5: BASELINE.json configs[4] scaled to what the
                        reference finishes in a minute
"""
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from conan_b200 import topology, xpm  # noqa: E402
from conan_b200.xtc import XtcFile  # noqa: E402
from oracle import ref_runner  # noqa: E402

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def parse_tree(root):
    """Every .dat/.txt under root -> {relative path: float array} (np.loadtxt semantics)."""
    res = {}
    for dp, _, files in os.walk(root):
        for fn in sorted(files):
            if not (fn.endswith(".dat") or fn.endswith(".txt")):
                continue
            p = os.path.join(dp, fn)
            try:
                arr = np.loadtxt(p)
            except Exception:
                continue
            res[os.path.relpath(p, root)] = arr
    return res


def reference_goldens():
    out = {}
    for test in ("1_ReadWrite", "2_Pearson", "3_ClusterTraj", "6_PCA", "8_Asymmetric"):
        tree = parse_tree(os.path.join(REF, "test", test, "correct_output"))
        for k, v in tree.items():
            if k.startswith("matrices/"):
                continue
            out["%s/%s" % (test, k)] = v
    lv = []
    for f in range(11):
        g = np.loadtxt(os.path.join(REF, "test/1_ReadWrite/correct_output/matrices/%05d.dat" % f))
        lv.append(np.rint(g[:, 2] * 1000).astype(np.uint16).reshape(28, 28))
    out["levels_sym"] = np.stack(lv)
    lv = []
    for f in range(11):
        g = np.loadtxt(os.path.join(REF, "test/8_Asymmetric/correct_output/matrices/%05d.dat" % f))
        lv.append(np.rint(g[:, 2] * 1000).astype(np.uint16).reshape(14, 14))
    out["levels_asym"] = np.stack(lv)
    out["observable"] = np.loadtxt(os.path.join(REF, "test/2_Pearson/observable.dat"), skiprows=2)
    np.savez_compressed(os.path.join(OUT, "ref_goldens.npz"), **out)
    print("ref_goldens.npz:", len(out), "arrays")


def peptide_fixture():
    shutil.copyfile(os.path.join(REF, "test/sample.xtc"), os.path.join(OUT, "sample.xtc"))
    top = topology.read_tpr(os.path.join(REF, "test/md.tpr"))
    x = XtcFile(os.path.join(OUT, "sample.xtc"))
    xyz0 = x.decode(frame_ids=[0])[0]
    box = x.boxes[0].diagonal()
    topology.write_gro(os.path.join(OUT, "peptide.gro"), top, xyz0, box, "ACE-(33..58)-NME peptide of the CONAN tests")
    print("peptide.gro:", top.natoms, "atoms")
    # the first 12 frames of the reference's ubiquitin unfolding trajectory (1231 atoms, written by gmx):
    # a second real compressed stream for the decoders
    data = open(os.path.join(REF, "applications/1_ubi_unfold/pull_500ps.xtc"), "rb").read()
    xu = XtcFile(data)
    with open(os.path.join(OUT, "ubi_12.xtc"), "wb") as f:
        f.write(data[:int(xu.offsets[12])])
    print("ubi_12.xtc:", int(xu.offsets[12]), "bytes,", xu.natoms, "atoms")


def synthetic_levels(seed, nf, nres, nlevels, lo, hi):
    """Random-walk distances that cross the interaction cut-offs often; some pairs never in range."""
    rng = np.random.default_rng(seed)
    iu = np.triu_indices(nres, 1)
    P = len(iu[0])
    base = rng.uniform(0.15, 1.3, size=P)
    walk = np.cumsum(rng.normal(0, 0.06, size=(nf, P)), axis=0)
    r = np.clip(base + walk - walk.mean(axis=0), 0.0, 1.0)
    r[:, rng.random(P) < 0.15] = 1.0           # pairs never within the truncation distance
    far = rng.random(P) < 0.2                   # pairs that appear late
    r[: nf // 3, far] = 1.0
    lv = np.rint(r * (nlevels - 1)).astype(np.uint16)
    full = np.zeros((nf, nres, nres), np.uint16)
    full[:, iu[0], iu[1]] = lv
    full[:, iu[1], iu[0]] = lv
    return full


def run_case(name, full, times, trunc, nlevels, keywords, obs=None, extra_files=None, store_levels=True, keep=None):
    work = tempfile.mkdtemp(prefix="conan_golden_")
    with open(os.path.join(work, "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, full, times, trunc, nlevels)
    text = "RUN_MDMAT no\nTRUNC %g\nNLEVEL %d\nMAKE_MOVIE no\n" % (trunc, nlevels) + keywords
    if obs is not None:
        with open(os.path.join(work, "obs.dat"), "w") as f:
            f.write("%d\n" % obs.shape[1])
            for k in range(obs.shape[1]):
                f.write("observable %d\n" % k)
            for t, row in zip(times, obs):
                f.write("%d %s\n" % (t, " ".join("%.6f" % v for v in row)))
        text += "PEARSON_OBS obs.dat\n"
    for fname, content in (extra_files or {}).items():
        os.makedirs(os.path.dirname(os.path.join(work, fname)), exist_ok=True)
        with open(os.path.join(work, fname), "w") as f:
            f.write(content)
    g = ref_runner.run_reference(work, text)
    out = {"in/times": np.asarray(times, np.float64), "in/trunc": np.float64(trunc),
           "in/nlevels": np.int64(nlevels), "in/keywords": np.array(text)}
    if store_levels:
        out["in/levels_full"] = full
    else:   # large stacks are recomputed by the test from the trajectory; a digest and three frames pin them
        import hashlib
        out["in/levels_sha256"] = np.array(hashlib.sha256(np.ascontiguousarray(full).tobytes()).hexdigest())
        out["in/levels_sample"] = full[[0, len(full) // 2, len(full) - 1]]
    if obs is not None:
        out["in/obs"] = obs
    for k, v in parse_tree(work).items():
        if k.startswith("input/") or k.startswith("oldpca/") or k == "ref.inp" or k == "obs.dat":
            continue
        if keep is not None and not keep(k):
            continue
        out["file/" + k] = v
    out["ret/vectors"] = np.asarray(g["vectors"])
    out["ret/pairs_list"] = np.asarray(g["pairs_list"], dtype=np.int64)
    out["ret/time_vec"] = np.asarray(g["time_vec"], dtype=np.float64)
    out["ret/trunc_inter"] = np.float64(g["trunc_inter"])
    out["ret/trunc_inter_high"] = np.float64(g["trunc_inter_high"])
    np.savez_compressed(os.path.join(OUT, "case_%s.npz" % name), **out)
    shutil.rmtree(work)
    print("case_%s.npz:" % name, len(out), "arrays;", "pairs", len(g["pairs_list"]))


UBI_STAGES = '#Three stages of the pulling run\n0 60000 "folded"\n60000 150000 "unfolding"\n150000 235500 "extended"\n'


def ubi_case():
    """BASELINE.json configs[1]: applications/1_ubi_unfold with the keywords of fine.inp (:3-8, :13) minus
    GNUS_PATH / COORD_PDB / DOMAINS, plus STAGES (conan.py:798-805, :961-990)."""
    from oracle import oracle
    app = os.path.join(REF, "applications/1_ubi_unfold")
    shutil.copyfile(os.path.join(app, "pull_500ps.xtc"), os.path.join(OUT, "ubi_pull_500ps.xtc"))
    shutil.copyfile(os.path.join(app, "ubi.pdb"), os.path.join(OUT, "ubi.pdb"))
    top = topology.read_topology(os.path.join(OUT, "ubi.pdb"))
    x = XtcFile(os.path.join(OUT, "ubi_pull_500ps.xtc"))
    sel = top.default_groups()[2][1]                       # Protein-H, what the reference's tests pick
    res_start = top.residue_runs(sel)
    xyz = x.decode()
    full = oracle.mdmat_levels(xyz, x.boxes, sel, res_start, 1.0, 1001, 1)
    print("ubiquitin: %d frames, %d atoms, %d selected, %d residues" % (len(xyz), xyz.shape[1], len(sel), len(res_start) - 1))
    sample = {"00000", "00235", "00471"}
    run_case("ubi", full, x.times, 1.0, 1001,
             "DT 500\nTRUNC_INTER 0.5\nTRUNC_DR 1.0\nDR_MODE init\nK_TRAJ_CLUSTERS 1-5\nSTAGES stages.txt\n",
             extra_files={"stages.txt": UBI_STAGES}, store_levels=False,
             keep=lambda k: not k.startswith("matrices/") or k.split("/")[1][:5] in sample or k.split("/")[1].startswith("stage_"))


def c5_case():
    """BASELINE.json configs[4] at a size the reference finishes in a minute: two chains of 50 and 100
    residues, 200 frames, START_X / NRES_X / START_Y / NRES_Y as SURVEY section 8(d) writes them."""
    import hashlib
    from conan_b200 import synth
    from oracle import oracle
    nx, ny, nf = 50, 100, 200
    xyz, res_start, _ = synth.make_trajectory(nx + ny, nf, 8, seed=5, two_chain=(nx, ny), closer=0.7)
    full = oracle.mdmat_levels(xyz, None, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0)
    run_case("c5", full, np.arange(nf) * 100, 1.0, 1001,
             "TRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nSTART_X 1\nNRES_X %d\nNTERM_X 1\nSTART_Y %d\nNRES_Y %d\nNTERM_Y %d\n"
             "MATRICES no\nK_TRAJ_CLUSTERS 2-5\n" % (nx, nx + 1, ny, nx + 1), store_levels=False)
    c = dict(np.load(os.path.join(OUT, "case_c5.npz")))
    c["in/xyz_sha256"] = np.array(hashlib.sha256(np.ascontiguousarray(xyz).tobytes()).hexdigest())
    c["in/synth"] = np.array([nx, ny, nf, 8, 5, 0.7])
    np.savez_compressed(os.path.join(OUT, "case_c5.npz"), **c)


def dimer_case():
    """DIMER yes on a 2 x 6 residue system whose X x Y block is handed to the reference transposed in a third
    of the frames (the two chains exchanged): read_frame's orientation (conan.py:553-565) and the dimer branch of
    rmsd_frame (:1420-1428)."""
    nf, n, nl = 40, 6, 101
    full = synthetic_levels(21, nf, 2 * n, nl, 0.5, 0.6)
    rng = np.random.default_rng(22)
    swap = rng.random(nf) < 0.35
    for f in np.flatnonzero(swap):
        blk = full[f, :n, n:].copy()
        full[f, :n, n:] = blk.T
        full[f, n:, :n] = blk
    run_case("dimer", full, np.arange(nf) * 100, 1.0, nl,
             "TRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nSTART_X 1\nNRES_X %d\nNTERM_X 1\nSTART_Y %d\nNRES_Y %d\nNTERM_Y 1\n"
             "DIMER yes\nDR_MODE both\n" % (n, n + 1, n))
    c = dict(np.load(os.path.join(OUT, "case_dimer.npz")))
    c["in/swapped_frames"] = swap
    np.savez_compressed(os.path.join(OUT, "case_dimer.npz"), **c)


def main():
    if not ref_runner.available():
        sys.exit("needs /root/reference")
    only = sys.argv[1] if len(sys.argv) > 1 else None
    if only == "dimer":
        return dimer_case()
    if only == "ubi":
        return ubi_case()
    if only == "c5":
        return c5_case()
    if not only:
        peptide_fixture()
        reference_goldens()
    nf, nres, nl = 48, 12, 101
    times = np.arange(nf) * 100
    full = synthetic_levels(11, nf, nres, nl, 0.5, 0.6)
    rng = np.random.default_rng(5)
    obs = np.stack([np.linspace(0, 1, nf) + rng.normal(0, 0.2, nf), rng.normal(0, 1, nf)], axis=1)
    if not only:
        run_case("hyst", full, times, 1.0, nl,
                 "TRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nPEARSON_TIME yes\nPEARSON_INTER yes\nPCA_MAKE 3\n"
                 "K_TRAJ_CLUSTERS 2-3\nDR_MODE both\n", obs)
        run_case("equalcut", full, times, 1.0, nl, "TRUNC_INTER 0.45\nMATRICES no\nPEARSON_TIME yes\n")
        run_case("default", full[:20], times[:20], 1.0, nl, "MATRICES no\n")
        run_case("asym", full, times, 1.0, nl,
                 "TRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nSTART_X 1\nNRES_X 5\nNTERM_X 1\nSTART_Y 6\nNRES_Y 7\nNTERM_Y 6\n"
                 "MATRICES no\nPEARSON_TIME yes\nK_TRAJ_CLUSTERS 2\n")
    if only in (None, "zoom"):
        run_case("zoom", full[:30], times[:30], 1.0, nl, "TRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nMATRICES no\nZOOM_LIST zoom.txt\n",
                 extra_files={"zoom.txt": "# pairs\n2 7\n9 3\n1 12\n"})
    if only in (None, "pcaread"):
        # PCA_READ: principal components written by the reference's own PCA_MAKE run (case_hyst), read back and projected
        hyst = np.load(os.path.join(OUT, "case_hyst.npz"))
        pcs = {}
        for k in (1, 2, 3):
            rows = hyst["file/pca/pc.%d.dat" % k]
            text, last = [], None
            for i, j, v in rows:
                if last is not None and i != last:
                    text.append("\n")
                text.append("%5i %5i %12.6f\n" % (i, j, v))
                last = i
            pcs["oldpca/pc.%d.dat" % k] = "".join(text) + "\n"
        run_case("pcaread", full, times, 1.0, nl,
                 "TRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nMATRICES no\nPCA_READ oldpca\n", extra_files=pcs)
    if only:
        return
    ubi_case()
    c5_case()
    dimer_case()
    # finer legend: 1001 levels, trunc 1.5 (dx = 0.0015, not a multiple of 1e-4)
    full2 = synthetic_levels(12, 30, 9, 1001, 0.5, 0.6)
    run_case("fine", full2, np.arange(30) * 500 + 1000, 1.5, 1001,
             "TRUNC_INTER 0.6\nTRUNC_INTER_HIGH 0.8\nMATRICES no\nPEARSON_TIME yes\nPCA_MAKE 2\n")


if __name__ == "__main__":
    main()
