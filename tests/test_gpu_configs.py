"""BASELINE.json configs[1] (ubiquitin unfolding, applications/1_ubi_unfold) and configs[4] (two-chain interface
with frame-frame clustering) on the GPU.

configs[1]: the reference's own trajectory (tests/golden/ubi_pull_500ps.xtc, 472 frames x 1231 atoms, box
27.9 x 6.97 x 6.97 nm) and structure file.  Stage 1 is compared bit by bit with the CPU restatement of gmx mdmat,
with and without the periodic minimum image; the command-line run is compared with what the UNMODIFIED
reference wrote for the same keywords (fine.inp plus a STAGES file), tests/golden/case_ubi.npz.

configs[4]: a small two-chain system against the live reference (tests/golden/case_c5.npz), and the full size
(200 x 400 residues, 20 000 frames) through properties that hold for any input.
"""
import hashlib
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
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def _tolerance(key, want):
    """Printed precision of each file: %9.4f columns differ by at most one unit in the last place."""
    if "assignment" in key or "linkage" in key or "rmsd_interframe" in key:
        return 2e-5
    if "rmsf" in key or "rmsd" in key or "native" in key or "matrices/" in key or "average_map" in key or "lifetime" in key:
        return 1.0001e-4
    return 2.0e-6


def _compare_tree(c, workdir, skip=()):
    checked = 0
    for key in c.files:
        if not key.startswith("file/") or any(s in key for s in skip):
            continue
        path = os.path.join(workdir, key[len("file/"):])
        assert os.path.exists(path), key
        got, want = np.loadtxt(path), c[key]
        assert got.shape == want.shape, key
        if key.endswith("linkage.txt"):
            # merge heights only: frames with identical contact maps are exactly tied in the reference's float64
            # distances, and which of two tied merges comes first is not defined by the data
            got, want = got[:, 2], want[:, 2]
        err = np.abs(got - want).max()
        assert err <= _tolerance(key, want), (key, err)
        checked += 1
    return checked


# ------------------------------------------------------------------------------------------------
# configs[1]: ubiquitin
# ------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ubiquitin():
    from conan_b200 import topology
    from conan_b200.xtc import XtcFile
    top = topology.read_topology(os.path.join(GOLDEN, "ubi.pdb"))
    x = XtcFile(os.path.join(GOLDEN, "ubi_pull_500ps.xtc"))
    sel = top.default_groups()[2][1]
    return x, x.decode(), sel, top.residue_runs(sel)


def test_ubiquitin_stage1_bit_exact_with_and_without_pbc(lib, orc, ubiquitin):
    import torch
    from conan_b200 import stage1
    x, xyz, sel, res_start = ubiquitin
    assert xyz.shape == (472, 1231, 3) and len(sel) == 602 and len(res_start) == 77
    c = np.load(os.path.join(GOLDEN, "case_ubi.npz"))
    geom = stage1.Geometry(res_start)
    assert geom.max_res_atoms > 8                       # real residues: ragged sizes, up to 14 heavy atoms
    picked = np.ascontiguousarray(xyz[:, sel])
    differ = 0
    for pbc in (1, 0):
        want_full = orc.mdmat_levels(xyz, x.boxes, sel, res_start, 1.0, 1001, pbc)
        if pbc:     # the stack the committed reference outputs were computed from
            assert hashlib.sha256(np.ascontiguousarray(want_full).tobytes()).hexdigest() == str(c["in/levels_sha256"])
            assert np.array_equal(want_full[[0, 236, 471]], c["in/levels_sample"])
        levels, _ = stage1.compute_levels(picked, x.boxes if pbc else None, geom, 1.0, 1001, pbc=bool(pbc))
        got = levels.view(torch.int16).cpu().numpy().view(np.uint16)[:, :geom.P]
        assert np.array_equal(got, orc.condensed(want_full)), "pbc=%d" % pbc
        if pbc:
            with_images = got.copy()
        else:
            differ = int((got != with_images).sum())
    # the box is 6.97 nm across and the chain is pulled to 25 nm: if images never mattered the two runs would
    # agree everywhere and the periodic path would be untested on this system
    print("levels that differ between pbc and no pbc:", differ)


def test_ubiquitin_cli_matches_live_reference(lib, tmp_path):
    """`conan.py` from the .xtc and the .pdb with fine.inp's keywords; expected files from the unmodified reference."""
    c = np.load(os.path.join(GOLDEN, "case_ubi.npz"))
    with open(str(tmp_path / "stages.txt"), "w") as f:
        f.write('#Three stages of the pulling run\n0 60000 "folded"\n60000 150000 "unfolding"\n150000 235500 "extended"\n')
    keywords = str(c["in/keywords"]).replace("RUN_MDMAT no\n", "TRAJ %s/ubi_pull_500ps.xtc\nCOORD %s/ubi.pdb\n" % (GOLDEN, GOLDEN))
    out = _run_cli(str(tmp_path), keywords)
    assert "Protein-H" in out and "472 distance matrices of 76 residues" in out
    n = _compare_tree(c, str(tmp_path))
    assert n >= 50
    # integer columns of the timeline exactly: first / last frame times, encounters
    tl = np.loadtxt(str(tmp_path / "aggregate" / "timeline.dat"))
    want = c["file/aggregate/timeline.dat"]
    assert np.array_equal(tl[:, [0, 1, 2, 3, 6]], want[:, [0, 1, 2, 3, 6]])
    for n in (1, 2, 3):
        assert (tmp_path / "matrices" / ("stage_%04d_dr.dat" % n)).exists()


# ------------------------------------------------------------------------------------------------
# configs[4]: two-chain interface
# ------------------------------------------------------------------------------------------------
def test_interface_small_matches_live_reference(lib, orc, tmp_path):
    """50 x 100 residues, 200 frames, K_TRAJ_CLUSTERS 2-5: stage 1 on the X x Y block, then the asymmetric
    stage 2 / clustering outputs against the unmodified reference."""
    import torch
    from conan_b200 import stage1, synth, xpm
    c = np.load(os.path.join(GOLDEN, "case_c5.npz"))
    nx, ny, nf, apr, seed, closer = c["in/synth"]
    nx, ny, nf = int(nx), int(ny), int(nf)
    xyz, res_start, _ = synth.make_trajectory(nx + ny, nf, int(apr), seed=int(seed), two_chain=(nx, ny), closer=float(closer))
    assert hashlib.sha256(np.ascontiguousarray(xyz).tobytes()).hexdigest() == str(c["in/xyz_sha256"]), \
        "the synthetic trajectory is not the one the fixture was made from"
    full = orc.mdmat_levels(xyz, None, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0)
    assert hashlib.sha256(np.ascontiguousarray(full).tobytes()).hexdigest() == str(c["in/levels_sha256"])
    # block mode computes exactly the X x Y entries
    geom = stage1.Geometry(res_start, 0, nx, nx, ny)
    levels, _ = stage1.compute_levels(xyz, None, geom, 1.0, 1001)
    got = levels.view(torch.int16).cpu().numpy().view(np.uint16)[:, :geom.P]
    assert np.array_equal(got, orc.block(full, 0, nx, nx, ny))
    # the symmetric stack of the same frames feeds the command line as dmf.xpm (RUN_MDMAT no, as the reference's
    # own asymmetric test does)
    gsym = stage1.Geometry(res_start)
    lsym, _ = stage1.compute_levels(xyz, None, gsym, 1.0, 1001)
    flat = lsym.view(torch.int16).cpu().numpy().view(np.uint16)[:, :gsym.P]
    sq = xpm.square_from_condensed(flat, nx + ny)
    assert np.array_equal(sq, full)
    with open(str(tmp_path / "dmf.xpm"), "w") as f:
        xpm.write_xpm(f, sq, c["in/times"], 1.0, 1001)
    _run_cli(str(tmp_path), str(c["in/keywords"]))
    assert _compare_tree(c, str(tmp_path)) >= 40


@pytest.fixture(scope="module")
def interface_fullsize(lib):
    from conan_b200 import stage1, synth
    nx, ny, nf = 200, 400, 20000
    xyz, res_start, _ = synth.make_trajectory(nx + ny, nf, 8, seed=5, two_chain=(nx, ny))
    geom = stage1.Geometry(res_start, 0, nx, nx, ny)
    levels, nb = stage1.compute_levels(xyz, None, geom, 1.0, 1001, batch=2500)
    return xyz, res_start, geom, levels, (nx, ny, nf)


def test_interface_fullsize_block_mode(interface_fullsize, orc):
    """200 x 400 residues x 20 000 frames in block mode: sampled frames against the oracle, stage-2 checksums,
    total_interactions.x / .y and the frame-frame distances with the asymmetric norm sqrt(xres * yres)
    (conan.py:2038-2039) against direct reductions of the stack."""
    import torch
    from conan_b200 import stage2, stage3
    from conan_b200.levels import legend_values, level_thresholds
    xyz, res_start, geom, levels, (nx, ny, nf) = interface_fullsize
    P = geom.P
    assert P == nx * ny
    frames = np.array([0, 3, 9999, nf - 1])
    full = orc.mdmat_levels(xyz[frames], None, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0)
    got = levels.view(torch.int16)[torch.from_numpy(frames).cuda()].cpu().numpy().view(np.uint16)[:, :P]
    assert np.array_equal(got, orc.block(full, 0, nx, nx, ny))
    values = legend_values(1.0, 1001)
    lo, hi, eq = level_thresholds(values, 0.5, 0.6)
    st = stage2.pair_stats(levels, P, 1001, lo, hi, eq)
    lv = levels[:, :P].view(torch.int16)
    assert int(st.sum_k.sum().item()) == int(lv.sum(dtype=torch.int64).item())
    assert int(st.n_int.sum(dtype=torch.int64).item()) == int((st.native.to(torch.int64) + st.nonnative).sum().item())
    f = 12345
    d = lv[f].to(torch.int64) - lv[f - 1].to(torch.int64)
    assert int(st.ssd_prev[f].item()) == int((d * d).sum().item())
    # per-residue interaction totals of the X and the Y chain (conan.py:937-947) are row / column sums of the
    # lifetime block
    life = (st.n_int.to(torch.float64) / nf).view(nx, ny)
    assert abs(float(life.sum(dim=1).sum().item()) - float(life.sum(dim=0).sum().item())) < 1e-9
    # frame-frame distances over the pairs that ever come within trunc, on a sample of frame pairs
    seen = torch.nonzero(st.sum_k < nf * 1000).flatten()
    assert 0 < len(seen) < P
    lut = stage3.lut16(values, levels.device)
    sub = torch.arange(0, nf, 40, device=levels.device)          # 500 frames
    rows = levels.view(torch.int16)[sub].contiguous().view(torch.uint16)
    mean_level = np.clip(np.rint(st.sum_k[seen].cpu().numpy() / nf), 0, 1000).astype(np.int64)
    centre = values.astype(np.float16).astype(np.float32)[mean_level]        # any centre serves a distance
    norm = float(np.sqrt(nx * ny))
    dist = stage3.frame_distances(rows, seen.cpu().numpy().astype(np.int32), 1001, lut, centre, norm)
    dist.fill_diagonal_(0.0)
    v = lut[rows.view(torch.int16)[:, seen].long()].to(torch.float64) / norm
    want = torch.cdist(v, v)
    err = (dist.to(torch.float64) - want).norm() / want.norm()
    assert float(err) < 1e-5, float(err)


def test_command_line_on_two_ranks_equals_one(lib, tmp_path):
    """`conan.py` started by torch.distributed.run with two ranks (frames sharded, per-pair sums merged, rank 0
    writes) gives the files of a single-process run.  Both ranks share this box's GPU, so the collectives go
    through gloo here; bench.py and tools/sharded_check.py exercise the same code over NCCL on several GPUs."""
    import socket
    text = ("TRAJ %s/ubi_pull_500ps.xtc\nCOORD %s/ubi.pdb\nGROUP 2\nNLEVEL 1001\nTRUNC 1.0\nTRUNC_INTER 0.5\n"
            "TRUNC_INTER_HIGH 0.6\nMATRICES no\nMAKE_MOVIE no\nPEARSON_TIME yes\nK_TRAJ_CLUSTERS 2\n" % (GOLDEN, GOLDEN))
    one, two = tmp_path / "one", tmp_path / "two"
    one.mkdir()
    two.mkdir()
    _run_cli(str(one), text)
    with open(str(two / "test.inp"), "w") as f:
        f.write(text)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    env = dict(os.environ, CONAN_DIST_BACKEND="gloo")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "conan.py"), "test.inp"], cwd=str(two), env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "Rank 1 of 2" in r.stdout
    for name in ("aggregate/timeline.dat", "aggregate/mdmat_average_rmsf.dat", "aggregate/native_contacts.dat",
                 "aggregate/time_mdmat_rmsd.dat", "aggregate/local_interactions.dat", "aggregate/pearson_data.dat",
                 "cluster_trj/2/assignment.dat", "cluster_trj/rmsd_interframe.dat"):
        a, b = np.loadtxt(str(one / name)), np.loadtxt(str(two / name))
        assert np.array_equal(a, b), name


def test_other_trajectory_formats_give_the_same_matrices(lib, tmp_path):
    """The reference's peptide trajectory rewritten as .trr (float and double), multi-frame .gro and multi-model
    .pdb: `conan.py` produces the same matrices and tables as from the .xtc (docs/index.md:143-149: whatever gmx
    reads).  The coordinates have three decimals, which every one of these formats keeps exactly."""
    from conan_b200 import topology, trajio
    from conan_b200.xtc import XtcFile
    x = XtcFile(os.path.join(GOLDEN, "sample.xtc"))
    xyz = x.decode()
    top = topology.read_gro(os.path.join(GOLDEN, "peptide.gro"))
    box = x.boxes
    files = {}
    files["trr"] = str(tmp_path / "t.trr")
    trajio.write_trr(files["trr"], xyz, box, x.times)
    files["trr64"] = str(tmp_path / "t64.trr")
    trajio.write_trr(files["trr64"], xyz.astype(np.float64), box, x.times, double=True)
    files["gro"] = str(tmp_path / "t.gro")
    with open(files["gro"], "w") as f:
        for k in range(len(xyz)):
            f.write("frame t= %.5f\n%5d\n" % (x.times[k], top.natoms))
            for a in range(top.natoms):
                f.write("%5d%-5s%5s%5d%8.3f%8.3f%8.3f\n" % (top.resids[a] % 100000, top.resnames[a][:5], top.names[a][:5], (a + 1) % 100000,
                                                           xyz[k, a, 0], xyz[k, a, 1], xyz[k, a, 2]))
            f.write("%10.5f%10.5f%10.5f\n" % (box[k, 0, 0], box[k, 1, 1], box[k, 2, 2]))
    files["pdb"] = str(tmp_path / "t.pdb")
    with open(files["pdb"], "w") as f:
        for k in range(len(xyz)):
            f.write("CRYST1%9.3f%9.3f%9.3f%7.2f%7.2f%7.2f P 1           1\nMODEL     %4d\n"
                    % (10 * box[k, 0, 0], 10 * box[k, 1, 1], 10 * box[k, 2, 2], 90, 90, 90, k + 1))
            for a in range(top.natoms):
                f.write("ATOM  %5d %-4s %-3s %1s%4d    %8.3f%8.3f%8.3f  1.00  0.00\n"
                        % ((a + 1) % 100000, top.names[a][:4], top.resnames[a][:3], "A", top.resids[a] % 10000,
                           10 * xyz[k, a, 0], 10 * xyz[k, a, 1], 10 * xyz[k, a, 2]))
            f.write("TER\nENDMDL\n")
    ref = tmp_path / "xtc"
    ref.mkdir()
    base = "COORD %s/peptide.gro\nTRUNC 1.0\nNLEVEL 1001\nMAKE_MOVIE no\n" % GOLDEN
    _run_cli(str(ref), "TRAJ %s/sample.xtc\n" % GOLDEN + base)
    for tag, path in files.items():
        d = tmp_path / tag
        d.mkdir()
        _run_cli(str(d), "TRAJ %s\n" % path + base)
        for name in ("matrices/00000.dat", "matrices/00007.dat", "aggregate/mdmat_average_rmsf.dat", "aggregate/native_contacts.dat"):
            a, b = np.loadtxt(str(ref / name)), np.loadtxt(str(d / name))
            if tag == "pdb" and "native" in name:
                a, b = a[:, 1:], b[:, 1:]                                  # .pdb models carry no times
            assert np.array_equal(a, b), (tag, name)
        tl_a, tl_b = np.loadtxt(str(ref / "aggregate/timeline.dat")), np.loadtxt(str(d / "aggregate/timeline.dat"))
        cols = [0, 1, 4, 6] if tag == "pdb" else [0, 1, 2, 3, 4, 6]          # .pdb models carry no times
        assert np.array_equal(tl_a[:, cols], tl_b[:, cols]), tag
