"""Host readers of the plain-number trajectory formats (conan_b200/trajio.py): .trr (float / double), multi-frame
.gro, multi-model .pdb.  CPU only: what the readers return is what goes to the GPU kernel unchanged
(tests/test_gpu_configs.py checks the matrices computed from them)."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


@pytest.fixture(scope="module")
def peptide():
    from conan_b200.xtc import XtcFile
    x = XtcFile(os.path.join(GOLDEN, "sample.xtc"))
    return x.decode(), x.boxes, x.times


@pytest.mark.parametrize("double", [False, True])
def test_trr_round_trip(tmp_path, peptide, double):
    from conan_b200 import trajio
    xyz, boxes, times = peptide
    p = str(tmp_path / "t.trr")
    trajio.write_trr(p, xyz.astype(np.float64) if double else xyz, boxes, times, double=double)
    t = trajio.open_trajectory(p)
    assert t.nframes == len(xyz) and t.natoms == xyz.shape[1]
    got = t.coords(np.arange(t.nframes))
    assert got.dtype == np.float32 and np.array_equal(got, xyz)
    assert np.allclose(t.times, times) and np.allclose(np.asarray(t.boxes).reshape(-1, 3, 3), boxes)
    # a subset of frames in arbitrary order
    ids = np.array([t.nframes - 1, 0, 3])
    assert np.array_equal(t.coords(ids), xyz[ids])
    sel = np.arange(2, xyz.shape[1], 5)
    assert np.array_equal(t.coords(ids, sel), xyz[ids][:, sel])


def test_gro_and_pdb_frames(tmp_path, peptide):
    from conan_b200 import topology, trajio
    xyz, boxes, times = peptide
    top = topology.read_gro(os.path.join(GOLDEN, "peptide.gro"))
    gro = str(tmp_path / "t.gro")
    with open(gro, "w") as f:
        for k in range(len(xyz)):
            f.write("frame t= %.5f\n%5d\n" % (times[k], top.natoms))
            for a in range(top.natoms):
                f.write("%5d%-5s%5s%5d%8.3f%8.3f%8.3f\n" % (top.resids[a] % 100000, top.resnames[a][:5], top.names[a][:5],
                                                           (a + 1) % 100000, xyz[k, a, 0], xyz[k, a, 1], xyz[k, a, 2]))
            f.write("%10.5f%10.5f%10.5f\n" % (boxes[k, 0, 0], boxes[k, 1, 1], boxes[k, 2, 2]))
    t = trajio.open_trajectory(gro)
    assert t.nframes == len(xyz) and t.natoms == top.natoms
    # (the xtc decoder multiplies integers by the float32 reciprocal of the precision, the text goes through the nearest
    #  float32 of the decimal: an ulp apart at most)
    assert np.abs(t.coords(np.arange(t.nframes)) - xyz).max() < 1e-6
    assert np.allclose(t.times, times)
    pdb = str(tmp_path / "t.pdb")
    with open(pdb, "w") as f:
        for k in range(len(xyz)):
            f.write("MODEL     %4d\n" % (k + 1))
            f.write("CRYST1%9.3f%9.3f%9.3f%7.2f%7.2f%7.2f P 1           1\n" % (10 * boxes[k, 0, 0], 10 * boxes[k, 1, 1], 10 * boxes[k, 2, 2], 90, 90, 90))
            for a in range(top.natoms):
                f.write("ATOM  %5d %-4s %3s A%4d    %8.3f%8.3f%8.3f  1.00  0.00\n" % ((a + 1) % 100000, top.names[a][:4], top.resnames[a][:3],
                                                                                   top.resids[a] % 10000, 10 * xyz[k, a, 0], 10 * xyz[k, a, 1], 10 * xyz[k, a, 2]))
            f.write("ENDMDL\n")
    t = trajio.open_trajectory(pdb)
    assert t.nframes == len(xyz) and t.natoms == top.natoms
    assert np.abs(t.coords(np.arange(t.nframes)) - xyz).max() < 1.5e-4          # Angstrom with three decimals


def test_unknown_extension_is_refused(tmp_path):
    from conan_b200 import trajio
    p = str(tmp_path / "t.dcd")
    open(p, "wb").write(b"\0" * 64)
    with pytest.raises(Exception):
        trajio.open_trajectory(p)
