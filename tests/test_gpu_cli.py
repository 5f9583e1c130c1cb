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


@pytest.mark.parametrize("case", ["hyst", "equalcut", "default", "asym", "fine"])
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
