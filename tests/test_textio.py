"""Native text writers (conan_write_matrices / conan_write_xpm) against straight Python formatting of
the reference's file grammar (conan.py:570-597 write_matrix; dmf.xpm as read by conan.py:455-567)."""
import io
import os

import numpy as np
import pytest

from conan_b200 import _lib, levels as lv, tables, xpm


def _py_matrix_text(mat, xterm, yterm, dr):
    head = ("# i  j   r_ij\n# r_ij is the shortest distance between any atom of residues  i, j.\n" if not dr else
            "# i  j   dr_ij\n# dr_ij is the change (compared to the previous frame)\n"
            "# in shortest distance between any atom of residues  i, j.\n")
    out = [head]
    for i in range(mat.shape[0]):
        for j in range(mat.shape[1]):
            out.append("%4d %4d %9.4f\n" % (i + xterm, j + yterm, mat[i, j]))
        out.append("\n")
    return "".join(out)


def _write(stack, sym, xterm, xres, yterm, yres, values, dr_mode, directory, index0=0, first=None, prev=None, threads=3):
    lib = _lib.load()
    stack = np.ascontiguousarray(stack, np.uint16)
    values = np.ascontiguousarray(values, np.float64)
    _lib.check(lib.conan_write_matrices(stack.ctypes.data, stack.shape[1], stack.shape[0], index0, sym, xterm, xres,
                                        yterm, yres, values.ctypes.data, len(values), dr_mode,
                                        None if first is None else first.ctypes.data,
                                        None if prev is None else prev.ctypes.data, os.fsencode(directory), threads))


@pytest.mark.parametrize("nlevels,trunc", [(101, 1.0), (1001, 1.5)])
def test_matrices_symmetric(tmp_path, nlevels, trunc):
    rng = np.random.default_rng(3)
    nres, nf = 9, 5
    P = nres * (nres - 1) // 2
    stride = lv.padded_stride(P)
    stack = np.zeros((nf, stride), np.uint16)
    stack[:, :P] = rng.integers(0, nlevels, size=(nf, P))
    values = lv.legend_values(trunc, nlevels)
    lay = tables.Layout(nres=nres, nterm=4)
    _write(stack, 1, 4, nres, 4, nres, values, 3, str(tmp_path))
    mats = [lay.to_matrix(values[stack[n, :P]], diag=0.0) for n in range(nf)]
    for n in range(nf):
        assert (tmp_path / ("%05d.dat" % n)).read_text() == _py_matrix_text(mats[n], 4, 4, False)
        if n == 0:
            assert not (tmp_path / "00000_dr.init.dat").exists()
            continue
        assert (tmp_path / ("%05d_dr.init.dat" % n)).read_text() == _py_matrix_text(mats[n] - mats[0], 4, 4, True)
        assert (tmp_path / ("%05d_dr.prev.dat" % n)).read_text() == _py_matrix_text(mats[n] - mats[n - 1], 4, 4, True)


def test_matrices_block_and_batches(tmp_path):
    rng = np.random.default_rng(4)
    xres, yres, nf, nlevels = 3, 5, 6, 41
    stack = rng.integers(0, nlevels, size=(nf, xres * yres)).astype(np.uint16)
    values = lv.legend_values(0.8, nlevels)
    # two batches, the second one continuing the file numbering and the dr references
    _write(stack[:2], 0, 2, xres, 11, yres, values, 2, str(tmp_path), threads=1)
    _write(stack[2:], 0, 2, xres, 11, yres, values, 2, str(tmp_path), index0=2, first=stack[0].copy(), prev=stack[1].copy())
    for n in range(nf):
        mat = values[stack[n]].reshape(xres, yres)
        assert (tmp_path / ("%05d.dat" % n)).read_text() == _py_matrix_text(mat, 2, 11, False)
        if n:
            prev = values[stack[n - 1]].reshape(xres, yres)
            assert (tmp_path / ("%05d_dr.prev.dat" % n)).read_text() == _py_matrix_text(mat - prev, 2, 11, True)
            assert not (tmp_path / ("%05d_dr.init.dat" % n)).exists()


@pytest.mark.parametrize("nlevels", [40, 1001])
def test_xpm_matches_python_writer_and_parses(tmp_path, nlevels):
    rng = np.random.default_rng(5)
    nres, nf = 13, 4
    P = nres * (nres - 1) // 2
    stack = np.zeros((nf, lv.padded_stride(P)), np.uint16)
    stack[:, :P] = rng.integers(0, nlevels, size=(nf, P))
    times = np.array([0.0, 100.0, 250.0, 12345678.0])
    path = tmp_path / "dmf.xpm"
    xpm.write_xpm_stack(str(path), stack[:3], True, nres, nres, times[:3], 1.0, nlevels, nthreads=2)
    xpm.write_xpm_stack(str(path), stack[3:], True, nres, nres, times[3:], 1.0, nlevels, append=True)
    ref = io.StringIO()
    xpm.write_xpm(ref, xpm.square_from_condensed(stack[:, :P], nres), times, 1.0, nlevels)
    assert path.read_text() == ref.getvalue()
    t, lev, trunc, nl = xpm.read_xpm(str(path))[:4]
    assert nl == nlevels and abs(trunc - 1.0) < 1e-6
    np.testing.assert_array_equal(t, times)
    np.testing.assert_array_equal(lev, xpm.square_from_condensed(stack[:, :P], nres))


def test_xpm_block(tmp_path):
    rng = np.random.default_rng(6)
    xres, yres, nf, nlevels = 4, 7, 3, 101
    stack = rng.integers(0, nlevels, size=(nf, xres * yres)).astype(np.uint16)
    times = np.arange(nf) * 10.0
    path = tmp_path / "b.xpm"
    xpm.write_xpm_stack(str(path), stack, False, xres, yres, times, 1.0, nlevels)
    ref = io.StringIO()
    # write_xpm wants [f, y, x]
    xpm.write_xpm(ref, stack.reshape(nf, xres, yres).transpose(0, 2, 1), times, 1.0, nlevels)
    assert path.read_text() == ref.getvalue()


def test_writer_errors(tmp_path):
    lib = _lib.load()
    stack = np.zeros((1, 64), np.uint16)
    values = np.linspace(0, 1, 11)
    rc = lib.conan_write_matrices(stack.ctypes.data, 64, 1, 0, 1, 1, 3, 1, 3, values.ctypes.data, 11, 0, None, None,
                                  os.fsencode(str(tmp_path / "missing_dir")), 1)
    assert rc == -5 and b"cannot open" in lib.conan_last_error()
    rc = lib.conan_write_matrices(stack.ctypes.data, 64, 1, 0, 1, 1, 3, 1, 4, values.ctypes.data, 11, 0, None, None,
                                  os.fsencode(str(tmp_path)), 1)
    assert rc == -1


def test_pair_table_matches_python_formatting(tmp_path):
    rng = np.random.default_rng(8)
    lay = tables.Layout(xterm=3, xres=5, yterm=40, yres=7)
    a = rng.normal(0, 50, size=(5, 7))
    b = rng.integers(-5, 2000000, size=(5, 7)).astype(np.float64) + 0.75      # printed with %12i: truncated
    c = rng.uniform(0, 1, size=(5, 7))
    c[1, 2] = 0.1234565                                                         # a rounding tie in binary terms
    c[0, 0] = -0.0
    fmt = "%4d %4d %12i %12.6f %9.4f\n"
    head = "# a header\n# of two lines\n"
    path = tmp_path / "t.dat"
    tables.write_pair_table(str(path), lay, [b, a, c], fmt, head)
    want = head
    for i in range(5):
        for j in range(7):
            want += fmt % (i + 3, j + 40, b[i, j], a[i, j], c[i, j])
        want += "\n"
    assert path.read_text() == want
    # no value columns, and a format the writer must refuse
    tables.write_pair_table(str(path), lay, [], "%5i %5i\n")
    assert path.read_text().splitlines()[0] == "%5i %5i" % (3, 40)
    with pytest.raises(_lib.ConanError):
        tables.write_pair_table(str(path), lay, [a], "%4d %4d %s\n")
    with pytest.raises(_lib.ConanError):
        tables.write_pair_table(str(path), lay, [a, c], "%4d %4d %9.4f\n")
