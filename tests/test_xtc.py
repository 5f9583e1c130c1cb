"""Host-side xtc index / decode / encode through the C-ABI, against the oracle's decoder."""
import os

import numpy as np
import pytest

from conftest import GOLDEN


def test_decode_matches_oracle_bit_for_bit(lib, orc):
    from conan_b200.xtc import XtcFile
    path = os.path.join(GOLDEN, "sample.xtc")
    x = XtcFile(path)
    times, boxes, xyz = orc.xtc_decode(open(path, "rb").read())
    assert (x.nframes, x.natoms) == (11, 399)
    assert np.array_equal(x.times, times) and np.array_equal(x.boxes, boxes)
    assert np.array_equal(x.decode(nthreads=1), xyz)
    assert np.array_equal(x.decode(nthreads=4), xyz)


def test_decode_frame_and_atom_selection(lib, orc):
    from conan_b200.xtc import XtcFile
    path = os.path.join(GOLDEN, "sample.xtc")
    x = XtcFile(path)
    _, _, xyz = orc.xtc_decode(open(path, "rb").read())
    sel = np.array([0, 5, 17, 398], np.int32)
    got = x.decode(frame_ids=[10, 0, 3], atom_sel=sel)
    assert np.array_equal(got, xyz[[10, 0, 3]][:, sel])
    from conan_b200._lib import ConanError
    with pytest.raises(ConanError):
        x.decode(frame_ids=[11])
    with pytest.raises(ConanError):
        x.decode(atom_sel=np.array([399], np.int32))


def test_encode_roundtrip_and_small_systems(lib, orc, tmp_path):
    from conan_b200.xtc import XtcFile, write_xtc
    rng = np.random.default_rng(0)
    for natoms in (3, 9, 10, 257):
        xyz = np.round(rng.uniform(-2, 9, size=(5, natoms, 3)) * 1000) / 1000
        p = str(tmp_path / ("t%d.xtc" % natoms))
        write_xtc(p, xyz, boxes=np.tile(np.eye(3) * 11.0, (5, 1, 1)), times=np.arange(5) * 2.0)
        x = XtcFile(p)
        assert (x.nframes, x.natoms) == (5, natoms)
        got = x.decode()
        assert np.abs(got - xyz).max() < 1e-6 + (0 if natoms <= 9 else 5e-7)
        # the oracle's independent decoder reads the same file identically
        _, _, oxyz = orc.xtc_decode(open(p, "rb").read())
        assert np.array_equal(got, oxyz)
        assert np.array_equal(x.times, np.arange(5) * 2.0)


def test_encode_uses_runs_on_clustered_atoms(lib, orc, tmp_path):
    """Protein-like coordinates: neighbouring atoms are close, so the encoder emits runs of small
    differences with an adapting code width; both decoders must read them back identically."""
    from conan_b200 import synth
    from conan_b200.xtc import XtcFile, write_xtc
    for seed, nres in ((5, 57), (9, 200)):
        xyz, _, L = synth.make_protein_like(nres, 7, seed=seed)
        p = str(tmp_path / ("p%d.xtc" % nres))
        nbytes = write_xtc(p, xyz, boxes=np.tile(np.eye(3) * L, (7, 1, 1)), times=np.arange(7) * 10.0)
        assert nbytes < 0.5 * xyz.size * 4          # compressed well below 12 bytes per atom
        x = XtcFile(p)
        got = x.decode()
        assert np.abs(got - xyz).max() < 1.5e-6
        _, _, oxyz = orc.xtc_decode(open(p, "rb").read())
        assert np.array_equal(got, oxyz)
    # wide systems at high precision: sizes above 2^24 switch to per-component bit fields (200 nm), sizes
    # between 2^23 and 2^24 give 70..72-bit packed triples and the widest small codes (100 nm)
    for width in (200.0, 100.0):
        xyz = np.round(np.random.default_rng(2).uniform(0, width, size=(2, 40, 3)) * 1000) / 1000
        p = str(tmp_path / "wide.xtc")
        write_xtc(p, xyz, precision=100000.0)
        got = XtcFile(p).decode()
        _, _, oxyz = orc.xtc_decode(open(p, "rb").read())
        assert np.array_equal(got, oxyz) and np.abs(got - xyz).max() < 5e-5


def test_truncated_file_is_an_error(lib, tmp_path):
    from conan_b200.xtc import XtcFile
    from conan_b200._lib import ConanError
    data = open(os.path.join(GOLDEN, "sample.xtc"), "rb").read()
    with pytest.raises(ConanError):
        XtcFile(data[:-5])


def test_encode_decode_property(lib, orc, tmp_path):
    """Randomised systems (clustered, spread, tiny and wide boxes, several precisions): the encoder's streams are
    read back identically by the host decoder and by the oracle's independent decoder, within half a quantum
    of the original coordinates."""
    from hypothesis import given, settings, strategies as st
    from conan_b200.xtc import XtcFile, write_xtc

    @settings(max_examples=40, deadline=None)
    @given(natoms=st.integers(1, 70), nframes=st.integers(1, 3), spread=st.sampled_from([0.02, 0.3, 3.0, 60.0]),
           cluster=st.sampled_from([0.0, 0.05, 0.4]), precision=st.sampled_from([10.0, 1000.0, 100000.0]),
           seed=st.integers(0, 2 ** 31 - 1))
    def check(natoms, nframes, spread, cluster, precision, seed):
        rng = np.random.default_rng(seed)
        centres = rng.uniform(-spread, spread, size=(nframes, max(1, natoms // 5 + 1), 3))
        xyz = centres[:, np.arange(natoms) // 5] + rng.normal(0, cluster, size=(nframes, natoms, 3))
        xyz = xyz.astype(np.float32)
        p = str(tmp_path / "h.xtc")
        write_xtc(p, xyz, precision=precision)
        x = XtcFile(p)
        got = x.decode()
        _, _, oxyz = orc.xtc_decode(open(p, "rb").read())
        assert np.array_equal(got, oxyz)
        tol = 0.0 if natoms <= 9 else 0.5 / precision * 1.0001 + np.abs(xyz).max() * 2.0e-7
        assert np.abs(got.astype(np.float64) - xyz.astype(np.float64)).max() <= tol
    check()
