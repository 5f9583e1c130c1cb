"""Stage 1 parity on the GPU: conan_mdmat_frames through the C-ABI vs the CPU oracle, bit-exact."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _levels(xyz, boxes, geom, trunc, nlevels, pbc=False, **kw):
    from conan_b200 import stage1
    lv, nb = stage1.compute_levels(xyz, boxes, geom, trunc, nlevels, pbc=pbc, **kw)
    import torch
    return lv.view(torch.int16).cpu().numpy().view(np.uint16)[:, :geom.P], nb


def test_peptide_matches_reference_goldens(lib, goldens, peptide):
    """The 11 x 28 x 28 golden levels of the reference's own test, straight from the .xtc."""
    from conan_b200 import stage1
    from conan_b200.xtc import XtcFile
    _, sel, res_start = peptide
    x = XtcFile(os.path.join(GOLDEN, "sample.xtc"))
    xyz = x.decode(atom_sel=sel)
    geom = stage1.Geometry(res_start)
    for pbc in (False, True):
        lv, nb = _levels(xyz, x.boxes, geom, 1.0, 1001, pbc=pbc)
        iu = np.triu_indices(28, 1)
        assert np.array_equal(lv, goldens["levels_sym"][:, iu[0], iu[1]])
    geom = stage1.Geometry(res_start, x0=0, xres=14, y0=14, yres=14)
    lv, _ = _levels(xyz, x.boxes, geom, 1.0, 1001)
    assert np.array_equal(lv.reshape(11, 14, 14), goldens["levels_asym"])


@pytest.mark.parametrize("nres,nf,nlevels,trunc", [(300, 6, 1001, 1.0), (70, 9, 40, 1.5), (130, 5, 7744, 1.2)])
def test_synthetic_symmetric_bit_exact(lib, orc, nres, nf, nlevels, trunc):
    from conan_b200 import stage1, synth
    xyz, res_start, L = synth.make_trajectory(nres, nf, seed=3)
    geom = stage1.Geometry(res_start)
    want = orc.condensed(orc.mdmat_levels(xyz, None, np.arange(xyz.shape[1]), res_start, trunc, nlevels, 0))
    got, nb = _levels(xyz, None, geom, trunc, nlevels, batch=4)
    assert got.shape == want.shape
    assert np.array_equal(got, want), "%d mismatches" % (got != want).sum()
    assert 0 <= nb <= got.size * nlevels * 2e-6 + 5


def test_variable_residue_sizes_and_block_mode(lib, orc):
    from conan_b200 import stage1, synth
    xyz, res_start, L = synth.make_protein_like(97, 4)
    full = orc.mdmat_levels(xyz, None, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0)
    got, _ = _levels(xyz, None, stage1.Geometry(res_start), 1.0, 1001)
    assert np.array_equal(got, orc.condensed(full))
    for (x0, xres, y0, yres) in [(0, 40, 40, 57), (70, 27, 3, 66), (10, 30, 10, 30)]:
        geom = stage1.Geometry(res_start, x0=x0, xres=xres, y0=y0, yres=yres)
        got, _ = _levels(xyz, None, geom, 1.0, 1001)
        assert np.array_equal(got, orc.block(full, x0, xres, y0, yres)), (x0, xres, y0, yres)


def test_minimum_image_across_the_box(lib, orc):
    """A small box so that many contacts go through the periodic boundary."""
    from conan_b200 import stage1, synth
    xyz, res_start, L = synth.make_trajectory(120, 3, seed=9)
    box = 4.0
    xyz = np.mod(xyz, box).astype(np.float32)          # wrap atoms into the box (residues get broken)
    boxes = np.tile(np.eye(3, dtype=np.float32) * box, (3, 1, 1))
    want = orc.condensed(orc.mdmat_levels(xyz, boxes, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 1))
    plain = orc.condensed(orc.mdmat_levels(xyz, boxes, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0))
    assert (want != plain).sum() > 50                   # the periodic images matter here
    got, _ = _levels(xyz, boxes, stage1.Geometry(res_start), 1.0, 1001, pbc=True)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("npt", [False, True])
def test_whole_residues_wrapped_into_a_box(lib, orc, npt):
    """Residues kept whole but wrapped into the box frame by frame, so they jump across it inside a frame group
    and many contacts need a lattice shift in the fast path; with npt the box changes from frame to frame
    (the pruning then takes the smallest image over the group's boxes)."""
    from conan_b200 import stage1, synth
    nf = 8
    xyz, res_start, _ = synth.make_trajectory(150, nf, seed=12)
    xyz = xyz + np.linspace(0.0, 1.3, nf, dtype=np.float32)[:, None, None]      # a drift, so residues cross the faces
    lengths = 5.0 * (1.0 + (0.002 * np.arange(nf) if npt else np.zeros(nf)))
    boxes = np.zeros((nf, 3, 3), np.float32)
    wrapped = np.empty_like(xyz)
    for f in range(nf):
        boxes[f] = np.eye(3, dtype=np.float32) * np.float32(lengths[f])
        L = float(boxes[f, 0, 0])
        for r in range(len(res_start) - 1):
            a, b = res_start[r], res_start[r + 1]
            c = xyz[f, a:b].mean(axis=0)
            wrapped[f, a:b] = xyz[f, a:b] - np.floor(c / L).astype(np.float32) * np.float32(L)
    wrapped = wrapped.astype(np.float32)
    want = orc.condensed(orc.mdmat_levels(wrapped, boxes, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 1))
    plain = orc.condensed(orc.mdmat_levels(wrapped, boxes, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 0))
    assert (want != plain).sum() > 200                   # the periodic images matter here
    got, _ = _levels(wrapped, boxes, stage1.Geometry(res_start), 1.0, 1001, pbc=True)
    assert np.array_equal(got, want)
    # and as an X x Y block
    geom = stage1.Geometry(res_start, x0=10, xres=70, y0=60, yres=90)
    got, _ = _levels(wrapped, boxes, geom, 1.0, 1001, pbc=True)
    full = orc.mdmat_levels(wrapped, boxes, np.arange(xyz.shape[1]), res_start, 1.0, 1001, 1)
    assert np.array_equal(got, orc.block(full, 10, 70, 60, 90))


def test_edge_sizes(lib, orc):
    from conan_b200 import stage1
    rng = np.random.default_rng(2)
    # two residues of one atom each; a single frame
    xyz = np.array([[[0.0, 0, 0], [0.3, 0.4, 0]]], np.float32)
    got, _ = _levels(xyz, None, stage1.Geometry([0, 1, 2]), 1.0, 11)
    assert got.tolist() == [[5]]
    # residues larger than one slot chunk, count not a multiple of four
    sizes = [13, 1, 22, 5, 9]
    rs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    xyz = np.round(rng.uniform(0, 1.5, size=(7, rs[-1], 3)) * 1000) / 1000
    want = orc.condensed(orc.mdmat_levels(xyz, None, np.arange(rs[-1]), rs, 0.8, 501, 0))
    got, _ = _levels(xyz.astype(np.float32), None, stage1.Geometry(rs), 0.8, 501)
    assert np.array_equal(got, want)


def test_bad_arguments_are_errors(lib):
    import torch
    from conan_b200 import stage1
    from conan_b200._lib import ConanError
    geom = stage1.Geometry([0, 2, 4])
    xyz = torch.zeros((1, 4, 3), device="cuda")
    lv = torch.zeros((1, geom.stride), dtype=torch.uint16, device="cuda")
    work = torch.empty(stage1.mdmat_workspace_bytes(1, geom), dtype=torch.uint8, device="cuda")
    with pytest.raises(ConanError):
        stage1.mdmat_frames(xyz, None, geom, 1.0, 8000, False, lv, work)          # too many levels
    with pytest.raises(ConanError):
        stage1.mdmat_frames(xyz, None, geom, 1.0, 100, True, lv, work)            # pbc without a box
    with pytest.raises(ConanError):
        stage1.mdmat_frames(xyz, None, geom, 1.0, 100, False, lv, work[:256])     # workspace too small
    with pytest.raises(ConanError):
        stage1.mdmat_frames(xyz.cpu(), None, geom, 1.0, 100, False, lv, work)     # host tensor: no CPU path
