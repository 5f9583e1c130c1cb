"""The .tpr topology reader (conan_b200/topology.py) on the two run-input files the reference ships: test/md.tpr
(tpx 83, written by GROMACS 4.6.5) and applications/1_ubi_unfold/pull.tpr (tpx 100, 5.0.7, with solvent) -- what
`COORD md.tpr` / `COORD pull.tpr` of the reference's input files (test/*/test.inp, fine.inp:2) resolve to.  The
files themselves stay in the reference tree (the GPU box has none: the tests that run there use the .gro / .pdb
fixtures), so this test runs only where /root/reference is mounted."""
import os

import numpy as np
import pytest

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree is not mounted here")


@pytest.mark.parametrize("tpr,plain", [("test/md.tpr", "peptide.gro"), ("applications/1_ubi_unfold/pull.tpr", "ubi.pdb")])
def test_tpr_gives_the_atoms_and_groups_of_the_plain_topology(tpr, plain):
    from conan_b200 import topology
    a = topology.read_topology(os.path.join(REF, tpr))
    b = topology.read_topology(os.path.join(GOLDEN, plain))
    # the .tpr of the peptide holds the whole solvated system; its protein comes first
    n = b.natoms if plain.endswith(".pdb") else None
    ga, gb = dict(a.default_groups()), dict(b.default_groups())
    for name in ("Protein", "Protein-H", "C-alpha", "Backbone"):
        if name in gb:
            assert name in ga, name
            sa, sb = np.asarray(ga[name]), np.asarray(gb[name])
            assert np.array_equal(sa, sb), name
            assert np.array_equal(a.residue_runs(sa), b.residue_runs(sb)), name
    if n is not None:
        assert a.natoms == n
        # (.pdb files write hydrogens with the digit in front: 1HE2 for HE21)
        norm = lambda nm: nm[1:] + nm[0] if nm[:1].isdigit() else nm
        assert [norm(x) for x in a.names] == [norm(x) for x in b.names] and list(a.resnames) == list(b.resnames)
