import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    """The C-ABI library, built on demand (nvcc cross-compiles without a GPU)."""
    from conan_b200 import _lib, build
    build.build()
    return _lib.load()


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.load()
    return oracle


@pytest.fixture(scope="session")
def goldens():
    import numpy as np
    return np.load(os.path.join(GOLDEN, "ref_goldens.npz"))


@pytest.fixture(scope="session")
def peptide():
    """(topology, Protein-H selection, res_start) of the reference's test peptide."""
    from conan_b200 import topology
    top = topology.read_gro(os.path.join(GOLDEN, "peptide.gro"))
    sel = top.default_groups()[2][1]
    return top, sel, top.residue_runs(sel)
