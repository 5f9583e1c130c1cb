"""The C-ABI library loads without a GPU and exports every symbol include/conan_b200.h declares."""
import ctypes
import os
import re

import numpy as np

from conftest import GOLDEN, ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "conan_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(conan_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from conan_b200 import _lib
    names = _declared_symbols()
    assert "conan_mdmat_frames" in names and "conan_pair_stats" in names and "conan_xtc_decode" in names
    for n in names:
        assert hasattr(lib, n), "library does not export %s" % n
        assert n in _lib.SIGNATURES, "no ctypes signature for %s" % n
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_text(lib):
    assert lib.conan_version() >= 100
    from conan_b200 import _lib
    nf, na = ctypes.c_int64(), ctypes.c_int32()
    junk = np.frombuffer(b"not an xtc file at all, just some bytes to scan over......", dtype=np.uint8)
    rc = lib.conan_xtc_index(junk.ctypes.data, junk.size, 0, None, None, None, None, ctypes.byref(nf), ctypes.byref(na))
    assert rc == -3
    assert b"malformed" in lib.conan_last_error()


def test_workspace_helpers_reject_bad_sizes(lib):
    assert lib.conan_mdmat_workspace_bytes(0, 10, 2, 5) == 0
    assert lib.conan_mdmat_workspace_bytes(4, 8000, 1000, 8) > 4 * 8000 * 12
    assert lib.conan_pair_stats_workspace_bytes(0, 10) == 0
    assert lib.conan_pair_stats_workspace_bytes(100, 499500) > 0


def test_missing_library_fails_loudly(monkeypatch):
    from conan_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIBPATH", "/nonexistent/libconan_b200.so")
    try:
        _lib.load()
    except ImportError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("load() must fail when the library is missing")
