"""Runs the UNMODIFIED reference (/root/reference/conan.py) in this container.

TEST INFRASTRUCTURE ONLY; needs /root/reference, so it is used by tests/golden/make_golden.py and by
`-m "not gpu"` tests that skip when the reference tree is absent (the GPU box has none).

The reference imports matplotlib (absent here) and `scipy.var` (removed from scipy); two import
shims are installed before runpy executes the script as __main__ in a scratch directory, exactly
as SURVEY.md section 8(c) describes.  Nothing of the reference is copied or modified.
"""
import contextlib
import io
import os
import runpy
import sys
import types

REFERENCE = "/root/reference/conan.py"


def available():
    return os.path.exists(REFERENCE)


def _install_shims():
    if "matplotlib.pyplot" not in sys.modules or getattr(sys.modules["matplotlib.pyplot"], "_conan_stub", False):
        mpl = types.ModuleType("matplotlib")
        plt = types.ModuleType("matplotlib.pyplot")
        plt._conan_stub = True
        for name in ("figure", "xlabel", "ylabel", "rc", "savefig", "show", "close"):
            setattr(plt, name, lambda *a, **k: None)
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    import numpy
    import scipy
    if not hasattr(scipy, "var"):
        scipy.var = numpy.var
    # dendrogram() would try to draw through the stub: ask for no plot
    import scipy.cluster.hierarchy as h
    if not getattr(h, "_conan_wrapped", False):
        orig = h.dendrogram
        h.dendrogram = lambda *a, **k: orig(*a, **{**k, "no_plot": True})
        h._conan_wrapped = True


def run_reference(workdir, input_text, stdin_text=""):
    """Execute conan.py with `input_text` as its keyword file inside workdir; return its globals."""
    _install_shims()
    os.makedirs(workdir, exist_ok=True)
    with open(os.path.join(workdir, "ref.inp"), "w") as f:
        f.write(input_text)
    old_cwd, old_argv, old_stdin = os.getcwd(), sys.argv, sys.stdin
    out = io.StringIO()
    try:
        os.chdir(workdir)
        sys.argv = [REFERENCE, "ref.inp"]
        sys.stdin = io.StringIO(stdin_text)
        with contextlib.redirect_stdout(out):
            g = runpy.run_path(REFERENCE, run_name="__main__")
    finally:
        os.chdir(old_cwd)
        sys.argv = old_argv
        sys.stdin = old_stdin
    g["_stdout"] = out.getvalue()
    return g
