"""Builds the C-ABI shared library in-tree with nvcc for sm_100a.

The library (conan_b200/lib/libconan_b200.so) travels to the GPU box with the snapshot; there is
no JIT cache and no pip install.  `python -m conan_b200.build` rebuilds when a source is newer.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIBPATH = os.path.join(LIBDIR, "libconan_b200.so")
SOURCES = ["api.cu", "xtc.cpp", "xtc_gpu.cu", "textio.cpp", "mdmat.cu", "pair_stats.cu", "dimer.cu", "dense_prep.cu", "eigen.cu", "gemm3x.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "--fmad=true", "-Xcompiler", "-fPIC,-O3,-pthread", "-cudart", "static",
]


def _stale():
    if not os.path.exists(LIBPATH):
        return True
    t = os.path.getmtime(LIBPATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "conan_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every translation unit for sm_100a and link the shared library."""
    if not force and not _stale():
        return LIBPATH
    nvcc = os.environ.get("NVCC", "nvcc")
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, os.path.splitext(src)[0] + ".o")
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
              ["-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj]
        subprocess.run(cmd, check=True)
        objs.append(obj)
    cmd = [nvcc, "-shared", "-cudart", "static", "-o", LIBPATH] + objs + ["-lpthread"]
    subprocess.run(cmd, check=True)
    return LIBPATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
