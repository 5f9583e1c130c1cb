#!/usr/bin/env python3
"""Wall-clock profile of a whole CLI run (conan.py input file -> all outputs) on a synthetic system.

usage: python tools/cli_profile.py [nres] [nframes] [extra keyword lines ...]"""
import cProfile
import io
import os
import pstats
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from conan_b200 import synth, topology  # noqa: E402
from conan_b200.xtc import write_xtc  # noqa: E402

nres = int(sys.argv[1]) if len(sys.argv) > 1 else 300
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
extra = sys.argv[3:]
work = tempfile.mkdtemp(prefix="conan_cli_")
os.chdir(work)
xyz, res_start, L = synth.make_trajectory(nres, nf, seed=3)
write_xtc("traj.xtc", xyz, boxes=np.tile(np.eye(3, dtype=np.float32) * L, (nf, 1, 1)), times=np.arange(nf) * 100.0)
natoms = xyz.shape[1]
with open("top.gro", "w") as f:
    f.write("synthetic\n%d\n" % natoms)
    for a in range(natoms):
        r = a // 8
        f.write("%5d%-5s%5s%5d%8.3f%8.3f%8.3f\n" % ((r + 1) % 100000, "ALA", "C%d" % (a % 8), (a + 1) % 100000, *xyz[0, a]))
    f.write("%10.5f%10.5f%10.5f\n" % (L, L, L))
with open("run.inp", "w") as f:
    f.write("TRAJ traj.xtc\nCOORD top.gro\nNLEVEL 1001\nTRUNC 1.0\nTRUNC_INTER 0.5\nTRUNC_INTER_HIGH 0.6\nMAKE_MOVIE no\nGROUP 0\n")
    for line in extra:
        f.write(line.replace("=", " ") + "\n")
from conan_b200 import frontend  # noqa: E402
pr = cProfile.Profile()
t0 = time.time()
pr.enable()
try:
    frontend.main(["conan.py", "run.inp"])
except SystemExit:
    pass
pr.disable()
print("total wall %.2f s for %d residues x %d frames" % (time.time() - t0, nres, nf))
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28)
print("\n".join(l for l in s.getvalue().splitlines() if l.strip())[:6000])
