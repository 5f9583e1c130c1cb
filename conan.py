#!/usr/bin/env python3
"""CONAN contact analysis on B200: `./conan.py input.inp` and `import conan`.

The keyword file, the module-level function names and the output files are those of HITS-MBM/conan;
the trajectory contact path underneath is conan_b200 (CUDA, sm_100a).  See DESIGN.md.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conan_b200.frontend import *  # noqa: F401,F403
from conan_b200.frontend import main  # noqa: F401
# the reference's module-level helpers that work on the caller's own Python objects (frame dictionaries,
# aggregates rows, vectors, condensed distances), and the observable-file readers
from conan_b200.compat import (call_hierarchical, get_best_center, get_inter, plot_frames, plot_pc,  # noqa: F401
                               prepare_dimer_transpose, print_hierarchical_clusters, project_pc, read_frame,
                               read_next_time, read_pca, rmsd_frame, update_aggregate)
from conan_b200.analyses import read_in_observables, read_in_observables_blind  # noqa: F401

if __name__ == '__main__':
    main(sys.argv)
