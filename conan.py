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

if __name__ == '__main__':
    main(sys.argv)
