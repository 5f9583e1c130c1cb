#!/usr/bin/env python3
"""Decode throughput of the .xtc feed: GPU kernel (conan_xtc_decode_device) vs the threaded host decoder.

usage: python tools/xtc_bench.py [nres] [nframes]"""
import os
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import _lib, synth  # noqa: E402
from conan_b200.xtc import XtcFile, write_xtc  # noqa: E402

nres = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 2500
xyz, _, L = synth.make_trajectory(nres, nf, seed=4)
path = os.path.join(tempfile.mkdtemp(), "t.xtc")
t = time.time()
nbytes = write_xtc(path, xyz, boxes=np.tile(np.eye(3, dtype=np.float32) * L, (nf, 1, 1)), times=np.arange(nf) * 100.0)
print("encoded %d frames x %d atoms: %.1f MB (%.2f B/atom) in %.1f s" % (nf, xyz.shape[1], nbytes / 1e6, nbytes / nf / xyz.shape[1], time.time() - t))
x = XtcFile(path)
out = np.empty((nf, x.natoms, 3), np.float32)
for nt in (1, os.cpu_count()):
    t = time.time()
    x.decode(out=out, nthreads=nt)
    dt = time.time() - t
    print("host decode, %2d threads: %.3f s  %.1f M atoms/s" % (nt, dt, nf * x.natoms / dt / 1e6))
dev = torch.device("cuda")
lib = _lib.load()
d_data = torch.from_numpy(np.array(x._buf)).to(dev)
d_off = torch.from_numpy(x.offsets).to(dev)
d_out = torch.empty((nf, x.natoms, 3), dtype=torch.float32, device=dev)
d_status = torch.zeros(nf, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(3):
    e0.record()
    _lib.check(lib.conan_xtc_decode_device(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), nf, x.natoms, None, x.natoms,
                                           d_out.data_ptr(), d_status.data_ptr(), st))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
print("GPU decode: %.3f ms  %.1f M atoms/s  (%.1f GB/s of coordinates out)" % (ms, nf * x.natoms / ms / 1e3, nf * x.natoms * 12 / ms / 1e6))
assert int(d_status.abs().sum().item()) == 0
assert np.array_equal(d_out.cpu().numpy(), out)
print("bit-identical to the host decoder")

# a real gmx-written stream (groups of ~7 atoms, a run flag on nine groups in ten): the reference's ubiquitin
# frames, each decoded many times over
real = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ubi_12.xtc")
if os.path.exists(real):
    xr = XtcFile(real)
    reps = 1000
    ids = np.tile(np.arange(xr.nframes), reps)
    d_data = torch.from_numpy(np.array(xr._buf)).to(dev)
    d_off = torch.from_numpy(xr.offsets[ids]).to(dev)
    d_out = torch.empty((len(ids), xr.natoms, 3), dtype=torch.float32, device=dev)
    d_status = torch.zeros(len(ids), dtype=torch.int32, device=dev)
    for rep in range(3):
        e0.record()
        _lib.check(lib.conan_xtc_decode_device(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), len(ids), xr.natoms, None,
                                               xr.natoms, d_out.data_ptr(), d_status.data_ptr(), st))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    t = time.time()
    hostr = xr.decode(frame_ids=ids)
    dt = time.time() - t
    print("ubiquitin stream x%d: GPU %.3f ms  %.1f M atoms/s;  host %d threads %.3f s  %.1f M atoms/s"
          % (reps, ms, len(ids) * xr.natoms / ms / 1e3, os.cpu_count(), dt, len(ids) * xr.natoms / dt / 1e6))
    assert np.array_equal(d_out.cpu().numpy(), hostr)
