#!/usr/bin/env python3
"""Stage-1 time on a protein-like system (4..14 atoms per residue) next to the uniform 8-atom benchmark system,
without and with the periodic minimum image."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import stage1, synth
nres = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 1500
dev = torch.device("cuda")
for name, (xyz, rs, L) in (("uniform 8", synth.make_trajectory(nres, nf, seed=4)), ("ragged 4..14", synth.make_protein_like(nres, nf, seed=4))):
    geom = stage1.Geometry(rs)
    d_xyz = torch.from_numpy(np.ascontiguousarray(xyz, np.float32)).to(dev)
    levels = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
    work = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
    counters = torch.zeros(3, dtype=torch.int64, device=dev)
    for pbc in (False, True):
        d_box = torch.from_numpy(np.tile(np.eye(3, dtype=np.float32).reshape(1, 9) * np.float32(L), (nf, 1))).to(dev) if pbc else None
        counters.zero_()
        stage1.mdmat_frames(d_xyz, d_box, geom, 1.0, 1001, pbc, levels, work, counters)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            stage1.mdmat_frames(d_xyz, d_box, geom, 1.0, 1001, pbc, levels, work, None)
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 3
        c = counters.cpu().numpy()
        print("%-14s %s natoms %6d: %.3f ms per %d frames (%.2f us/frame)  survivors %.1f%%  evals %.3g  -> %.2f G atom pairs/s"
              % (name, "pbc" if pbc else "   ", xyz.shape[1], ms, nf, ms * 1e3 / nf, 100.0 * c[1] / (geom.P * nf), c[2], c[2] / ms / 1e6))
