#!/usr/bin/env python3
"""One stage-1 launch on the uniform benchmark system (target for ncu); periodic unless the argument is nopbc."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import stage1, synth
nres, nf = 1000, 900
xyz, rs, L = synth.make_trajectory(nres, nf, seed=4)
geom = stage1.Geometry(rs)
dev = torch.device("cuda")
d_xyz = torch.from_numpy(np.ascontiguousarray(xyz, np.float32)).to(dev)
pbc = not (len(sys.argv) > 1 and sys.argv[1] == "nopbc")
d_box = torch.from_numpy(np.tile(np.eye(3, dtype=np.float32).reshape(1, 9) * np.float32(L), (nf, 1))).to(dev) if pbc else None
levels = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
work = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
for _ in range(2):
    stage1.mdmat_frames(d_xyz, d_box, geom, 1.0, 1001, pbc, levels, work, None)
torch.cuda.synchronize()
print("ok")
