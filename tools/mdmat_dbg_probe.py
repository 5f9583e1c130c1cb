#!/usr/bin/env python3
"""Where mdmat_tile_kernel's time goes: the kernel timed with parts switched off (CONAN_MDMAT_DEBUG bits:
1 no exact path, 2 no fill, 4 no survivor items, 8 no pruning).  Results are wrong with any bit set."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from conan_b200 import stage1, synth
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
dev = torch.device("cuda", 0)
xyz, res_start, _ = synth.make_trajectory(1000, nf, 8, seed=4)
geom = stage1.Geometry(res_start)
d_xyz = torch.from_numpy(xyz).to(dev)
work = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
levels = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
stream = torch.cuda.current_stream(dev)
os.environ["CONAN_MDMAT_SWZ"] = "1"
for dbg in (0, 1, 2, 3, 4, 6, 12, 14):
    os.environ["CONAN_MDMAT_DEBUG"] = str(dbg)
    stage1.mdmat_frames(d_xyz, None, geom, 1.0, 1001, False, levels, work, None)
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(3):
        stage1.mdmat_frames(d_xyz, None, geom, 1.0, 1001, False, levels, work, None, phases=2)
    b.record(stream)
    torch.cuda.synchronize(dev)
    print(json.dumps({"debug_bits": dbg, "ms_per_12500": a.elapsed_time(b) / 3 * 12500 / nf}), flush=True)
