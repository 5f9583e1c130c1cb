#!/usr/bin/env python3
"""Multi-GPU check (run under torchrun, one rank per GPU): frame shards + NCCL merge == one GPU.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/sharded_check.py
Every rank builds the same synthetic trajectory, processes its own contiguous frame range through
stage 1 and the sharded stage 2, and rank 0 compares the merged per-pair statistics and the
gathered per-frame lines with a single-GPU run over all frames (bit-exact integers).
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import sharded, stage1, stage2, synth  # noqa: E402
from conan_b200.levels import legend_values, level_thresholds  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    nres, nf, nlevels, trunc = 220, 96 * world + 17, 1001, 1.0
    xyz, res_start, _ = synth.make_trajectory(nres, nf, seed=31)
    geom = stage1.Geometry(res_start)
    lo, hi, eq = level_thresholds(legend_values(trunc, nlevels), 0.5, 0.6)
    bounds = np.linspace(0, nf, world + 1).astype(int)
    f0, f1 = int(bounds[rank]), int(bounds[rank + 1])
    mine, _ = stage1.compute_levels(xyz[f0:f1], None, geom, trunc, nlevels, device=dev)
    st = sharded.pair_stats_sharded(mine, geom.P, nlevels, lo, hi, eq, frame0=f0)
    torch.cuda.synchronize(dev)
    # gather the per-frame outputs on rank 0
    frame_fields = {}
    for name in ("native", "nonnative", "ssd_prev", "ssd_first"):
        t = getattr(st, name)
        sizes = [int(bounds[r + 1] - bounds[r]) for r in range(world)]
        parts = [torch.empty(s, dtype=t.dtype, device=dev) for s in sizes]
        dist.all_gather(parts, t) if len(set(sizes)) == 1 else [dist.broadcast(p if r != rank else t, src=r) for r, p in enumerate(parts)]
        parts[rank] = t
        frame_fields[name] = torch.cat(parts).cpu().numpy()
    ok = True
    if rank == 0:
        whole, _ = stage1.compute_levels(xyz, None, geom, trunc, nlevels, device=dev)
        ref = stage2.pair_stats(whole, geom.P, nlevels, lo, hi, eq).host()
        got = st.host()
        for k in ("sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f", "status"):
            same = np.array_equal(got[k], ref[k])
            ok &= same
            print("%-10s %s" % (k, "ok" if same else "MISMATCH"))
        for k, v in frame_fields.items():
            same = np.array_equal(v, ref[k])
            ok &= same
            print("%-10s %s" % (k, "ok" if same else "MISMATCH"))
        print("sharded_check world=%d: %s (n_enc max %d)" % (world, "PASS" if ok else "FAIL", int(ref["n_enc"].max())))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
