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

    # ---- stage 3 on the shards: covariance (contraction over frames, all-reduce) and frame-frame rows ----
    from conan_b200 import stage3
    values = legend_values(trunc, nlevels)
    lut = stage3.lut16(values, dev)
    merged = st.host()
    lut_h = np.asarray(values).astype(np.float16)
    shift = lut_h[np.clip(np.rint(merged["sum_k"] / nf).astype(np.int64), 0, nlevels - 1)].astype(np.float32)
    mean, nf_total = sharded.global_mean(mine, geom.P, nlevels, lut, shift)
    lo_level = np.empty(geom.P, np.int64)
    kmin, _ = stage3.level_range(mine, geom.P)
    t_kmin = torch.from_numpy(np.asarray(kmin, np.int64)).to(dev)
    dist.all_reduce(t_kmin, op=dist.ReduceOp.MIN)
    sel = np.flatnonzero(t_kmin.cpu().numpy() < nlevels - 1).astype(np.int32)        # pairs ever below trunc
    cov = sharded.covariance_sharded(mine, sel, nlevels, lut, mean[sel], nf_total)
    norm = nres / np.sqrt(2.0)
    block, row0 = sharded.frame_distances_sharded(mine, sel, nlevels, lut, mean[sel], norm)
    torch.cuda.synchronize(dev)
    if rank == 0:
        full_mean = stage3.linregress_all(whole, geom.P, nlevels, lut, np.zeros((1, nf)), shift)["mean"]
        ok3 = nf_total == nf and np.abs(full_mean - mean).max() < 1e-12
        cov1 = stage3.covariance(whole, sel, nlevels, lut, full_mean[sel])
        e_cov = (torch.linalg.norm((cov - cov1).double()) / torch.linalg.norm(cov1.double())).item()
        d1 = stage3.frame_distances(whole, sel, nlevels, lut, full_mean[sel], norm)
        e_d = (torch.linalg.norm((block - d1[row0:row0 + block.shape[0]]).double()) / torch.linalg.norm(d1[:block.shape[0]].double())).item()
        # and against float64 numpy on the float16 vectors
        vec = lut_h[whole.view(torch.int16).cpu().numpy().view(np.uint16)[:, sel]].astype(np.float64).T
        e_np = np.linalg.norm(cov.cpu().numpy() - np.cov(vec)) / np.linalg.norm(np.cov(vec))
        ok3 &= e_cov < 2e-6 and e_d < 2e-6 and e_np < 1e-5
        print("stage 3 shards: %d pairs; covariance vs one GPU %.2e, vs numpy float64 %.2e; frame-frame rows vs one GPU %.2e: %s"
              % (len(sel), e_cov, e_np, e_d, "PASS" if ok3 else "FAIL"))
        ok &= ok3
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
