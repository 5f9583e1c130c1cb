#!/usr/bin/env python3
"""Multi-GPU check (run under torchrun, one rank per GPU): frame shards + NCCL merge == one GPU.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/sharded_check.py
Every rank builds the same synthetic trajectory, processes its own contiguous frame range (the last shard
is shorter than the others) through stage 1 and the sharded stage 2, and rank 0 compares the merged per-pair
statistics and the gathered per-frame lines with a single-GPU run over all frames (bit-exact integers); a
band of pairs is forced to sit inside the hysteresis buffer for whole shards, so the entering status has to
travel across more than one shard.  Then stage 3 on the shards (covariance by contraction + all-reduce,
frame-frame rows by broadcast) against the one-GPU result.  bench.py runs check() before its timed region
whenever it is started on more than one GPU.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from conan_b200 import sharded, stage1, stage2, synth  # noqa: E402
from conan_b200.levels import legend_values, level_thresholds  # noqa: E402


def check(dev, verbose=False):
    """Needs an initialised process group; returns the result record on rank 0 (None elsewhere) and raises
    nothing: the caller decides what a mismatch means."""
    rank, world = dist.get_rank(), dist.get_world_size()
    nres, nlevels, trunc = 220, 1001, 1.0
    per = 96
    nf = per * world - 29                                   # the last shard is shorter
    xyz, res_start, _ = synth.make_trajectory(nres, nf, seed=31)
    geom = stage1.Geometry(res_start)
    P = geom.P
    lo, hi, eq = level_thresholds(legend_values(trunc, nlevels), 0.5, 0.6)
    f0, f1 = rank * per, min(nf, (rank + 1) * per)
    mine, _ = stage1.compute_levels(xyz[f0:f1], None, geom, trunc, nlevels, device=dev)
    whole = None
    if rank == 0:
        whole, _ = stage1.compute_levels(xyz, None, geom, trunc, nlevels, device=dev)

    def buffered(levels, g0):
        # every 9th pair: on in global frame 0 (half of them), then inside the buffer [lo, hi) until frame
        # 2.5 shards in, where a quarter of them leave it upwards
        lv = levels.view(torch.int16)
        fr = torch.arange(g0, g0 + lv.shape[0], device=dev)
        col = torch.arange(0, P, 9, device=dev)
        vals = torch.full((lv.shape[0], len(col)), (lo + hi) // 2, dtype=torch.int16, device=dev)
        vals[(fr == 0)[:, None] & (col % 2 == 0)[None, :]] = lo // 2
        vals[(fr >= int(2.5 * per))[:, None] & (col % 4 == 0)[None, :]] = nlevels - 1
        lv[:, col] = vals
        return levels

    mine = buffered(mine, f0)
    # once with the merge on a stream of its own (what the step loop of bench.py does), once in line: same numbers
    side = torch.cuda.Stream(dev)
    st2 = sharded.pair_stats_sharded(mine, P, nlevels, lo, hi, eq, frame0=f0, merge_stream=side)
    st2.merged.synchronize()
    st = sharded.pair_stats_sharded(mine, P, nlevels, lo, hi, eq, frame0=f0)
    torch.cuda.synchronize(dev)
    same_on_side_stream = all(bool(torch.equal(getattr(st, n), getattr(st2, n)))
                              for n in ("sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f", "status", "decided"))
    flag = torch.tensor([1 if same_on_side_stream else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    same_on_side_stream = bool(int(flag.item()))
    # the per-frame outputs of all ranks on rank 0
    sizes = [min(nf, (r + 1) * per) - r * per for r in range(world)]
    frame_fields = {}
    for name in ("native", "nonnative", "ssd_prev", "ssd_first"):
        t = getattr(st, name)
        padded = torch.zeros(per, dtype=t.dtype, device=dev)
        padded[:len(t)] = t
        parts = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(parts, padded)
        frame_fields[name] = torch.cat([p[:n] for p, n in zip(parts, sizes)]).cpu().numpy()
    rec = None
    if rank == 0:
        whole = buffered(whole, 0)
        ref = stage2.pair_stats(whole, P, nlevels, lo, hi, eq).host()
        got = st.host()
        bad = [k for k in ("sum_k", "sum_k2", "n_int", "n_enc", "first_f", "last_f", "status") if not np.array_equal(got[k], ref[k])]
        bad += [k for k, v in frame_fields.items() if not np.array_equal(v, ref[k])]
        carried = int((ref["n_int"][::18] >= min(nf, int(2.5 * per))).sum())      # pairs that stayed on through the buffer
        rec = {"world": world, "frames": nf, "pairs": P, "stage2": "bit-exact" if not bad else "MISMATCH: " + ",".join(bad),
               "status_carried_across_shards": carried, "n_enc_max": int(ref["n_enc"].max()),
               "merge_on_side_stream": "same" if same_on_side_stream else "DIFFERENT"}
        if verbose:
            print("stage 2 shards:", rec)

    # ---- stage 3 on the shards: covariance (contraction over frames, all-reduce) and frame-frame rows ----
    from conan_b200 import stage3
    values = legend_values(trunc, nlevels)
    lut = stage3.lut16(values, dev)
    merged = st.host()
    lut_h = np.asarray(values).astype(np.float16)
    shift = lut_h[np.clip(np.rint(merged["sum_k"] / nf).astype(np.int64), 0, nlevels - 1)].astype(np.float32)
    mean, nf_total = sharded.global_mean(mine, P, nlevels, lut, shift)
    kmin, _ = stage3.level_range(mine, P)
    t_kmin = torch.from_numpy(np.asarray(kmin, np.int64)).to(dev)
    dist.all_reduce(t_kmin, op=dist.ReduceOp.MIN)
    sel = np.flatnonzero(t_kmin.cpu().numpy() < nlevels - 1).astype(np.int32)        # pairs ever below trunc
    cov = sharded.covariance_sharded(mine, sel, nlevels, lut, mean[sel], nf_total)
    norm = nres / np.sqrt(2.0)
    block, row0 = sharded.frame_distances_sharded(mine, sel, nlevels, lut, mean[sel], norm)
    torch.cuda.synchronize(dev)
    if rank == 0:
        full_mean = stage3.linregress_all(whole, P, nlevels, lut, np.zeros((1, nf)), shift)["mean"]
        cov1 = stage3.covariance(whole, sel, nlevels, lut, full_mean[sel])
        e_cov = (torch.linalg.norm((cov - cov1).double()) / torch.linalg.norm(cov1.double())).item()
        d1 = stage3.frame_distances(whole, sel, nlevels, lut, full_mean[sel], norm)
        e_d = (torch.linalg.norm((block - d1[row0:row0 + block.shape[0]]).double()) / torch.linalg.norm(d1[:block.shape[0]].double())).item()
        # and against float64 numpy on the float16 vectors
        vec = lut_h[whole.view(torch.int16).cpu().numpy().view(np.uint16)[:, sel]].astype(np.float64).T
        e_np = np.linalg.norm(cov.cpu().numpy() - np.cov(vec)) / np.linalg.norm(np.cov(vec))
        rec.update({"mean_max_abs_diff": float(np.abs(full_mean - mean).max()), "cov_rel_frob": e_cov,
                    "cov_rel_frob_vs_numpy_f64": float(e_np), "ff_rel_frob": e_d, "stage3_pairs": int(len(sel))})
        rec["ok"] = bool(rec["stage2"] == "bit-exact" and same_on_side_stream and nf_total == nf and rec["mean_max_abs_diff"] < 1e-12 and
                         e_cov < 2e-6 and e_d < 2e-6 and e_np < 1e-5 and carried > 0)
        if verbose:
            print("stage 3 shards:", {k: rec[k] for k in ("cov_rel_frob", "cov_rel_frob_vs_numpy_f64", "ff_rel_frob", "stage3_pairs")})
            print("sharded_check world=%d: %s" % (world, "PASS" if rec["ok"] else "FAIL"))
    dist.barrier()
    return rec


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rec = check(dev, verbose=True)
    ok = torch.tensor([1 if (rec is None or rec["ok"]) else 0], device=dev)
    dist.broadcast(ok, src=0)
    dist.destroy_process_group()
    sys.exit(0 if int(ok.item()) else 1)


if __name__ == "__main__":
    main()
