#!/usr/bin/env python3
"""Times conan_pair_stats in both forms of its streaming pass on the benchmark shard's level stack and
checks that every output field agrees.

    python tools/stats_probe.py [--frames 4000] [--nres 1000]

CONAN_STATS_V1=1 selects the first form (every pair, every frame); the default form works only on the
groups of pairs that are not on the last level.
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=4000)
    ap.add_argument("--nres", type=int, default=1000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--debug", type=int, default=0, help="CONAN_STATS_DEBUG bits for the second form (timing experiments)")
    args = ap.parse_args()
    import torch
    from conan_b200 import stage1, stage2, synth
    from conan_b200.levels import legend_values, level_thresholds
    dev = torch.device("cuda", 0)
    xyz, res_start, _ = synth.make_trajectory(args.nres, args.frames, 8, seed=4)
    geom = stage1.Geometry(res_start)
    levels, _ = stage1.compute_levels(xyz, None, geom, 1.0, 1001, device=dev)
    nf, P = args.frames, geom.P
    lo, hi, eq = level_thresholds(legend_values(1.0, 1001), 0.5, 0.6)
    lv = levels.view(torch.int16)
    frac_max = float((lv[:, :P] == 1000).float().mean().item())
    # how much of the stack the second form has to work on: groups of 4 adjacent pairs x 16-frame stages
    # that show anything but the last level
    if P % 4 == 0 and nf % 16 == 0:
        nm = (lv[:, :P] != 1000).view(nf // 16, 16, P // 4, 4)
        print(json.dumps({"active_group_stages": float(nm.any(dim=3).any(dim=1).float().mean().item()),
                          "active_pair_stages": float(nm.any(dim=1).float().mean().item()),
                          "active_groups_per_frame": float(nm.any(dim=3).float().mean().item())}), flush=True)
        del nm
    work = torch.empty(stage2.pair_stats_workspace_bytes(nf, P), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)
    res = {}
    for form in ("1", "0"):
        os.environ["CONAN_STATS_V1"] = form[0]
        os.environ["CONAN_STATS_DEBUG"] = str(args.debug) if form == "0" else "0"
        out = stage2.PairStats(nf, P, dev)
        stage2.pair_stats(levels, P, 1001, lo, hi, eq, out=out, work=work)
        torch.cuda.synchronize(dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(args.reps):
            stage2.pair_stats(levels, P, 1001, lo, hi, eq, out=out, work=work)
        b.record(stream)
        torch.cuda.synchronize(dev)
        ms = a.elapsed_time(b) / args.reps
        res[form] = (ms, out.host())
        print(json.dumps({"form": form, "ms": ms, "GB/s": 2.0 * P * nf / ms / 1e6,
                          "ms_per_12500": ms * 12500 / nf, "frac_on_last_level": frac_max}), flush=True)
    bad = [k for k in res["1"][1] if not np.array_equal(res["1"][1][k], res["0"][1][k])]
    os.environ["CONAN_STATS_DEBUG"] = "0"
    print("fields that differ between the forms:", bad)
    if bad and not args.debug:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
