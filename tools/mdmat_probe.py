#!/usr/bin/env python3
"""Times mdmat_tile_kernel under its CTA shapes / tile layouts on the benchmark shard and checks that
every variant writes the same level stack.

    python tools/mdmat_probe.py [--frames 4000] [--nres 1000] [--out gpurun_out/mdmat_probe.json]

CONAN_MDMAT_CFG (0: 256 threads x 2 CTAs per SM with 128 registers, 1: 256 x 3 with 80) and CONAN_MDMAT_SWZ (32-byte TMA swizzle
of the tiles) are read by the library on every call, so one process can walk through them.
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
    ap.add_argument("--atoms", type=int, default=8)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from conan_b200 import stage1, synth
    dev = torch.device("cuda", 0)
    xyz, res_start, _ = synth.make_trajectory(args.nres, args.frames, args.atoms, seed=4)
    geom = stage1.Geometry(res_start)
    d_xyz = torch.from_numpy(xyz).to(dev)
    nf = args.frames
    work = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
    counters = torch.zeros(3, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)
    ref = None
    rows = []
    for cfg in ("0", "1"):
        for swz in ("1", "0"):
            os.environ["CONAN_MDMAT_CFG"] = cfg
            os.environ["CONAN_MDMAT_SWZ"] = swz
            levels = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
            counters.zero_()
            stage1.mdmat_frames(d_xyz, None, geom, 1.0, 1001, False, levels, work, counters)
            torch.cuda.synchronize(dev)
            cnt = counters.cpu().numpy()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(args.reps):
                stage1.mdmat_frames(d_xyz, None, geom, 1.0, 1001, False, levels, work, None, phases=2)
            b.record(stream)
            torch.cuda.synchronize(dev)
            ms = a.elapsed_time(b) / args.reps
            a.record(stream)
            for _ in range(args.reps):
                stage1.mdmat_frames(d_xyz, None, geom, 1.0, 1001, False, levels, work, None, phases=1)
            b.record(stream)
            torch.cuda.synchronize(dev)
            ms_prep = a.elapsed_time(b) / args.reps
            lv = levels.view(torch.int16)
            same = True if ref is None else bool(torch.equal(lv, ref))
            if ref is None:
                ref = lv.clone()
            row = {"cfg": int(cfg), "swz": int(swz), "tile_ms": ms, "prep_ms": ms_prep, "frames": nf,
                   "us_per_frame": 1e3 * ms / nf, "ms_per_12500": ms * 12500 / nf, "same_levels": same,
                   "survivor_frac": float(cnt[1]) / (geom.P * nf), "atom_pairs": int(cnt[2]), "near": int(cnt[0]),
                   "tflops_8flop": 8.0 * cnt[2] / (ms * 1e-3) / 1e12}
            print(json.dumps(row), flush=True)
            rows.append(row)
            del levels
    if args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)
    if not all(r["same_levels"] for r in rows):
        raise SystemExit("variants disagree")


if __name__ == "__main__":
    main()
