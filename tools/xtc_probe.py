#!/usr/bin/env python3
"""Times the two GPU .xtc decoders on the benchmark system at several batch sizes.

    python tools/xtc_probe.py [--nres 1000] [--frames 4200] [--out gpurun_out/xtc_probe.json]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nres", type=int, default=1000)
    ap.add_argument("--frames", type=int, default=4200)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from conan_b200 import _lib, synth
    from conan_b200.xtc import XtcFile, encode_xtc
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    xyz, _, _ = synth.make_trajectory(args.nres, args.frames, 8, seed=4)
    blob = encode_xtc(xyz, times=np.arange(args.frames, dtype=np.float32) * 100.0)
    x = XtcFile(bytes(blob))
    natoms = x.natoms
    d_data = torch.from_numpy(np.array(x._buf)).to(dev)
    d_off_all = torch.from_numpy(x.offsets).to(dev)
    host = x.decode(frame_ids=np.arange(min(64, args.frames)))
    stream = torch.cuda.current_stream(dev)
    rows = []
    for n in sorted({min(args.frames, v) for v in (32, 300, 1250, 2500, args.frames)}):
        d_off = d_off_all[:n].contiguous()
        d_out = torch.empty((n, natoms, 3), dtype=torch.float32, device=dev)
        d_status = torch.zeros(n, dtype=torch.int32, device=dev)
        d_work = torch.empty(int(lib.conan_xtc_decode_workspace_bytes(n, natoms)), dtype=torch.uint8, device=dev)

        def one():
            _lib.check(lib.conan_xtc_decode_device(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), n, natoms, None,
                                                   natoms, d_out.data_ptr(), d_status.data_ptr(), stream.cuda_stream))

        def two():
            _lib.check(lib.conan_xtc_decode_device_ws(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), n, natoms, None,
                                                      natoms, d_out.data_ptr(), d_status.data_ptr(), d_work.data_ptr(),
                                                      d_work.numel(), stream.cuda_stream))

        def timed(fn, reps=5):
            fn()
            torch.cuda.synchronize(dev)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(reps):
                fn()
            b.record(stream)
            torch.cuda.synchronize(dev)
            return a.elapsed_time(b) / reps
        ms1 = timed(one)
        ref = d_out.clone()
        d_out.zero_()
        ms2 = timed(two)
        ok = bool(torch.equal(ref, d_out)) and int(d_status.abs().sum().item()) == 0
        ok_host = bool(np.array_equal(d_out[:len(host)].cpu().numpy(), host[:n]))
        row = {"frames": n, "natoms": natoms, "one_pass_ms": ms1, "two_pass_ms": ms2, "same": ok, "same_as_host": ok_host,
               "two_pass_atoms_per_s": n * natoms / (ms2 * 1e-3), "ms_per_12500": ms2 * 12500 / n}
        print(json.dumps(row), flush=True)
        rows.append(row)
        del d_out, d_work, ref
    if args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)
    if not all(r["same"] and r["same_as_host"] for r in rows):
        raise SystemExit("decoders disagree")


if __name__ == "__main__":
    main()
