#!/usr/bin/env python3
"""Benchmark of the CONAN contact hot path: residue-pair-frames/s (min-dist + contact stats).

    python bench.py --gpus N --steps K --warmup W            this implementation (CUDA, sm_100a)
    python bench.py --impl reference --gpus N ...             the CPU restatement on the host cores

A step is one pass of the hot path over one batch of synthetic input: stage 1 (coordinates ->
uint16 level stack, conan_mdmat_frames) followed by stage 2 (streaming per-pair statistics,
conan_pair_stats).  The workload is the per-GPU shard of BASELINE.json's configs[3]: a synthetic
1000-residue system (8 heavy atoms per residue), 12 500 frames per GPU; at --gpus 8 this is the
full 100 000-frame configuration, frames sharded across GPUs (weak scaling), per-pair moments
merged with an NCCL all-reduce and the hysteresis state carried across shards.

One JSON line is printed by rank 0 (see the repository's bench contract):
  value      whole-job pair-frames/s with the coordinates already resident in HBM
  e2e        same metric through the C-ABI from HOST memory, starting where the reference starts: the shard's
             compressed .xtc bytes (pinned) are copied H2D batch by batch, decoded on the GPU
             (conan_xtc_decode_device), and the per-pair / per-frame results are read back, all inside
             the timed region
  roofline   the dominant kernel (mdmat_tile_kernel), timed alone with CUDA events
  cpu_baseline  the CPU oracle (a port: gmx is unavailable) on a bounded sample, one thread
  c2, c5     BASELINE.json configs[1] (the reference's ubiquitin unfolding run, whole) and configs[4]
             (200 x 400 residue interface, 20 000 frames, frame-frame distances), rank 0 only
On more than one GPU the frame shards are consecutive ranges of ONE trajectory, and before anything is
timed tools/sharded_check.py compares the sharded stage 2 / stage 3 with a single-GPU run ("sharded_check").
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NRES = 1000
ATOMS_PER_RES = 8
FRAMES_PER_GPU = 12500
TRUNC, NLEVELS = 1.0, 1001
TRUNC_INTER, TRUNC_INTER_HIGH = 0.5, 0.6
# the reference's own stage 2 (read_process_frames, one thread), measured in the survey container where
# /root/reference exists; the box that runs this file has no reference tree
REFERENCE_CONAN_PY = {"us_per_pair_frame": 32.0, "pair_frames_per_s": 1.0e6 / 32.0, "threads": 1,
                      "source": "BASELINE.md section 3: read_process_frames on 330 frames x 378 pairs, 3.96 s"}
METRIC = "residue-pair-frames/s (min-dist + contact stats)"
UNIT = "pair-frames/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=FRAMES_PER_GPU, help="frames per GPU")
    ap.add_argument("--nres", type=int, default=NRES)
    ap.add_argument("--batch", type=int, default=1250, help="frames per H2D batch of the e2e leg")
    ap.add_argument("--e2e-batches", type=int, default=3, help="batches of the e2e leg: the first of --batch frames, the rest equal")
    ap.add_argument("--cpu-frames", type=int, default=48, help="frames of the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stage3", action="store_true", help="skip the stage-3 (Pearson / covariance / frame-frame) leg")
    ap.add_argument("--stage3-frames", type=int, default=50000)
    ap.add_argument("--no-xtc", action="store_true", help="skip the trajectory-feed leg (xtc decode on the GPU vs host threads)")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs[1] (ubiquitin) and configs[4] (interface) legs")
    ap.add_argument("--config", default=None, choices=["c2", "c5"], help="run only that configuration's leg and print its line")
    ap.add_argument("--xtc-frames", type=int, default=2500)
    return ap.parse_args()


def measured_traffic(kernel, nres, pair_frames):
    """dram__bytes_read + dram__bytes_write of `kernel` per launch, scaled from the committed ncu capture of this
    build (profiles/r2_traffic.json, written by tools/summarize_profiles.py from an `ncu --set full` run of the
    benchmark shard); None when there is no capture for this system size."""
    path = os.path.join(ROOT, "profiles", "r2_traffic.json")
    try:
        with open(path) as f:
            rec = json.load(f).get(kernel)
    except (OSError, ValueError):
        return None
    if not rec or rec.get("nres") != nres:
        return None
    return rec["dram_bytes"] / rec["pair_frames"] * pair_frames


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def thresholds():
    from conan_b200.levels import legend_values, level_thresholds
    return level_thresholds(legend_values(TRUNC, NLEVELS), TRUNC_INTER, TRUNC_INTER_HIGH)


# --------------------------------------------------------------------------------------------
# CPU legs (oracle port; the one place besides tests/smoke that may execute oracle/)
# --------------------------------------------------------------------------------------------
def cpu_pass(xyz, res_start, threads):
    """Stage 1 + stage 2 of the CPU restatement on `threads` host threads; returns seconds."""
    from oracle import oracle
    oracle.load()
    lo, hi, eq = thresholds()
    nf, natoms = xyz.shape[0], xyz.shape[1]
    sel = np.arange(natoms)
    t0 = time.perf_counter()
    if threads <= 1:
        full = oracle.mdmat_levels(xyz, None, sel, res_start, TRUNC, NLEVELS, 0)
    else:
        parts = [None] * threads
        bounds = np.linspace(0, nf, threads + 1).astype(int)

        def work(i):
            if bounds[i + 1] > bounds[i]:
                parts[i] = oracle.mdmat_levels(xyz[bounds[i]:bounds[i + 1]], None, sel, res_start, TRUNC, NLEVELS, 0)

        ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        full = np.concatenate([p for p in parts if p is not None])
    lv = oracle.condensed(full)
    if threads <= 1:
        oracle.pair_stats(lv, lo, hi, eq)
    else:
        P = lv.shape[1]
        bounds = np.linspace(0, P, threads + 1).astype(int)
        ts = [threading.Thread(target=lambda a, b: oracle.pair_stats(np.ascontiguousarray(lv[:, a:b]), lo, hi, eq),
                               args=(bounds[i], bounds[i + 1])) for i in range(threads) if bounds[i + 1] > bounds[i]]
        [t.start() for t in ts]
        [t.join() for t in ts]
    return time.perf_counter() - t0


def run_reference(args):
    """--impl reference: the CPU restatement with every host core, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from conan_b200 import synth
    cores = os.cpu_count() or 1
    threads = min(cores, 64)
    nf = max(threads, min(args.cpu_frames * 4, 256))
    xyz, res_start, _ = synth.make_trajectory(args.nres, nf, ATOMS_PER_RES, seed=4)
    P = args.nres * (args.nres - 1) // 2
    # every step is one pass over the same bounded sample (sized so that --steps 20 --warmup 3 ends in minutes)
    for _ in range(max(0, args.warmup)):
        cpu_pass(xyz, res_start, threads)
    times = [cpu_pass(xyz, res_start, threads) for _ in range(max(1, args.steps))]
    t = float(np.mean(times))
    value = P * nf / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": max(0, args.warmup), "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32+u16", "data": "synthetic",
        "config": {"workload": "C4 shard sample: %d residues x %d frames (of %d per GPU)" % (args.nres, nf, args.frames),
                   "nres": args.nres, "atoms_per_res": ATOMS_PER_RES, "trunc": TRUNC, "nlevels": NLEVELS},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "%d frames of the %d-residue system, frames split over %d threads; gmx mdmat and "
                                   "conan.py themselves are unavailable on this box" % (nf, args.nres, threads),
                         "reference_conan_py": REFERENCE_CONAN_PY},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# GPU legs
# --------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from conan_b200 import _lib, stage1, stage2, synth, sharded

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for --impl ours)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    def mark(what):      # progress on stderr (CONAN_BENCH_TRACE=1): where a hung run stopped
        if os.environ.get("CONAN_BENCH_TRACE"):
            print("[bench rank %d %.1fs] %s" % (rank, time.time() - T0, what), file=sys.stderr, flush=True)
    T0 = time.time()

    nf, nres = args.frames, args.nres
    shard_check = None
    if world > 1:
        # correctness of the sharded path on this many GPUs, before anything is timed
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import sharded_check
        mark("sharded check")
        shard_check = sharded_check.check(dev)
        mark("sharded check done")
        ok = torch.tensor([1 if (shard_check is None or shard_check["ok"]) else 0], device=dev)
        dist.broadcast(ok, src=0)
        if not int(ok.item()):
            if rank == 0:
                print(json.dumps({"sharded_check": shard_check, "error": "sharded results differ from one GPU"}), flush=True)
            dist.destroy_process_group()
            raise SystemExit(1)
    # one long trajectory of nf * world frames (one seed); every rank takes its own consecutive frame range
    # (the slow global modes of the synthetic system repeat every shard length, so every rank's shard holds the same
    #  amount of work as the one-GPU run: weak scaling with fixed work per GPU)
    xyz, res_start, _ = synth.make_trajectory(nres, nf * world, ATOMS_PER_RES, seed=4, frame_range=(rank * nf, (rank + 1) * nf),
                                              mode_period=nf)
    geom = stage1.Geometry(res_start)
    P = geom.P
    lo, hi, eq = thresholds()

    h_xyz = torch.from_numpy(xyz).pin_memory()
    d_xyz = h_xyz.to(dev)
    levels = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
    counters = torch.zeros(3, dtype=torch.int64, device=dev)
    work1 = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
    work2 = torch.empty(stage2.pair_stats_workspace_bytes(nf, P), dtype=torch.uint8, device=dev)
    out = stage2.PairStats(nf, P, dev)
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    launches = [0]

    # The merge behind a shard's pass (packing kernels and two all-reduces, all small) runs on a stream of its own,
    # so the next step's distance kernel does not queue behind it; the pass of the next step waits for it (it
    # rewrites `out`), and so does the end of every timed region.
    merge_stream = torch.cuda.Stream(dev) if world > 1 else None
    merged = [None]

    def sharded_stats():
        if merged[0] is not None:
            stream.wait_event(merged[0])
        st = sharded.pair_stats_sharded(levels, P, NLEVELS, lo, hi, eq, frame0=rank * nf, out=out, work=work2,
                                        merge_stream=merge_stream)
        merged[0] = st.merged

    def join_merge():
        if merged[0] is not None:
            stream.wait_event(merged[0])

    def step_resident():
        stage1.mdmat_frames(d_xyz, None, geom, TRUNC, NLEVELS, False, levels, work1, counters)
        launches[0] += 2
        if world > 1:
            sharded_stats()
            launches[0] += 4           # shard summary + entering status + streaming pass + combine (collectives are NCCL's)
        else:
            stage2.pair_stats(levels, P, NLEVELS, lo, hi, eq, out=out, work=work2)
            launches[0] += 3

    # ---- device-resident throughput --------------------------------------------------------
    mark("resident warm-up")
    for _ in range(max(args.warmup, 3)):
        step_resident()
    barrier()
    mark("resident timing")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches[0] = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_resident()
    join_merge()                 # the timed region ends when the last step's merge has finished
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    n_launch = launches[0]

    # ---- kernel-level timings for the roofline (same stream, CUDA events) -------------------
    def time_region(fn, reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream)
        torch.cuda.synchronize(dev)
        return a.elapsed_time(b) / reps

    reps = max(2, min(args.steps, 5))
    counters.zero_()
    stage1.mdmat_frames(d_xyz, None, geom, TRUNC, NLEVELS, False, levels, work1, counters)
    torch.cuda.synchronize(dev)
    cnt = counters.cpu().numpy().astype(np.float64)
    ms_prep = time_region(lambda: stage1.mdmat_frames(d_xyz, None, geom, TRUNC, NLEVELS, False, levels, work1, None, phases=1), reps)
    ms_tile = time_region(lambda: stage1.mdmat_frames(d_xyz, None, geom, TRUNC, NLEVELS, False, levels, work1, None, phases=2), reps)
    ms_stats = time_region(lambda: stage2.pair_stats(levels, P, NLEVELS, lo, hi, eq, out=out, work=work2), reps)
    clocks = sampler.stop() if rank == 0 else None

    # ---- end to end from host memory: the shard's .xtc bytes in, the statistics out -----------------------
    from concurrent.futures import ThreadPoolExecutor
    from conan_b200.xtc import XtcFile, encode_xtc
    # A decode launch is bound by the length of one frame's flag walk (a few tenths of a millisecond however few
    # frames it has), and it finds little room beside the distance kernel (whose CTAs hold the whole register
    # file).  So the shard goes in few, large batches: a first one of --batch frames that gets the pipeline going,
    # the rest in two halves.
    first = max(1, min(args.batch, nf))
    rest = max(1, args.e2e_batches - 1)
    edges = [0, first] + [first + ((nf - first) * q + rest - 1) // rest for q in range(1, rest + 1)]
    edges = sorted(set(min(e, nf) for e in edges))
    nb = len(edges) - 1
    batch = max(edges[b + 1] - edges[b] for b in range(nb))
    def encode(b):      # one .xtc stream per batch (what a trajectory file split at those frames would hold)
        b0, b1 = edges[b], edges[b + 1]
        return encode_xtc(xyz[b0:b1], times=(np.arange(b0, b1, dtype=np.float32) + rank * nf) * 100.0)
    with ThreadPoolExecutor(max_workers=min(nb, os.cpu_count() or 1)) as pool:
        streams = list(pool.map(encode, range(nb)))
    h_bytes = [torch.from_numpy(b).pin_memory() for b in streams]
    h_offs = [torch.from_numpy(XtcFile(b.tobytes()).offsets.copy()).pin_memory() for b in streams]
    del streams
    max_bytes = max(t.numel() for t in h_bytes)
    # staging buffers and decoder scratch for two steps' worth of batches: the bytes of a step travel, and the run
    # flags of its frames are walked, while the step before is still computing
    NBUF = 2 * nb
    stage_buf = [torch.empty(max_bytes, dtype=torch.uint8, device=dev) for _ in range(NBUF)]
    stage_off = [torch.empty(batch, dtype=torch.int64, device=dev) for _ in range(NBUF)]
    dec_works = [torch.empty(int(lib.conan_xtc_decode_workspace_bytes(edges[b % nb + 1] - edges[b % nb], geom.natoms)),
                             dtype=torch.uint8, device=dev) for b in range(NBUF)]
    d_coords = [torch.empty((batch, geom.natoms, 3), dtype=torch.float32, device=dev) for _ in range(2)]
    d_status = torch.zeros(nf, dtype=torch.int32, device=dev)
    work_b = torch.empty(stage1.mdmat_workspace_bytes(batch, geom), dtype=torch.uint8, device=dev)
    copy_stream, dec_stream = torch.cuda.Stream(dev), torch.cuda.Stream(dev, priority=-1)
    walk_stream = torch.cuda.Stream(dev, priority=-1)
    d2h_stream = torch.cuda.Stream(dev)
    host_out = {n: torch.empty(getattr(out, n).shape, dtype=getattr(out, n).dtype).pin_memory()
                for n, _ in stage2.PairStats.PAIR_FIELDS + stage2.PairStats.FRAME_FIELDS}
    d2h_bytes = sum(t.numel() * t.element_size() for t in host_out.values())
    h2d_bytes = sum(t.numel() for t in h_bytes) + sum(t.numel() * 8 for t in h_offs)
    float_bytes = h_xyz.numel() * 4

    # streams: copies and walks run a step ahead, the decode of a batch's groups just ahead of its distance kernel
    bytes_free = [None] * NBUF   # event after the decode that read each staging buffer (also across steps)
    coords_free = [None, None]   # event after the stage-1 launch that read each coordinate buffer

    # timing experiments only (CONAN_BENCH_E2E_SKIP=copy,decode): leave parts of the feed out to see what they cost
    skip = []          # (filled in after the warm-up steps, which always run the whole feed)

    # The decoder's first pass (the walk over the run flags: one thread per frame, a millisecond of latency however
    # many frames, few warps) needs only the bytes, so the walks of a step's batches are started when the LAST distance
    # kernel of the step before has finished, i.e. beside its stage-2 kernel, which leaves room for them on every SM;
    # beside the distance kernel (which holds all registers) every walk CTA would keep one of its CTAs out.  The second
    # pass (one thread per group of atoms) runs per batch, just ahead of the batch's distance kernel.
    distances_done = [None]      # event behind the last distance kernel of the step before
    step_no = [0]

    def step_e2e():
        base = (step_no[0] % 2) * nb
        step_no[0] += 1
        walked = []
        for b in range(nb):
            b0, b1 = edges[b], edges[b + 1]
            k = base + b
            with torch.cuda.stream(copy_stream):
                if bytes_free[k] is not None:
                    copy_stream.wait_event(bytes_free[k])
                if "copy" not in skip:
                    stage_buf[k][: h_bytes[b].numel()].copy_(h_bytes[b], non_blocking=True)
                    stage_off[k][: b1 - b0].copy_(h_offs[b], non_blocking=True)
                copied = torch.cuda.Event()
                copied.record(copy_stream)
            with torch.cuda.stream(walk_stream):
                walk_stream.wait_event(copied)
                if distances_done[0] is not None:
                    walk_stream.wait_event(distances_done[0])
                if "decode" not in skip:
                    _lib.check(lib.conan_xtc_walk_device(stage_buf[k].data_ptr(), h_bytes[b].numel(), stage_off[k].data_ptr(), b1 - b0,
                                                         geom.natoms, None, geom.natoms, d_coords[b % 2].data_ptr(),
                                                         d_status[b0:b1].data_ptr(), dec_works[k].data_ptr(), dec_works[k].numel(),
                                                         walk_stream.cuda_stream))
                ev = torch.cuda.Event()
                ev.record(walk_stream)
                walked.append(ev)
        for b in range(nb):
            b0, b1 = edges[b], edges[b + 1]
            k, c = base + b, b % 2
            with torch.cuda.stream(dec_stream):
                dec_stream.wait_event(walked[b])
                if coords_free[c] is not None:
                    dec_stream.wait_event(coords_free[c])
                if "decode" not in skip:
                    _lib.check(lib.conan_xtc_groups_device(stage_buf[k].data_ptr(), h_bytes[b].numel(), b1 - b0, geom.natoms, None,
                                                           geom.natoms, d_coords[c].data_ptr(), dec_works[k].data_ptr(),
                                                           dec_works[k].numel(), dec_stream.cuda_stream))
                decoded = torch.cuda.Event()
                decoded.record(dec_stream)
            bytes_free[k] = decoded
            stream.wait_event(decoded)
            stage1.mdmat_frames(d_coords[c][: b1 - b0], None, geom, TRUNC, NLEVELS, False, levels[b0:b1], work_b, None)
            done = torch.cuda.Event()
            done.record(stream)
            coords_free[c] = done
        distances_done[0] = coords_free[(nb - 1) % 2]
        if results_home[0] is not None:
            stream.wait_event(results_home[0])      # the previous step's results have left `out`
        if world > 1:
            sharded_stats()
            join_merge()                     # (the results that go home are the merged ones)
        else:
            stage2.pair_stats(levels, P, NLEVELS, lo, hi, eq, out=out, work=work2)
        # the results go home on a stream of their own, so the next step's kernels do not queue behind the transfer
        computed = torch.cuda.Event()
        computed.record(stream)
        with torch.cuda.stream(d2h_stream):
            d2h_stream.wait_event(computed)
            for n in host_out:
                host_out[n].copy_(getattr(out, n), non_blocking=True)
            fetched = torch.cuda.Event()
            fetched.record(d2h_stream)
        results_home[0] = fetched

    results_home = [None]
    mark("e2e warm-up")
    for _ in range(2):
        step_e2e()
    barrier()
    assert int(d_status.abs().sum().item()) == 0, "the GPU decoder rejected a frame of its own encoder's stream"
    skip[:] = [w for w in os.environ.get("CONAN_BENCH_E2E_SKIP", "").split(",") if w]
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(args.steps):
        step_e2e()
    stream.wait_event(results_home[0])          # the timed region ends when the last step's results are on the host
    b.record(stream)
    barrier()
    ms_e2e = a.elapsed_time(b)
    mark("e2e done")

    # ---- reduce over ranks -----------------------------------------------------------------
    t = torch.tensor([ms, ms_e2e, ms_tile, ms_stats, ms_prep], dtype=torch.float64, device=dev)
    per_rank = [t.clone() for _ in range(world)]
    if world > 1:
        dist.all_gather(per_rank, t)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_tile, ms_stats, ms_prep = [float(v) for v in t.cpu()]
    per_rank = [[float(v) for v in r.cpu()] for r in per_rank]

    if rank == 0:
        pk, pk_kind = peaks()
        fp32_peak = ctypes.c_double()
        _lib.check(lib.conan_measure_fp32_peak(ctypes.byref(fp32_peak), stream.cuda_stream))
        units = float(P) * nf * world                      # pair-frames per step, whole job
        value = units * args.steps / (ms * 1e-3)
        e2e_value = units * args.steps / (ms_e2e * 1e-3)
        # stage 1 algorithmic work per launch (one GPU): 8 flop per evaluated atom pair (3 sub,
        # 3 mul, 2 add) for the residue pairs that survive pruning + 2 B written per pair-frame
        flops = 8.0 * cnt[2]
        tile_tflops = flops / (ms_tile * 1e-3) / 1e12
        stats_gbs = 2.0 * P * nf / (ms_stats * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+u16", "data": "synthetic",
            "config": {
                "workload": "C4 shard: %d residues x %d frames per GPU (BASELINE configs[3] = 8 such shards)" % (nres, nf),
                "nres": nres, "atoms_per_res": ATOMS_PER_RES, "pairs": P, "frames_per_gpu": nf,
                "trunc": TRUNC, "nlevels": NLEVELS, "trunc_inter": TRUNC_INTER, "trunc_inter_high": TRUNC_INTER_HIGH,
                "sharding": "frames x%d (consecutive ranges of one trajectory whose slow global modes repeat every %d frames)" % (world, nf),
                "l2": "inputs larger than L2 (%.1f GB coordinates, %.1f GB levels per GPU); no flush needed"
                      % (float_bytes / 1e9, levels.numel() * 2 / 1e9),
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "ms_per_step": ms_e2e / args.steps,
                    "input": "compressed .xtc bytes of the shard (%.2f B per atom instead of 12 as float32), pinned host memory, "
                             "decoded on the GPU; per rank" % (h2d_bytes / float(nf * geom.natoms)),
                    "h2d_gbs_per_rank": h2d_bytes / (ms_e2e / args.steps * 1e-3) / 1e9},
            "per_rank_ms": {"step": [r[0] / args.steps for r in per_rank], "e2e_step": [r[1] / args.steps for r in per_rank],
                            "mdmat_tile_kernel": [r[2] for r in per_rank], "pair_stats": [r[3] for r in per_rank]},
            "gpu_launches": n_launch,
            "clocks": clocks,
            "roofline": {
                "kernel": "mdmat_tile_kernel", "bound": "fp32", "achieved": tile_tflops, "peak": fp32_peak.value,
                "unit": "TFLOP/s", "frac": tile_tflops / fp32_peak.value if fp32_peak.value else None,
                "traffic": measured_traffic("mdmat_tile_kernel", nres, float(P) * nf), "traffic_unit": "bytes per launch",
                "ms": ms_tile,
                "peak_source": "FP32 FMA rate measured in this run (conan_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 "
                               "entry, profiles/r2_fp32_peak.json keeps the value of the round's runs",
                "algorithmic": "8 flop x %.4g atom pairs (%.1f%% of residue pairs survive pruning)"
                               % (cnt[2], 100.0 * cnt[1] / (float(P) * nf)),
                "hbm_view": {"achieved": 2.0 * P * nf / (ms_tile * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "unit": "GB/s",
                             "frac": 2.0 * P * nf / (ms_tile * 1e-3) / 1e9 / pk["hbm_gbs"], "peak_kind": pk_kind},
            },
            "roofline_stats": {"kernel": "pair_stats_kernel (+combine)", "bound": "hbm", "achieved": stats_gbs,
                               "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": stats_gbs / pk["hbm_gbs"],
                               "traffic": measured_traffic("pair_stats_kernel", nres, float(P) * nf), "traffic_unit": "bytes per launch",
                               "ms": ms_stats, "peak_kind": pk_kind},
            "stage_ms": {"prep_slots_kernel": ms_prep, "mdmat_tile_kernel": ms_tile, "pair_stats": ms_stats},
            "boundary_results": int(cnt[0]),
        }
        if shard_check is not None:
            line["sharded_check"] = shard_check
        if not args.no_cpu_baseline:
            ncpu = max(4, args.cpu_frames)
            t_cpu = cpu_pass(xyz[:ncpu], res_start, 1)
            line["cpu_baseline"] = {
                "value": float(P) * ncpu / t_cpu, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": "first %d frames of the same %d-residue shard (%.1f s of one host core; host has %d cores); "
                          "gmx mdmat itself: unavailable on this box" % (ncpu, nres, t_cpu, os.cpu_count() or 0)}
        if not args.no_xtc:
            line["xtc_feed"] = run_xtc_feed(args, dev, xyz)
        del d_xyz, levels, work1, work2, stage_buf, d_coords, work_b
        torch.cuda.empty_cache()
        if not args.no_stage3:
            line["stage3"] = run_stage3(args, dev)
        if not args.no_configs:
            line["c2"] = run_c2(dev)
            line["c5"] = run_c5(dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_xtc_feed(args, dev, xyz):
    """The trajectory feed of stage 1: the shard's first frames written as a real .xtc, decoded by the GPU
    kernel (compressed bytes copied from pinned host memory inside the timed region) and by the host decoder
    on all cores for comparison."""
    import tempfile
    import torch
    from conan_b200 import _lib
    from conan_b200.xtc import XtcFile, write_xtc
    lib = _lib.load()
    n = max(1, min(args.xtc_frames, xyz.shape[0]))
    path = os.path.join(tempfile.mkdtemp(prefix="conan_bench_"), "shard.xtc")
    nbytes = write_xtc(path, xyz[:n], times=np.arange(n, dtype=np.float32) * 100.0)
    x = XtcFile(path)
    natoms = x.natoms
    h_data = torch.from_numpy(np.array(x._buf)).pin_memory()
    d_data = torch.empty_like(h_data, device=dev)
    d_off = torch.from_numpy(x.offsets).to(dev)
    d_out = torch.empty((n, natoms, 3), dtype=torch.float32, device=dev)
    d_status = torch.zeros(n, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev)

    d_work = torch.empty(int(lib.conan_xtc_decode_workspace_bytes(n, natoms)), dtype=torch.uint8, device=dev)

    def decode():
        _lib.check(lib.conan_xtc_decode_device_ws(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), n, natoms, None, natoms,
                                                  d_out.data_ptr(), d_status.data_ptr(), d_work.data_ptr(), d_work.numel(),
                                                  stream.cuda_stream))

    def decode_one_pass():
        _lib.check(lib.conan_xtc_decode_device(d_data.data_ptr(), d_data.numel(), d_off.data_ptr(), n, natoms, None, natoms,
                                               d_out.data_ptr(), d_status.data_ptr(), stream.cuda_stream))

    def timed(fn, reps=3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn()
        torch.cuda.synchronize(dev)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream)
        torch.cuda.synchronize(dev)
        return a.elapsed_time(b) / reps

    d_data.copy_(h_data)
    ms_one_pass = timed(decode_one_pass)
    one_pass_out = d_out.clone()
    d_out.zero_()
    ms_kernel = timed(decode)
    same_decoders = bool(torch.equal(one_pass_out, d_out))
    del one_pass_out

    def feed():
        d_data.copy_(h_data, non_blocking=True)
        decode()
    ms_feed = timed(feed)
    assert int(d_status.abs().sum().item()) == 0
    threads = os.cpu_count() or 1
    host = np.empty((n, natoms, 3), np.float32)
    t0 = time.time()
    x.decode(out=host, nthreads=threads)
    t_host = time.time() - t0
    same = bool(np.array_equal(d_out.cpu().numpy(), host))
    os.remove(path)
    return {"workload": "first %d frames of the shard as .xtc (%d atoms, %.2f bytes per atom)" % (n, natoms, nbytes / n / natoms),
            "kernel": "xtc_walk_kernel + xtc_groups_kernel (conan_xtc_decode_device_ws)", "decode_ms": ms_kernel,
            "atoms_per_s": n * natoms / (ms_kernel * 1e-3),
            "one_pass": {"kernel": "xtc_decode_kernel (conan_xtc_decode_device)", "decode_ms": ms_one_pass,
                         "atoms_per_s": n * natoms / (ms_one_pass * 1e-3), "same_coordinates": same_decoders},
            "scratch_bytes": int(d_work.numel()),
            "h2d_plus_decode_ms": ms_feed, "compressed_bytes": int(nbytes), "float_bytes": int(n * natoms * 12),
            "host_decode": {"ms": t_host * 1e3, "threads": threads, "atoms_per_s": n * natoms / t_host},
            "bit_identical_to_host_decoder": same}


def run_stage3(args, dev):
    """BASELINE configs[2] (300 residues x 50k frames): Pearson vs 3 observables + time, covariance of the
    contacting pairs, and configs[4]-style frame-frame distances on a frame subset; CUDA-event timings."""
    import torch
    from conan_b200 import stage1, stage2, stage3, synth
    from conan_b200.levels import legend_values, level_thresholds
    nres, nf = 300, args.stage3_frames
    xyz, res_start, _ = synth.make_trajectory(nres, nf, ATOMS_PER_RES, seed=3)
    geom = stage1.Geometry(res_start)
    levels, _ = stage1.compute_levels(xyz, None, geom, TRUNC, NLEVELS, device=dev)
    del xyz
    values = legend_values(TRUNC, NLEVELS)
    lo, hi, eq = level_thresholds(values, TRUNC_INTER, TRUNC_INTER_HIGH)
    st = stage2.pair_stats(levels, geom.P, NLEVELS, lo, hi, eq).host()
    P = geom.P
    seen = np.flatnonzero(st["sum_k"] < nf * (NLEVELS - 1)).astype(np.int32)
    lut = stage3.lut16(values, dev)
    lut_h = values.astype(np.float16)
    shift = lut_h[np.clip(np.rint(st["sum_k"] / nf).astype(int), 0, NLEVELS - 1)].astype(np.float32)
    rng = np.random.default_rng(0)
    t = np.arange(nf, dtype=np.float64) * 100.0
    obs = np.stack([np.ones(nf), t - t.mean(), rng.normal(size=nf), np.cumsum(rng.normal(size=nf))]).astype(np.float32)
    stream = torch.cuda.current_stream(dev)

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize(dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream)
        torch.cuda.synchronize(dev)
        return a.elapsed_time(b) / reps

    d_shift = torch.from_numpy(shift).to(dev)
    d_obs = torch.from_numpy(obs).to(dev)
    T = torch.empty((4, P), dtype=torch.float64, device=dev)
    Q = torch.empty(P, dtype=torch.float64, device=dev)
    from conan_b200 import _lib
    lib = _lib.load()
    ms_pearson = timed(lambda: _lib.check(lib.conan_pearson_sums(levels.data_ptr(), geom.stride, nf, P, lut.data_ptr(), NLEVELS,
                                                                 d_obs.data_ptr(), 4, d_shift.data_ptr(), T.data_ptr(),
                                                                 Q.data_ptr(), stream.cuda_stream)))
    mean = (shift.astype(np.float64) + T[0].cpu().numpy() / nf)[seen]
    d_sel = torch.from_numpy(seen).to(dev)
    d_mean = torch.from_numpy(mean.astype(np.float32)).to(dev)
    ms_prep = timed(lambda: stage3.build_operand(levels, d_sel, NLEVELS, lut, d_mean, rows_are_frames=False), reps=2)
    hi_, lo_, _, sc = stage3.build_operand(levels, d_sel, NLEVELS, lut, d_mean, rows_are_frames=False)
    cov = torch.empty((len(seen), len(seen)), dtype=torch.float32, device=dev)
    ms_cov = timed(lambda: stage3.gemm_3xtf32(hi_, lo_, hi_, lo_, nf, alpha=1.0 / ((nf - 1) * sc * sc), symmetric=True, out=cov), reps=2)
    del hi_, lo_
    # top eigenpairs of that covariance (PCA_MAKE 3): restarted block Krylov vs the dense float64 factorisation
    stage3.top_eigenpairs(cov, 3)                  # (warm-up, as for every other kernel)
    torch.cuda.synchronize(dev)
    t0 = time.time()
    w_top, _ = stage3.top_eigenpairs(cov, 3)
    torch.cuda.synchronize(dev)
    t_eig = time.time() - t0
    t0 = time.time()
    w_dense = torch.linalg.eigvalsh(cov.double()).flip(0)[:3]
    torch.cuda.synchronize(dev)
    t_dense = time.time() - t0
    eig = {"k": 3, "ms": t_eig * 1e3, "dense_eigvalsh_float64_ms": t_dense * 1e3,
           "max_rel_diff": float(((w_top - w_dense).abs() / w_dense[0]).max().item())}
    del cov, w_dense
    # frame-frame distances on the first 20 000 frames (configs[4] has 20k frames)
    nff = min(nf, 20000)
    fh, fl, n2, sf = stage3.build_operand(levels[:nff], d_sel, NLEVELS, lut, d_mean, rows_are_frames=True, scale=1.0 / (nres / np.sqrt(2.0)))
    dist_out = torch.empty((nff, nff), dtype=torch.float32, device=dev)
    ms_ff = timed(lambda: stage3.gemm_3xtf32(fh, fl, fh, fl, len(seen), epilogue=1, alpha=1.0 / sf, row_norm2=n2, col_norm2=n2, symmetric=True,
                                             out=dist_out), reps=2)
    pk, kind = peaks()
    tf32_peak = pk["bf16_tflops"]             # the split parts are binary16: the MMAs run on the f16 / bf16 path
    # MMA flops issued: upper-triangle tiles (diagonal ones whole) of the kernel in use: 256 x 256 per CTA pair,
    # or 128 x 128 with CONAN_GEMM_CFG=0
    tile = 128 if os.environ.get("CONAN_GEMM_CFG", "1") == "0" else 256
    ntile = lambda n: (n + tile - 1) // tile
    mma_cov = 3 * 2.0 * (ntile(len(seen)) * (ntile(len(seen)) + 1) / 2) * tile * tile * ((nf + 63) // 64 * 64)
    mma_ff = 3 * 2.0 * (ntile(nff) * (ntile(nff) + 1) / 2) * tile * tile * ((len(seen) + 63) // 64 * 64)
    return {
        "workload": "C3: %d residues x %d frames, %d of %d pairs ever in contact" % (nres, nf, len(seen), P),
        "pearson": {"ms": ms_pearson, "observables": 3, "bound": "hbm", "achieved": 2.0 * P * nf / ms_pearson / 1e6, "unit": "GB/s",
                    "peak": pk["hbm_gbs"], "frac": 2.0 * P * nf / ms_pearson / 1e6 / pk["hbm_gbs"]},
        "covariance": {"ms": ms_cov, "operand_ms": ms_prep, "rows": int(len(seen)), "k": nf, "bound": "tensor",
                       "achieved": mma_cov / ms_cov / 1e9, "unit": "TFLOP/s (f16 MMA flops issued, 3 products)",
                       "peak": tf32_peak, "frac": mma_cov / ms_cov / 1e9 / tf32_peak,
                       "useful_tflops": 1.0 * len(seen) * len(seen) * nf / ms_cov / 1e9,
                       "peak_note": "measured bf16 burst (%s)" % kind},
        "pca_eigenpairs": eig,
        "frame_frame": {"ms": ms_ff, "rows": nff, "k": int(len(seen)), "bound": "tensor", "achieved": mma_ff / ms_ff / 1e9,
                        "unit": "TFLOP/s (f16 MMA flops issued, 3 products)", "peak": tf32_peak,
                        "frac": mma_ff / ms_ff / 1e9 / tf32_peak, "useful_tflops": 1.0 * nff * nff * len(seen) / ms_ff / 1e9},
    }


def run_c2(dev):
    """BASELINE.json configs[1]: the reference's ubiquitin unfolding trajectory (tests/golden/ubi_pull_500ps.xtc,
    472 frames x 1231 atoms, 76 residues / 602 heavy atoms) with fine.inp's settings: trajectory bytes -> periodic
    stage 1 -> stage 2 -> frame-frame distances for K_TRAJ_CLUSTERS, whole run, next to the CPU port of the same
    stages on one core and to the command-line run (`conan.py`, text outputs included)."""
    import shutil
    import tempfile
    import torch
    from conan_b200 import stage1, stage2, stage3, topology
    from conan_b200.levels import legend_values, level_thresholds
    from conan_b200.xtc import XtcFile
    golden = os.path.join(ROOT, "tests", "golden")
    traj, pdb = os.path.join(golden, "ubi_pull_500ps.xtc"), os.path.join(golden, "ubi.pdb")
    if not (os.path.exists(traj) and os.path.exists(pdb)):
        return {"unavailable": "tests/golden/ubi_pull_500ps.xtc missing"}
    values = legend_values(TRUNC, NLEVELS)
    lo, hi, eq = level_thresholds(values, 0.5, 0.5)        # TRUNC_INTER 0.5, no upper cut-off: no hysteresis buffer
    stream = torch.cuda.current_stream(dev)

    def whole_run():
        top = topology.read_topology(pdb)
        x = XtcFile(traj)
        sel = top.default_groups()[2][1]
        geom = stage1.Geometry(top.residue_runs(sel))
        d_xyz = x.decode_device(atom_sel=sel, device=dev)
        d_box = torch.from_numpy(np.ascontiguousarray(x.boxes).reshape(-1, 9)).to(dev)
        levels = torch.zeros((x.nframes, geom.stride), dtype=torch.uint16, device=dev)
        work = torch.empty(stage1.mdmat_workspace_bytes(x.nframes, geom), dtype=torch.uint8, device=dev)
        stage1.mdmat_frames(d_xyz, d_box, geom, TRUNC, NLEVELS, True, levels, work, None)
        st = stage2.pair_stats(levels, geom.P, NLEVELS, lo, hi, eq)
        h = st.host()
        seen = np.flatnonzero(h["sum_k"] < x.nframes * (NLEVELS - 1)).astype(np.int32)
        centre = values.astype(np.float16).astype(np.float32)[np.clip(np.rint(h["sum_k"][seen] / x.nframes).astype(int), 0, NLEVELS - 1)]
        D = stage3.frame_distances(levels, seen, NLEVELS, stage3.lut16(values, dev), centre, geom.nres / np.sqrt(2.0))
        D.cpu()
        return x.nframes, geom.P, x, sel, geom

    whole_run()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        nf, P, x, sel, geom = whole_run()
    torch.cuda.synchronize(dev)
    t_gpu = (time.perf_counter() - t0) / reps
    # the CPU port of stages 1 and 2 on the same frames, one core
    from oracle import oracle
    xyz = x.decode()
    t0 = time.perf_counter()
    full = oracle.mdmat_levels(xyz, x.boxes, sel, geom.res_start, TRUNC, NLEVELS, 1)
    oracle.pair_stats(oracle.condensed(full), lo, hi, eq)
    t_cpu = time.perf_counter() - t0
    # the command line, fine.inp's keywords (MATRICES no: 472 x 2 text files of 5776 rows are formatting, not analysis)
    work = tempfile.mkdtemp(prefix="conan_c2_")
    with open(os.path.join(work, "fine.inp"), "w") as f:
        f.write("TRAJ %s\nCOORD %s\nNLEVEL 1001\nDT 500\nTRUNC 1.0\nTRUNC_INTER 0.5\nTRUNC_DR 1.0\nDR_MODE init\n"
                "K_TRAJ_CLUSTERS 1-5\nMATRICES no\nMAKE_MOVIE no\n" % (traj, pdb))
    t0 = time.perf_counter()
    # a single-process run of the command line: under torchrun the launcher's variables would make it wait for
    # ranks that never come
    env = {k: v for k, v in os.environ.items()
           if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE", "GROUP_RANK", "ROLE_RANK", "ROLE_WORLD_SIZE",
                        "MASTER_ADDR", "MASTER_PORT", "GROUP_WORLD_SIZE", "ROLE_NAME", "OMP_NUM_THREADS")
           and not k.startswith("TORCHELASTIC_")}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "conan.py"), "fine.inp"], cwd=work, input="2\n", capture_output=True,
                       text=True, env=env, timeout=600)
    t_cli = time.perf_counter() - t0
    shutil.rmtree(work, ignore_errors=True)
    return {"workload": "C2: applications/1_ubi_unfold, %d frames x %d pairs (76 residues, 602 heavy atoms), periodic box" % (nf, P),
            "whole_run_s": t_gpu, "pair_frames_per_s": float(P) * nf / t_gpu,
            "stages": "xtc index + GPU decode + stage 1 (minimum image) + stage 2 + frame-frame distances, results on the host",
            "cpu_port_s": t_cpu, "cpu_port": "oracle stage 1 + stage 2, one core (gmx mdmat and conan.py: unavailable on this box; "
                                             "conan.py's stage 2 alone is %.0f s at 32 us per pair-frame)" % (32e-6 * P * nf),
            "cli_wall_s": t_cli, "cli_rc": r.returncode,
            "cli": "conan.py with fine.inp's keywords, K_TRAJ_CLUSTERS 1-5, MATRICES no; process start, import torch and "
                   "all text outputs included"}


def run_c5(dev):
    """BASELINE.json configs[4]: two chains of 200 and 400 residues, 20 000 frames, X x Y block only (80 000 pairs),
    stage 1 + stage 2 and the frame-frame distance matrix of K_TRAJ_CLUSTERS over the pairs that ever touch."""
    import torch
    from conan_b200 import stage1, stage2, stage3, synth
    from conan_b200.levels import legend_values, level_thresholds
    nx, ny, nf = 200, 400, 20000
    xyz, res_start, _ = synth.make_trajectory(nx + ny, nf, ATOMS_PER_RES, seed=5, two_chain=(nx, ny))
    geom = stage1.Geometry(res_start, 0, nx, nx, ny)
    P = geom.P
    values = legend_values(TRUNC, NLEVELS)
    lo, hi, eq = level_thresholds(values, TRUNC_INTER, TRUNC_INTER_HIGH)
    d_xyz = torch.from_numpy(xyz).to(dev)
    del xyz
    levels = torch.zeros((nf, geom.stride), dtype=torch.uint16, device=dev)
    work1 = torch.empty(stage1.mdmat_workspace_bytes(nf, geom), dtype=torch.uint8, device=dev)
    work2 = torch.empty(stage2.pair_stats_workspace_bytes(nf, P), dtype=torch.uint8, device=dev)
    out = stage2.PairStats(nf, P, dev)
    counters = torch.zeros(3, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev)

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize(dev)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            fn()
        b.record(stream)
        torch.cuda.synchronize(dev)
        return a.elapsed_time(b) / reps

    stage1.mdmat_frames(d_xyz, None, geom, TRUNC, NLEVELS, False, levels, work1, counters)
    ms1 = timed(lambda: stage1.mdmat_frames(d_xyz, None, geom, TRUNC, NLEVELS, False, levels, work1, None))
    # the stack of an interface is sparse (survivor_frac below): stage 2 runs its variant that skips stretches of
    # nothing but last levels, as the command line does for such stacks; the dense variant is timed beside it and must
    # give the same numbers
    ms2_dense = timed(lambda: stage2.pair_stats(levels, P, NLEVELS, lo, hi, eq, out=out, work=work2, sparse=False))
    h_dense = out.host()
    ms2 = timed(lambda: stage2.pair_stats(levels, P, NLEVELS, lo, hi, eq, out=out, work=work2, sparse=True))
    del d_xyz, work1
    h = out.host()
    same_stats = all(np.array_equal(h[k], h_dense[k]) for k in h)
    del h_dense
    seen = np.flatnonzero(h["sum_k"] < nf * (NLEVELS - 1)).astype(np.int32)
    centre = values.astype(np.float16).astype(np.float32)[np.clip(np.rint(h["sum_k"][seen] / nf).astype(int), 0, NLEVELS - 1)]
    lut = stage3.lut16(values, dev)
    norm = float(np.sqrt(nx * ny))                                       # conan.py:2038-2039
    d_sel = torch.from_numpy(seen).to(dev)
    d_mean = torch.from_numpy(centre).to(dev)
    ms_op = timed(lambda: stage3.build_operand(levels, d_sel, NLEVELS, lut, d_mean, rows_are_frames=True, scale=1.0 / norm), reps=2)
    fh, fl, n2, sf = stage3.build_operand(levels, d_sel, NLEVELS, lut, d_mean, rows_are_frames=True, scale=1.0 / norm)
    D = torch.empty((nf, nf), dtype=torch.float32, device=dev)
    ms_ff = timed(lambda: stage3.gemm_3xtf32(fh, fl, fh, fl, len(seen), epilogue=1, alpha=1.0 / sf, row_norm2=n2, col_norm2=n2,
                                             symmetric=True, out=D), reps=2)
    cnt = counters.cpu().numpy().astype(np.float64)
    total = ms1 + ms2 + ms_op + ms_ff
    return {"workload": "C5: %d x %d residue interface, %d frames, %d pairs (block mode), %d of them ever within trunc" % (nx, ny, nf, P, len(seen)),
            "stage1_ms": ms1, "stage2_ms": ms2, "stage2_dense_variant_ms": ms2_dense, "stage2_variants_agree": bool(same_stats),
            "operand_ms": ms_op, "frame_frame_ms": ms_ff, "frame_frame_rows": nf, "frame_frame_k": int(len(seen)),
            "survivor_frac": cnt[1] / (float(P) * nf), "pair_frames_per_s_stage12": float(P) * nf / ((ms1 + ms2) * 1e-3),
            "pair_frames_per_s_with_frame_frame": float(P) * nf / (total * 1e-3),
            "frame_frame_useful_tflops": 1.0 * nf * nf * len(seen) / ms_ff / 1e9}


def run_config_only(args):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for --impl ours)")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    rec = run_c2(dev) if args.config == "c2" else run_c5(dev)
    if args.config == "c2":
        value, ms = rec.get("pair_frames_per_s"), 1e3 * rec.get("whole_run_s", 0.0)
    else:
        value, ms = rec["pair_frames_per_s_with_frame_frame"], rec["stage1_ms"] + rec["stage2_ms"] + rec["operand_ms"] + rec["frame_frame_ms"]
    print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": 3, "warmup": 1, "ms_per_step": ms,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32+u16",
                      "data": "reference trajectory" if args.config == "c2" else "synthetic",
                      "config": {"workload": rec["workload"]}, args.config: rec}), flush=True)


def main():
    args = parse_args()
    # a run that stops making progress dumps every thread's Python stack and exits instead of holding the GPUs
    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("CONAN_BENCH_WATCHDOG", "1500")), exit=True, file=sys.stderr)
    if args.impl == "reference":
        run_reference(args)
    elif args.config:
        run_config_only(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
