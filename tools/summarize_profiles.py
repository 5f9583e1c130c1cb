#!/usr/bin/env python3
"""Turns the ncu outputs that tools/profile_round.sh left in gpurun_out/ into the small summaries kept under
profiles/ (run here, after the gpurun call has brought them back):

    python tools/summarize_profiles.py r2

writes profiles/<tag>_launch_shares.txt, _launches.csv, _ncu_full_summary.txt, _traffic.json (read by bench.py),
_sass_markers.txt, _by_line_<kernel>.txt, and copies the bench lines of the same run.
"""
import collections
import csv
import glob
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")
LIB = os.path.join(ROOT, "conan_b200", "lib", "libconan_b200.so")

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
MARKERS = ["UTCHMMA", "LDTM", "UTCBAR", "UTMALDG", "UBLKCP", "LDGSTS", "SYNCS", "FFMA2", "FMUL2", "FMNMX3", "REDUX", "DFMA",
           "LDS.128", "ATOMS", "BAR.SYNC"]


def to_bytes(value, unit):
    v = float(value.replace(",", ""))
    return v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1.0)


def launch_shares(tag):
    src = os.path.join(OUT, "%s_launches.csv" % tag)
    rows = list(csv.reader(open(src, errors="replace")))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[start]
    tot = collections.OrderedDict()
    for r in rows[start + 1:]:
        if len(r) < len(hdr):
            continue
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("<unnamed>::", "").replace("void ", "")
        e = tot.setdefault(name, [0.0, 0])
        e[0] += float(r[hdr.index("Metric Value")].replace(",", ""))
        e[1] += 1
    s = sum(v[0] for v in tot.values())
    with open(os.path.join(PROF, "%s_launch_shares.txt" % tag), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare shares)\n")
        f.write("# command: python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-xtc (full-size step: "
                "12 500 frames x 499 500 pairs; stage-3 leg at C3 size)\n# kernel, launches, total_us, share\n")
        for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0]):
            f.write("%s, %d, %.1f, %.3f\n" % (k, v[1], v[0] / 1e3, v[0] / s))
    shutil.copyfile(src, os.path.join(PROF, "%s_launches.csv" % tag))


def full_summaries(tag):
    traffic = {}
    with open(os.path.join(PROF, "%s_ncu_full_summary.txt" % tag), "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on -k regex:<kernel> -c 1, one capture per kernel, first "
                "launch of\n# python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-xtc [--no-stage3]   "
                "(tools/profile_round.sh)\n# i.e. at the benched size: 12 500 frames x 499 500 pairs for stages 1-2, 300 residues "
                "x 50 000 frames for stage 3;\n# the decoder kernels on the first batch of the end-to-end leg.  Times are "
                "cold-cache and serialised.\n")
        for rep in sorted(glob.glob(os.path.join(OUT, "%s_*.ncu-rep" % tag))):
            raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
            rr = list(csv.reader(raw.splitlines()))
            if len(rr) < 3:
                continue
            h, units = rr[0], rr[1]
            for r in rr[2:]:
                name = r[h.index("Kernel Name")]
                f.write("== %s  grid %s block %s\n" % (name[:110], r[h.index("Grid Size")], r[h.index("Block Size")]))
                keys = KEYS + [k for k in h if "issue_stalled" in k and "per_issue_active" in k]
                for k in keys:
                    if k not in h:
                        continue
                    v = r[h.index(k)]
                    try:
                        if "issue_stalled" in k and float(v) < 0.15:
                            continue
                    except ValueError:
                        pass
                    f.write("   %s = %s %s\n" % (k, v, units[h.index(k)]))
                short = name.split("(")[0].replace("<unnamed>::", "").replace("void ", "").split("<")[0]
                if short in ("mdmat_tile_kernel", "pair_stats_kernel"):
                    rd = to_bytes(r[h.index("dram__bytes_read.sum")], units[h.index("dram__bytes_read.sum")])
                    wr = to_bytes(r[h.index("dram__bytes_write.sum")], units[h.index("dram__bytes_write.sum")])
                    grid = [int(x) for x in re.findall(r"\d+", r[h.index("Grid Size")])]
                    traffic[short] = {"dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr, "grid": grid}
    # the captures are of the benchmark shard: 1000 residues x 12 500 frames
    for k in traffic:
        traffic[k].update({"pair_frames": 499500.0 * 12500.0, "nres": 1000,
                           "source": "profiles/%s_ncu_full_summary.txt (ncu --set full of the benched build at the benched size)" % tag})
    if traffic:
        with open(os.path.join(PROF, "%s_traffic.json" % tag), "w") as f:
            json.dump(traffic, f, indent=1)


def sass_markers(tag):
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, counts = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = name.replace("(anonymous namespace)::", "").replace("void ", "").split("(")[0]
            cur = counts.setdefault(name, collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["total"] += 1
            for mk in MARKERS:
                if op == mk or op.startswith(mk + ".") or (mk in ("LDS.128", "BAR.SYNC") and op.startswith(mk)):
                    cur[mk] += 1
    with open(os.path.join(PROF, "%s_sass_markers.txt" % tag), "w") as f:
        f.write("# cuobjdump -sass conan_b200/lib/libconan_b200.so: SASS instruction counts per kernel (static, not executed "
                "counts)\n# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = cp.async.bulk.tensor "
                "(TMA tile load),\n# UBLKCP = cp.async.bulk, LDGSTS = cp.async, SYNCS = mbarrier ops, FFMA2 / FMUL2 = packed FP32 "
                "pairs, FMNMX3 = three-input min,\n# REDUX = warp reduction, DFMA = float64 FMA\n")
        for name, c in counts.items():
            f.write("%-36s total %5d  %s\n" % (name[:36], c["total"], "  ".join("%s %d" % (mk, c[mk]) for mk in MARKERS if c[mk])))


def by_line(tag):
    tmp = os.path.join("/tmp", "conan_cubins_%d" % os.getpid())
    os.makedirs(tmp, exist_ok=True)
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, capture_output=True)
    # (kernel, the mangled fragment that names the captured instantiation in the cubin, source file)
    for kernel, mangled, src in (("mdmat_tile_kernel", "mdmat_tile_kernelILb0ELi256ELi2E", "mdmat"),
                                 ("pair_stats_kernel", "pair_stats_kernelILb0ELb0E", "pair_stats"),
                                 ("xtc_groups_kernel", "xtc_groups_kernel", "xtc_gpu"), ("xtc_walk_kernel", "xtc_walk_kernel", "xtc_gpu"),
                                 ("pearson_sums_kernel", "pearson_sums_kernelILi4E", "dense_prep"),
                                 ("gemm3x_pair_kernel", "gemm3x_pair_kernel", "gemm3x")):
        rep = os.path.join(OUT, "%s_%s.ncu-rep" % (tag, kernel))
        cubin = os.path.join(tmp, "%s.sm_100a.cubin" % src)
        if not (os.path.exists(rep) and os.path.exists(cubin)):
            continue
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_by_line.py"), rep, mangled, cubin],
                           capture_output=True, text=True)
        if r.returncode != 0:
            print("by-line failed for", kernel, r.stderr[-300:])
            continue
        lines = open(os.path.join(ROOT, "conan_b200", "csrc", src + ".cu")).read().splitlines()
        with open(os.path.join(PROF, "%s_by_line_%s.txt" % (tag, src if kernel in ("mdmat_tile_kernel", "pair_stats_kernel") else kernel)), "w") as f:
            f.write("# share of executed warp instructions / of stall samples per source line of conan_b200/csrc/%s.cu (%s; ncu "
                    "--set full --import-source on at the benched size, tools/ncu_by_line.py)\n" % (src, kernel))
            for ln in r.stdout.splitlines()[:70]:
                m = re.match(r"\s*(\d+) exec\s+([\d.]+) stall\s+([\d.]+)", ln)
                if m and 0 < int(m.group(1)) <= len(lines):
                    f.write("%5d exec %6s stall %6s  %s\n" % (int(m.group(1)), m.group(2), m.group(3), lines[int(m.group(1)) - 1].strip()[:110]))
                else:
                    f.write(ln + "\n")
    shutil.rmtree(tmp, ignore_errors=True)


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    launch_shares(tag)
    full_summaries(tag)
    sass_markers(tag)
    by_line(tag)
    for name in ("bench_n1.json", "bench_reference.json"):
        src = os.path.join(OUT, "%s_%s" % (tag, name))
        if os.path.exists(src):
            lines = [ln for ln in open(src) if ln.startswith("{")]
            if lines:
                open(os.path.join(PROF, "%s_%s" % (tag, name)), "w").write(lines[-1])
    print(open(os.path.join(PROF, "%s_launch_shares.txt" % tag)).read()[:1500])


if __name__ == "__main__":
    main()
