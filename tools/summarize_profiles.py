#!/usr/bin/env python3
"""Turns the ncu outputs brought back in gpurun_out/ into the small text summaries kept under profiles/."""
import collections
import csv
import subprocess
import sys

launches, rep, tag = sys.argv[1:4]
rows = list(csv.reader(open(launches, errors="replace")))
for i, r in enumerate(rows):
    if r and r[0] == "ID":
        hdr, start = r, i + 1
        break
tot = collections.OrderedDict()
for r in rows[start:]:
    if len(r) < len(hdr):
        continue
    name = r[hdr.index("Kernel Name")].split("(")[0].replace("<unnamed>::", "").replace("void ", "")
    v = float(r[hdr.index("Metric Value")].replace(",", ""))
    e = tot.setdefault(name, [0.0, 0])
    e[0] += v
    e[1] += 1
s = sum(v[0] for v in tot.values())
with open("profiles/%s_launch_shares.txt" % tag, "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare shares)\n")
    f.write("# command: python bench.py --frames 1000 --steps 1 --warmup 1 --no-cpu-baseline --stage3-frames 4000\n")
    f.write("# kernel, launches, total_us, share\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0]):
        f.write("%s, %d, %.1f, %.3f\n" % (k, v[1], v[0] / 1e3, v[0] / s))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h, units = rr[0], rr[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
keys += [k for k in h if "issue_stalled" in k and "per_issue_active" in k]
with open("profiles/%s_ncu_full_summary.txt" % tag, "w") as f:
    f.write("# ncu --set full --clock-control none --import-source on; one block per captured launch\n")
    for r in rr[2:]:
        f.write("== %s  grid %s block %s\n" % (r[h.index("Kernel Name")], r[h.index("Grid Size")], r[h.index("Block Size")]))
        for k in keys:
            if k in h:
                v = r[h.index(k)]
                try:
                    if "issue_stalled" in k and float(v) < 0.15:
                        continue
                except ValueError:
                    pass
                f.write("   %s = %s %s\n" % (k, v, units[h.index(k)]))
print(open("profiles/%s_launch_shares.txt" % tag).read())
