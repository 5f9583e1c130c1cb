#!/usr/bin/env python3
"""Attribute ncu per-SASS instruction counts to CUDA source lines (needs -lineinfo builds).

usage: ncu_by_line.py report.ncu-rep kernel_substring cubin
Joins `ncu --page source --csv` (per-SASS counts, in program order) with `nvdisasm -g -c` line markers
by instruction order inside the kernel."""
import collections
import csv
import re
import subprocess
import sys

rep, kname, cubin = sys.argv[1:4]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# split per kernel
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
blk = [b for b in blocks if kname.split("IL")[0] in b["name"]][0]
hdr = blk["rows"][0]
data = blk["rows"][1:]
iI, iS, iW = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# find function
start = None
for i, l in enumerate(dis):
    if l.startswith(".text.") and kname in l:
        start = i
        break
lines, cur_line = [], 0
for l in dis[start + 1:]:
    if l.startswith(".text.") or l.startswith("\t.section") and lines:
        if l.startswith(".text."):
            break
    m = re.search(r'//## File ".*?", line (\d+)', l)
    if m:
        cur_line = int(m.group(1))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l):
        lines.append(cur_line)
print("sass in ncu %d, in disasm %d" % (len(data), len(lines)))
n = min(len(data), len(lines))
ex, st = collections.Counter(), collections.Counter()
for i in range(n):
    ex[lines[i]] += int(data[i][iI])
    st[lines[i]] += int(data[i][iW])
te, ts = sum(ex.values()), max(sum(st.values()), 1)
src = open(sys.argv[4]).read().splitlines() if len(sys.argv) > 4 else None
for ln, v in sorted(ex.items(), key=lambda kv: -kv[1])[:40]:
    print("%5d exec %6.3f stall %6.3f  %s" % (ln, v / te, st[ln] / ts, src[ln - 1].strip()[:90] if src and 0 < ln <= len(src) else ""))
