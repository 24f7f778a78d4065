#!/usr/bin/env python
"""Dynamic instruction mix and hottest SASS of one kernel from an .ncu-rep (source page).
Usage: tools/ncu_hot.py <rep> <kernel-substring> [top]"""
import csv
import io
import subprocess
import sys
from collections import Counter

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
i, cur, hdr, data = 0, None, None, {}
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Kernel Name":
        cur = r[1]; hdr = rows[i + 1]; data[cur] = (hdr, []); i += 2; continue
    if cur and r:
        data[cur][1].append(r)
    i += 1
for name, (hdr, rs) in data.items():
    if pat not in name:
        continue
    si, ei, wi = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
    ops, samples, tot_e, tot_s = Counter(), Counter(), 0, 0
    lines = []
    for r in rs:
        try:
            e, s = int(r[ei]), int(r[wi])
        except ValueError:
            continue
        src = r[si].strip()
        op = src.split()[1] if src.startswith("@") else src.split()[0]
        ops[op] += e; samples[op] += s; tot_e += e; tot_s += s
        lines.append((s, e, src))
    print("KERNEL", name[:100], " warp-instr executed:", tot_e, " samples:", tot_s)
    print("  dynamic mix (share of executed warp instructions | share of stall samples):")
    for op, e in ops.most_common(top):
        print("   %-22s %6.2f%%  | %6.2f%%" % (op, 100.0 * e / tot_e, 100.0 * samples[op] / max(tot_s, 1)))
    print("  hottest instructions by samples:")
    for s, e, src in sorted(lines, reverse=True)[:12]:
        print("   %6.2f%%  x%-10d %s" % (100.0 * s / max(tot_s, 1), e, src[:90]))
