#!/usr/bin/env python
"""Instruction mix of one kernel from an .ncu-rep's source page: executed warp instructions per opcode,
divided by the number of messages the launch processed.  Usage: tools/ncu_src_mix.py <rep> <n_msgs>"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, n_msgs = sys.argv[1], float(sys.argv[2])
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(r for r in rows if "Instructions Executed" in r)
ia, isrc, ist = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
iex, iwf = hdr.index("L1 Wavefronts Shared Excessive"), hdr.index("L1 Wavefronts Shared")
ops, stall, exc, wf, tot = collections.Counter(), collections.Counter(), collections.Counter(), collections.Counter(), 0
for r in rows:
    if len(r) <= ia or not r[ia].isdigit():
        continue
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[isrc])
    op = (m.group(1) if m else "?").split(".")[0]
    n = int(r[ia])
    ops[op] += n; tot += n
    stall[op] += int(r[ist] or 0)
    exc[op] += int(r[iex] or 0); wf[op] += int(r[iwf] or 0)
print("warp instructions per message: %.1f" % (tot / n_msgs))
for op, n in ops.most_common(30):
    print("%-10s %8.1f per message   stall samples %d" % (op, n / n_msgs, stall[op]))
print("shared-memory wavefronts (excessive / all) by opcode:")
for op, n in wf.most_common(6):
    print("  %-8s %d / %d" % (op, exc[op], n))
