#!/usr/bin/env python
"""Executed warp instructions per SOURCE LINE of one kernel: joins the executed counts of an .ncu-rep's
source page with nvdisasm's line table of the cubin the kernel came from (same instruction order).
Usage: tools/ncu_line_mix.py <rep> <cubin> <kernel-substring> <n_msgs> <source-file>"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, cubin, kern, n, srcfile = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4]), sys.argv[5]
txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
infn, cur, seq = False, None, []
for ln in txt:
    if ".section" in ln and kern in ln and ".text" in ln:
        infn = True
        continue
    if infn and ln.startswith(".section") and seq:
        break
    if not infn:
        continue
    m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        seq.append((cur, m.group(2)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(r for r in rows if "Instructions Executed" in r)
ia = hdr.index("Instructions Executed")
ex = [int(r[ia]) for r in rows if len(r) > ia and r[ia].isdigit()]
assert len(ex) == len(seq), (len(ex), len(seq))
by, byop = collections.Counter(), collections.defaultdict(collections.Counter)
for (loc, ins), cnt in zip(seq, ex):
    by[loc] += cnt
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ins)
    byop[loc][m.group(1) if m else "?"] += cnt
src = open(srcfile).read().split("\n")
base = srcfile.split("/")[-1]
print("warp instructions per message by source line (top 40)")
for loc, c in by.most_common(40):
    if loc is None:
        continue
    f, l = loc
    text = src[l - 1].strip()[:64] if f == base and l <= len(src) else f
    print("%7.1f  %s:%d  %-64s %s" % (c / n, f, l, text, dict((k, round(v / n, 1)) for k, v in byop[loc].most_common(4))))
