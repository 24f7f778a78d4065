#!/usr/bin/env python
"""SASS evidence for profiles/: per kernel of libgpuaudio.so the static instruction histogram (by
mnemonic, and by the pipe the SHA-256 analysis cares about) and the lines that show the TMA /
mbarrier / packed-FP32 forms.  Usage: tools/sass_excerpt.py [kernel-regex] > profiles/<name>.txt"""
import collections
import re
import subprocess
import sys

SO = "loqa_hub_b200/libgpuaudio.so"
pat = re.compile(sys.argv[1] if len(sys.argv) > 1 else "ingest_kernel|resample48_kernel|score_kernel")
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
kern, body = None, collections.OrderedDict()
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        kern = m.group(1)
        body[kern] = []
    elif kern and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        body[kern].append(ln)
ALU = ("LOP3", "SHF", "IADD3", "IADD", "LEA", "PRMT", "ISETP", "IMNMX", "VIMNMX", "SEL", "MOV", "POPC", "FLO", "BMSK", "SGXT", "LOP")
SHOW = ("UTMALDG", "UBLKCP", "SYNCS", "FFMA2", "FMUL2", "FADD2", "UTMAPF", "LDGSTS")
demangle = subprocess.run(["cu++filt"] + list(body), capture_output=True, text=True).stdout.splitlines()
for (k, lines), name in zip(body.items(), demangle):
    if not pat.search(name):
        continue
    ops = collections.Counter()
    for ln in lines:
        m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m:
            ops[m.group(1).split(".")[0]] += 1
    total = sum(ops.values())
    print("=" * 100)
    print(name[:160])
    print("static instructions: %d   (%.1f KB of code)" % (total, total * 16 / 1024))
    print("ALU-pipe mnemonics (%s): %d" % (", ".join(ALU), sum(v for o, v in ops.items() if o in ALU)))
    print("top mnemonics: " + ", ".join("%s %d" % kv for kv in ops.most_common(24)))
    print("TMA / mbarrier / packed FP32: " + ", ".join("%s %d" % (o, ops[o]) for o in SHOW if ops.get(o)))
    shown = 0
    for ln in lines:
        if any(s in ln for s in ("UTMALDG", "UBLKCP", "SYNCS", "FFMA2", "FMUL2")) and shown < 14:
            print("   " + ln.strip()[:150])
            shown += 1
