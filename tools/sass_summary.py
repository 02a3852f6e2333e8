#!/usr/bin/env python
"""Opcode histogram per kernel of libromsb200.so (cuobjdump -sass): python tools/sass_summary.py > profiles/rN_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ucla-roms_b200", "libromsb200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
COLS = ["LDG.E.64", "LDG.E.128", "STG.E.64", "STG.E.128", "LDS", "STS", "UBLKCP", "UTMALDG", "SYNCS", "LDGSTS", "BAR", "DADD", "DMUL", "DFMA", "MUFU.RCP64H", "SHFL", "total"]
fn, rows = None, collections.OrderedDict()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"\(.*", "", fn).replace("void ", "")
        rows[fn] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
    if m and fn:
        op = m.group(1)
        rows[fn]["total"] += 1
        for c in COLS[:-1]:
            if op == c or op.startswith(c + ".") or (c in ("LDS", "STS", "BAR", "SHFL", "SYNCS", "UBLKCP", "UTMALDG", "LDGSTS", "DADD", "DMUL", "DFMA") and op.split(".")[0] == c):
                rows[fn][c] += 1
                break
print("# static SASS opcode counts per kernel (sm_100a), %s" % os.path.basename(so))
print("# UBLKCP = cp.async.bulk (TMA unit, 1-D bulk copy), SYNCS = mbarrier operations; fp64 arithmetic is compiled -fmad=false, so DFMA only comes from division/sqrt sequences")
print("%-46s " % "kernel" + " ".join("%9s" % c for c in COLS))
for fn, c in rows.items():
    if c["total"] < 40:
        continue
    print("%-46s " % fn[:46] + " ".join("%9d" % c[k] for k in COLS))
