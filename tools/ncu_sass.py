#!/usr/bin/env python
"""Dumps the SASS of one kernel of an .ncu-rep with executed counts and stall samples (development aid).

    python tools/ncu_sass.py x.ncu-rep [kernel-index] [min-executed]
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
mine = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
b = blocks[idx]
print("#", b["name"])
hdr = None
for r in b["rows"]:
    if r and r[0] == "Address":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    ex = int(float(r[hdr.index("Instructions Executed")] or 0))
    if ex < mine:
        continue
    smp = r[hdr.index("# Samples")]
    print(f"{r[0][-5:]} {ex:10d} {smp:>6s}  {r[1]}")
