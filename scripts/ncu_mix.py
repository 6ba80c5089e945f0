#!/usr/bin/env python
"""Instruction mix + top stall lines from an ncu report's source page.  usage: ncu_mix.py report.ncu-rep [kernel-index]"""
import csv, subprocess, sys
from collections import Counter
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
# split per kernel
kern, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "rows": []}; kern.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None: cur["rows"].append(r)
k = kern[int(sys.argv[2]) if len(sys.argv) > 2 else 0]
h = k["hdr"]; S, IE, ST = h.index("Source"), h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
tot = sum(int(r[IE]) for r in k["rows"]); stot = sum(int(r[ST]) for r in k["rows"])
c = Counter()
for r in k["rows"]:
    p = r[S].split(); op = p[1] if p and p[0].startswith("@") else (p[0] if p else "")
    c[op.split(".")[0]] += int(r[IE])
print(k["name"], "total warp-instructions", tot, "stall samples", stot)
for op, v in c.most_common(22): print(f"  {op:12s} {v:10d} {100*v/tot:5.1f}%")
print("top stall lines:")
for r in sorted(k["rows"], key=lambda r: -int(r[ST]))[:18]: print(f"  {int(r[ST]):6d} {100*int(r[ST])/max(stot,1):5.1f}%  x{int(r[IE]):8d}  {r[S].strip()[:90]}")
