#!/usr/bin/env python
"""SASS opcode histogram of every kernel object (cuobjdump -sass build/*.o): the committed proof of which Blackwell instructions each
kernel file uses (UTCHMMA/UTCIMMA = tcgen05.mma, UTMALDG = TMA tensor loads, UBLKCP = cp.async.bulk, LDTM/STTM = tcgen05.ld/st,
IMMA/HMMA = mma.sync, IDP = dp4a).  Runs without a GPU.  usage: sass_hist.py [--out profiles/r2_sass_histogram.txt]"""
import argparse, collections, glob, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser(); ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_sass_histogram.txt")); a = ap.parse_args()
KEY = ("UTCHMMA", "UTCIMMA", "UTCQMMA", "UTCOMMA", "UTMALDG", "UTMAPF", "UBLKCP", "UBLKPF", "LDTM", "STTM", "UTCBAR", "SYNCS", "IMMA", "HMMA", "IDP", "LDSM", "REDG", "ATOMG", "ACQBULK", "ELECT")
lines = ["# cuobjdump -sass opcode counts per object (sm_100a), key Blackwell / tensor / async opcodes first, then the ten most frequent", ""]
for obj in sorted(glob.glob(os.path.join(ROOT, "super-llama.cpp_b200", "build", "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", obj], stdout=subprocess.PIPE, text=True).stdout
    funcs = re.findall(r"Function : (\S+)", sass)
    ops = collections.Counter()
    for m in re.finditer(r"^\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", sass, re.M):
        ops[m.group(1)] += 1
    if not ops:
        continue
    key = "  ".join(f"{k} {sum(v for o, v in ops.items() if o.startswith(k))}" for k in KEY if any(o.startswith(k) for o in ops))
    top = "  ".join(f"{o} {v}" for o, v in ops.most_common(10))
    lines += [f"{os.path.basename(obj)}: {len(funcs)} kernels, {sum(ops.values())} instructions", f"    key: {key or '-'}", f"    top: {top}", ""]
open(a.out, "w").write("\n".join(lines))
print("\n".join(lines))
