#!/usr/bin/env python
"""per-launch-shape time budget from gemv_trace.py's json: python scripts/trace_agg.py tag [tag ...]"""
import json, sys, os
from collections import defaultdict
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for f in sys.argv[1:]:
    rows = json.load(open(os.path.join(ROOT, "gpurun_out", f + ".json")))
    agg = defaultdict(lambda: [0, 0.0, 0])
    for i, r in enumerate(rows):
        nxt = rows[i + 1]["wait"][0] if i + 1 < len(rows) else r["end"][2]
        k = (r["M"], r["K"], "mma" if r["mma"] else "dp4a", r["ctas"]); agg[k][0] += 1; agg[k][1] += nxt - r["wait"][0]; agg[k][2] += r["dq_cycles"]
    tot = sum(v[1] for v in agg.values())
    print(f, "total %.1f us" % tot)
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("   %-28s x%-3d %6.2f us each  %5.1f %%  dq %d" % (k, v[0], v[1] / v[0], 100 * v[1] / tot, v[2] // v[0]))
