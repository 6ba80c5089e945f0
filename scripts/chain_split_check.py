#!/usr/bin/env python
"""one GPU: the row-split form of the decode chain at world = 1 (flag-in-data hand-over everywhere, b200_decode_chain_split) against the
separate-launch model and the default chain, same seeded weights and inputs; then the time of one step of each as a CUDA graph"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B
import bench

ctx = B.Context(0)
cfg = bench.MODELS[os.environ.get("MODEL", "8b")]
names = ("q", "k", "v", "o", "d", "logits")
res, ms = {}, {}
FORMS = [{"split": "split", "separate": False, "default": True}[f] for f in os.environ.get("FORMS", "split,separate,default").split(",")]
for form in FORMS:
    m = bench.HotPathModel(B, ctx, cfg, 1, 0, 0, 1, chain=form)
    assert (m.chains is not None) == bool(form), (form, B.lib().b200_last_error().decode())
    m.reset_state(); m.step(); ctx.synchronize()
    err = ctx.decode_chain_error()
    assert err == 0, hex(err)
    res[form] = {n: ctx.download(getattr(m, n)).reshape(-1).copy() for n in names}
    for _ in range(3):
        m.step()
    ctx.synchronize()
    ctx.graph_begin(); m.step(); g = ctx.graph_end()
    ctx.graph_launch(g); ctx.synchronize()
    e0, e1 = ctx.event(), ctx.event()
    ctx.record(e0)
    for _ in range(20):
        ctx.graph_launch(g)
    ctx.record(e1)
    ms[form] = ctx.elapsed_ms(e0, e1) / 20
    del m
if len(FORMS) < 3:
    print({str(k): round(v, 4) for k, v in ms.items()}, "ms per step (graph)"); sys.exit(0)
nm = lambda a, b: float(((a.astype(np.float64) - b) ** 2).sum() / max((a.astype(np.float64) ** 2).sum(), 1e-300))
for n in names:
    print(f"{n:7s} nmse split-form vs default chain {nm(res[True][n], res['split'][n]):.3e}   finite {np.isfinite(res['split'][n]).all()}   (default chain vs separate launches, which do not pass data from node to node: {nm(res[False][n], res[True][n]):.1e})")
print({str(k): round(v, 4) for k, v in ms.items()}, "ms per step (graph)")
bad = max(nm(res[True][n], res["split"][n]) for n in names)
assert bad < 1e-5, bad
print("OK")
