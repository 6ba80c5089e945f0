#!/usr/bin/env python
"""Micro-benchmark of single MUL_MAT shapes through the C-ABI: cycles through enough distinct weight buffers to defeat
the 126 MB L2, times with CUDA events on the ctx stream (graph replay), prints GB/s per shape.
usage: gemv_bench.py [--types q4_K,q6_K] [--n 1] [--iters 200] [--ncu]   (--ncu: few launches, for profiling)"""
import argparse, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B

T = {"q4_0": B.Q4_0, "q8_0": B.Q8_0, "q4_K": B.Q4_K, "q5_K": B.Q5_K, "q6_K": B.Q6_K}
SH = [(4096, 4096), (4096, 1024), (4096, 14336), (14336, 4096), (4096, 128256)]      # (K, M)

def blocks(rng, t, n):
    raw = rng.integers(0, 256, size=(n, B.TSIZE[t]), dtype=np.uint8)
    h = lambda v: np.asarray(v, dtype=np.float16).view(np.uint8).reshape(n, 2)
    if t in (B.Q4_0, B.Q8_0): raw[:, 0:2] = h(rng.uniform(-1, 1, n) * 1e-2)
    elif t in (B.Q4_K, B.Q5_K): raw[:, 0:2] = h(rng.uniform(.1, 1, n) * 1e-3); raw[:, 2:4] = h(rng.uniform(.1, 1, n) * 1e-3)
    else: raw[:, 208:210] = h(rng.uniform(-1, 1, n) * 1e-3)
    return raw.reshape(-1)

ap = argparse.ArgumentParser(); ap.add_argument("--types", default="q4_K,q6_K"); ap.add_argument("--n", type=int, default=1)
ap.add_argument("--iters", type=int, default=200); ap.add_argument("--same", action="store_true", help="one weight buffer: L2-resident weights (126 MB L2)"); ap.add_argument("--ncu", action="store_true"); ap.add_argument("--shapes", default="")
a = ap.parse_args()
ctx = B.Context(0); rng = np.random.default_rng(0); out = []
shapes = [tuple(int(v) for v in s.split("x")) for s in a.shapes.split(",")] if a.shapes else SH
for tn in a.types.split(","):
    t = T[tn]
    for K, M in shapes:
        wb = K // B.BLCK[t] * B.TSIZE[t] * M
        nbuf = 1 if a.same else 2 if a.ncu else max(2, min(64, int(400e6 // wb) + 1))
        pool = blocks(rng, t, min(wb // B.TSIZE[t], 1 << 18))
        ws = []
        for _ in range(nbuf):
            w = ctx.empty(t, [K, M])
            for off in range(0, wb, pool.size): ctx.h2d(w.ptr + off, pool[: min(pool.size, wb - off)])
            ctx.repack(w); ws.append(w)
        x = ctx.upload(B.F32, [K, a.n], rng.standard_normal(K * a.n).astype(np.float32)); d = ctx.empty(B.F32, [M, a.n])
        for w in ws: ctx.mul_mat(w, x, d)
        ctx.synchronize()
        iters = 3 if a.ncu else a.iters
        ctx.profile(True)
        for i in range(iters): ctx.mul_mat(ws[i % nbuf], x, d)
        prof = ctx.profile_get(); ctx.profile(False)
        ctx.graph_begin()
        for i in range(max(nbuf, 8)): ctx.note_write(x); ctx.mul_mat(ws[i % nbuf], x, d)
        g = ctx.graph_end(); nrep = max(nbuf, 8); ctx.graph_launch(g); ctx.synchronize()
        e0, e1 = ctx.event(), ctx.event(); reps = max(1, iters // max(nbuf, 8))
        ctx.record(e0)
        for _ in range(reps): ctx.graph_launch(g)
        ctx.record(e1); ms = ctx.elapsed_ms(e0, e1) / (reps * nrep)
        kern = {k: v[0] / v[1] * 1e3 for k, v in prof.items()}
        gbs = wb / (ms * 1e-3) / 1e9
        if a.n > 8: print(f"      n={a.n}: {2.0 * M * K * a.n / (ms * 1e-3) / 1e12:8.1f} TFLOP/s", flush=True)
        print(f"{tn:5s} K={K:6d} M={M:6d} n={a.n}  {wb/1e6:8.2f} MB  graph {ms*1e3:8.2f} us/op {gbs:7.1f} GB/s ({gbs/6551:.2f})  kernel us: " +
              ", ".join(f"{k} {v:.2f}" for k, v in kern.items()), flush=True)
        out.append(dict(type=tn, K=K, M=M, n=a.n, us=ms * 1e3, gbs=gbs, kernel_us=kern))
        del ws
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gemv_bench.json"), "w"), indent=1)
