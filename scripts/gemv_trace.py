#!/usr/bin/env python
"""Device-side timeline of the decode MUL_MAT chain (b200_debug_trace): every CTA of the streaming GEMV kernels stamps
%globaltimer / clock64 at CTA start, prologue prefetch issued, griddepcontrol.wait returned, activations quantized, first stage
landed, CTA end.  Replays the MUL_MAT-only graph of a few Llama-3-8B Q4_K_M layers + the output head and prints, per launch,
where its time goes and how it overlaps its neighbours.  usage: gemv_trace.py [--layers 4] [--tag name]"""
import argparse, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B
import bench

ap = argparse.ArgumentParser(); ap.add_argument("--layers", type=int, default=4); ap.add_argument("--tag", default="trace")
ap.add_argument("--head", type=int, default=1); ap.add_argument("--show", type=int, default=14)
a = ap.parse_args()
ctx = B.Context(0)
cfg = dict(bench.LLAMA3_8B, n_layer=a.layers)
if not a.head: cfg["n_vocab"] = 4096
# the Q4_K_M type mix of the first layers of a 32-layer model (use_more_bits sees the real depth)
types32, _ = bench.q4_k_m_layer_types(bench.LLAMA3_8B)
orig = bench.q4_k_m_layer_types
bench.q4_k_m_layer_types = lambda c, q="q4_k_m": ((types32[2:2 + a.layers] if a.layers < 30 else types32[:a.layers]), B.Q6_K)      # layers 2.. : q4_K, q4_K, q6_K-promoted, ...
m = bench.HotPathModel(B, ctx, cfg, 1, 0)
bench.q4_k_m_layer_types = orig
step = lambda: m.step(attention=False)
for _ in range(3): step()
ctx.synchronize()
# untraced timing of the same graph first
ctx.graph_begin(); step(); g0 = ctx.graph_end()
for _ in range(3): ctx.graph_launch(g0)
ctx.synchronize()
e0, e1 = ctx.event(), ctx.event()
ctx.record(e0)
for _ in range(20): ctx.graph_launch(g0)
ctx.record(e1); ctx.synchronize()
us_plain = ctx.elapsed_ms(e0, e1) / 20 * 1e3
wb = m.weight_bytes
print(f"## {a.tag}: MUL_MAT graph of {a.layers} layers{' + head' if a.head else ''}: {us_plain:.1f} us per replay, {wb / 1e6:.1f} MB -> {wb / us_plain / 1e3:.0f} GB/s = {wb / us_plain / 1e3 / 6551:.3f} of HBM peak; env "
      + " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("GGML_B200")), flush=True)
CAP = 1 << 16
buf = ctx.malloc((1 + CAP) * 256)
ctx.memset(buf, 0, (1 + CAP) * 256)
B.lib().b200_debug_trace(ctx.h, buf, CAP)
ctx.graph_begin(); step(); g = ctx.graph_end()
B.lib().b200_debug_trace(ctx.h, None, 0)
for _ in range(3): ctx.graph_launch(g)
ctx.synchronize()
e0, e1 = ctx.event(), ctx.event()
ctx.memset(buf, 0, 256); ctx.synchronize()
ctx.record(e0); ctx.graph_launch(g); ctx.record(e1); ctx.synchronize()
ms = ctx.elapsed_ms(e0, e1)
raw = np.zeros((1 + CAP) * 32, dtype=np.uint64)
ctx.d2h(raw, buf)
n = int(raw[0]); rec = raw[32:32 * (1 + n)].reshape(n, 32)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
if os.environ.get("TRACE_SAVE_NPY"): np.save(os.path.join(ROOT, "gpurun_out", f"{a.tag}.npy"), rec)
key = rec[:, 0]; smid = (rec[:, 1] >> np.uint64(32)).astype(np.int64)
gt = rec[:, 2::2].astype(np.int64)[:, :10]; ck = rec[:, 3::2].astype(np.int64)[:, :10]
t0 = gt[:, 0].min()
gt = (gt - t0) / 1e3                                   # us
# launches = runs of equal key in slot order
bounds = [0] + [i for i in range(1, n) if key[i] != key[i - 1]] + [n]
dgt = np.diff(np.unique(rec[:, 2::2].astype(np.int64)[:, :10].ravel())); dgt = dgt[dgt > 0]
print(f"# {a.tag}: graph replay {ms * 1e3:.1f} us, {n} CTA records, {len(bounds) - 1} launches; globaltimer step ~{dgt.min() if dgt.size else -1} ns; "
      f"env " + " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("GGML_B200")))
print("# per launch, us since the first CTA start.  start/wait/quant/first/end = min..max over CTAs of: CTA start, wait returned, quantized, first stage landed, CTA end")
print("# dq = median cycles wait->quantized (clock64), gap = this launch's first wait-return minus the previous launch's last CTA end")
print(f"{'#':>3} {'M':>6} {'K':>5} {'kern':>4} {'ctas':>4} {'MB':>6} | {'start':>13} | {'wait':>13} | {'quant':>13} | {'first':>13} | {'end':>21} | {'dur':>5} {'gap':>5} {'dq_cyc':>6} {'GB/s':>6}")
prev_end = None; rows = []
for li in range(len(bounds) - 1):
    s, e = bounds[li], bounds[li + 1]
    M = int(key[s] >> np.uint64(32)); K = int(key[s] & np.uint64(0x3fffffff)); mma = bool(int(key[s]) & (1 << 30))
    G = gt[s:e]; Cc = ck[s:e]
    f = lambda c: f"{G[:, c].min():6.1f}..{G[:, c].max():6.1f}"
    end = G[:, 5]
    dq = int(np.median(Cc[:, 3] - Cc[:, 2]))
    dur = end.max() - G[:, 2].min()
    gap = (G[:, 2].min() - prev_end) if prev_end is not None else 0.0
    mb = M * (K // 256) * 144 / 1e6                      # q4_K bytes (q6_K launches: x 210/144)
    rows.append(dict(launch=li, M=M, K=K, mma=mma, ctas=e - s, start=[G[:, 0].min(), G[:, 0].max()], wait=[G[:, 2].min(), G[:, 2].max()],
                     quant=[G[:, 3].min(), G[:, 3].max()], first=[G[:, 4].min(), G[:, 4].max()], end=[end.min(), float(np.median(end)), end.max()], dq_cycles=dq))
    if li < a.show or li >= len(bounds) - 3: print(f"{li:3d} {M:6d} {K:5d} {'mma' if mma else 'dp4a':>4} {e - s:4d} {mb:6.1f} | {f(0)} | {f(2)} | {f(3)} | {f(4)} | {end.min():6.1f} {np.median(end):6.1f} {end.max():6.1f} | {dur:5.1f} {gap:5.1f} {dq:6d} {mb / dur * 1e3 if dur > 0 else 0:6.0f}")
    if li < a.show or li >= len(bounds) - 3:
        md = lambda x, y: int(np.median((Cc[:, x] - Cc[:, y])[(Cc[:, x] > 0) & (Cc[:, y] > 0)])) if ((Cc[:, x] > 0) & (Cc[:, y] > 0)).any() else -1
        print(f"      cycles (median over CTAs, thread 0): wait->x landed {md(8, 2)}, ->quantized+sync {md(3, 8)}, ->first item done {md(9, 3)}, ->warp 0 items done {md(6, 9)}, "
              + (f"->all warps done {md(7, 6)}, ->reduced+stored {md(5, 7)}" if mma else f"->all warps done+stored {md(5, 6)}"))
    prev_end = end.max()
json.dump(rows, open(os.path.join(ROOT, "gpurun_out", f"{a.tag}.json"), "w"))
