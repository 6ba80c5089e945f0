#!/usr/bin/env python
"""Micro-benchmark of the persistent decode chain (b200_decode_chain) on Llama-shaped layers: NL distinct layers (weights >> 126 MB L2)
in one CUDA graph, CUDA events on the ctx stream; prints us per layer chain and GB/s over the chain's weight bytes, next to the same
matmuls as separate launches of the flat API (b200_mul_mat / _multi / gate_up_swiglu; no glue ops there, so the chain does MORE work).
usage: chain_bench.py [--model 8b|70b] [--mix q4_k|q4_k_m] [--layers 8] [--reps 20] [--ncu] [--tag name]
Environment knobs of the kernel (measurement only): GGML_B200_CHAIN_DBG (1 no dot products, 2 no weight traffic), GGML_B200_CHAIN_SLOTS."""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B
from decode_chain import LayerChain

DIMS = {"8b": (4096, 14336, 1024), "70b": (8192, 28672, 1024), "tiny": (2048, 6144, 512)}


def blocks(rng, t, n):
    raw = rng.integers(0, 256, size=(n, B.TSIZE[t]), dtype=np.uint8)
    h = lambda v: np.asarray(v, dtype=np.float16).view(np.uint8).reshape(n, 2)
    if t == B.Q4_K: raw[:, 0:2] = h(rng.uniform(.1, 1, n) * 1e-3); raw[:, 2:4] = h(rng.uniform(.1, 1, n) * 1e-3)
    else: raw[:, 208:210] = h(rng.uniform(-1, 1, n) * 1e-3)
    return raw.reshape(-1)


ap = argparse.ArgumentParser()
ap.add_argument("--model", default="8b"); ap.add_argument("--mix", default="q4_k_m"); ap.add_argument("--layers", type=int, default=8)
ap.add_argument("--reps", type=int, default=20); ap.add_argument("--ncu", action="store_true"); ap.add_argument("--tag", default="")
ap.add_argument("--no-separate", action="store_true"); ap.add_argument("--trace", action="store_true")
ap.add_argument("--norm-outputs", action="store_true", help="also store the RMS_NORM / MUL node values (dead in the real graph): the kernel then takes the plain hand-over")
a = ap.parse_args()
n_embd, n_ff, n_kv = DIMS[a.model]
ctx = B.Context(0); rng = np.random.default_rng(0)
pool = {t: blocks(rng, t, 1 << 16) for t in (B.Q4_K, B.Q6_K)}


def weight(t, K, M):
    nb = K // 256 * M
    return (t, np.resize(pool[t].reshape(-1, B.TSIZE[t]), (nb, B.TSIZE[t])).reshape(-1))


layers = []
for il in range(2 if a.ncu else a.layers):
    more = a.mix == "q4_k_m" and il % 2 == 0                    # use_more_bits layers: attn_v and ffn_down are q6_K
    tv = td = B.Q6_K if more else B.Q4_K
    W = {"wo": weight(B.Q4_K, n_embd, n_embd), "gate": weight(B.Q4_K, n_embd, n_ff), "up": weight(B.Q4_K, n_embd, n_ff), "down": weight(td, n_ff, n_embd),
         "q": weight(B.Q4_K, n_embd, n_embd), "k": weight(B.Q4_K, n_embd, n_kv), "v": weight(tv, n_embd, n_kv)}
    lc = LayerChain(ctx, n_embd, n_ff, n_kv, W, np.ones(n_embd, np.float32), np.ones(n_embd, np.float32), norm_outputs=a.norm_outputs)
    lc.set_inputs(rng.standard_normal(n_embd).astype(np.float32), rng.standard_normal(n_embd).astype(np.float32))
    layers.append(lc)
ctx.synchronize()
nbytes = sum(l.weight_bytes() for l in layers)


def timed(fn, reps):
    fn(); ctx.synchronize()                                     # eager once (sizes scratch), then one graph of all layers
    ctx.graph_begin(); fn(); g = ctx.graph_end()
    for _ in range(3): ctx.graph_launch(g)
    ctx.synchronize()
    e0, e1 = ctx.event(), ctx.event()
    ctx.record(e0)
    for _ in range(reps): ctx.graph_launch(g)
    ctx.record(e1)
    return ctx.elapsed_ms(e0, e1) / reps


def run_chain():
    for l in layers: l.run()


def run_separate():
    for l in layers:
        for t in (l.attn_out, l.ffn_x, l.glu, l.attn_x): ctx.note_write(t)
        ctx.mul_mat(l.w["wo"], l.attn_out, l.wo_out)
        try: ctx.mul_mat_gate_up_swiglu(l.w["gate"], l.w["up"], l.ffn_x, l.gate, l.up, l.glu)
        except B.B200Error: ctx.mul_mat_multi([l.w["gate"], l.w["up"]], l.ffn_x, [l.gate, l.up])
        ctx.note_write(l.glu)
        ctx.mul_mat(l.w["down"], l.glu, l.down_out)
        ctx.mul_mat_multi([l.w["q"], l.w["k"], l.w["v"]], l.attn_x, [l.q, l.k, l.v])


reps = 2 if a.ncu else a.reps
res = {"model": a.model, "mix": a.mix, "layers": len(layers), "weight_MB_per_layer": nbytes / len(layers) / 1e6, "tag": a.tag,
       "env": {k: v for k, v in os.environ.items() if k.startswith("GGML_B200_")}}
ms = timed(run_chain, reps)
res["chain_us_per_layer"] = ms * 1e3 / len(layers); res["chain_gbs"] = nbytes / (ms * 1e-3) / 1e9
print(f"chain    {a.model} {a.mix}: {res['chain_us_per_layer']:7.2f} us/layer  {res['chain_gbs']:7.1f} GB/s  ({nbytes/len(layers)/1e6:.1f} MB/layer) {res['env']}", flush=True)
if a.trace:
    # device-side timeline (b200_debug_trace): per phase, medians over the CTAs of the last layer's launch, ns
    CAP = 148 * len(layers) + 8
    buf = ctx.malloc((1 + CAP) * 256 + 16384); ctx.memset(buf, 0, (1 + CAP) * 256 + 16384)
    B.lib().b200_debug_trace(ctx.h, buf, CAP)
    run_chain(); ctx.synchronize()
    B.lib().b200_debug_trace(ctx.h, None, 0)
    raw = np.empty((1 + CAP) * 32 + 2048, np.uint64); ctx.d2h(raw, buf); ctx.synchronize()
    stage_t = raw[(1 + CAP) * 32:].astype(np.int64).reshape(2, 1024)
    n = int(raw[0]); recs = raw[32:32 * (1 + n)].reshape(n, 32)
    per_launch = recs.reshape(len(layers), -1, 32) if n == 148 * len(layers) else recs[None]
    names = ["wait input", "build image", "stages", "publish"]
    for li in (0, len(per_launch) - 1):
        L = per_launch[li]; st = L[:, 2:32].astype(np.int64).reshape(len(L), 6, 5)
        t0 = st[:, 0, 0].min()
        print(f"trace launch {li}: {len(L)} CTAs; first CTA start 0, last CTA start {int(st[:, 0, 0].max() - t0)} ns")
        for p in range(len(layers[0].phases)):
            d = [np.median(st[:, p, j + 1] - st[:, p, j]) for j in range(4)]
            print(f"  phase {p}: enters at {int(np.median(st[:, p, 0]) - t0):6d} ns (spread {int(st[:, p, 0].max() - st[:, p, 0].min())})  " +
                  "  ".join(f"{nm} {int(v):5d}" for nm, v in zip(names, d)) + f"   max stages {int((st[:, p, 3] - st[:, p, 2]).max())}", flush=True)
        nph = len(layers[0].phases)
        ex = L[:, 26:29].astype(np.int64)
        if (ex > 0).all():
            print("  phase 1 image, medians after 'input complete': loads landed %d, norm sum known %d, quantized %d, (barrier) %d ns" %
                  tuple(int(np.median(v - st[:, 1, 1])) for v in (ex[:, 0], ex[:, 1], ex[:, 2], st[:, 1, 2])))
        print(f"  last CTA ends at {int(st[:, nph - 1, 4].max() - t0)} ns")
        if li == len(per_launch) - 1:
            ns = int((stage_t[0] > 0).sum())
            c0 = L[L[:, 1] % (1 << 32) == 0][0, 2:32].astype(np.int64).reshape(6, 5)
            print(f"  CTA 0 of this launch: {ns} stages; per stage: issued / seen by its consumer (ns after CTA 0 start {int(c0[0,0]-t0)}); phases enter at", [int(c0[q, 0] - c0[0, 0]) for q in range(nph)], "stage loops start at", [int(c0[q, 2] - c0[0, 0]) for q in range(nph)])
            for i0 in range(0, ns, 12):
                print("   ", "  ".join(f"{i}:{int(stage_t[1, i] - c0[0, 0])}/{int(stage_t[0, i] - c0[0, 0])}" for i in range(i0, min(ns, i0 + 12))))
if not a.no_separate and not a.ncu:
    ms = timed(run_separate, reps)
    res["separate_us_per_layer"] = ms * 1e3 / len(layers); res["separate_gbs"] = nbytes / (ms * 1e-3) / 1e9
    print(f"separate {a.model} {a.mix}: {res['separate_us_per_layer']:7.2f} us/layer  {res['separate_gbs']:7.1f} GB/s  (matmuls only, no glue)", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "chain_bench.jsonl"), "a") as f:
    f.write(json.dumps(res) + "\n")
