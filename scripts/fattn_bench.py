#!/usr/bin/env python
"""Micro-benchmark of FLASH_ATTN_EXT through the C-ABI on KV-cache views (Llama-3-8B head geometry): prefill ubatches at a given
context depth (config 3 of BASELINE.json: pp2048 @ 8k -> ubatches of 512 at n_kv 8704..10240) and decode rows."""
import argparse, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B

ap = argparse.ArgumentParser(); ap.add_argument("--cases", default="512:512,512:2048,512:8704,512:10240,1:8192,1:32768")
ap.add_argument("--iters", type=int, default=20); ap.add_argument("--kv", default="f16", help="KV cache type: f16, q8_0, q4_0 (-ctk / -ctv)")
a = ap.parse_args()
TKV = {"f16": B.F16, "q8_0": B.Q8_0, "q4_0": B.Q4_0}[a.kv]
ctx = B.Context(0); rng = np.random.default_rng(0)
H, Hk, D = 32, 8, 128
out = []
for case in a.cases.split(","):
    N, n_kv = (int(v) for v in case.split(":"))
    kv_size = n_kv
    kvd = Hk * D
    q = ctx.upload(B.F32, [D * H, N], rng.standard_normal(D * H * N).astype(np.float32))
    esz = B.TSIZE[TKV]; bsz = B.BLCK[TKV]
    tok, head = kvd // bsz * esz, D // bsz * esz                # bytes of a token row (all kv heads) / of one head's row
    if TKV == B.F16:
        blk = (rng.standard_normal(kvd * 1024) * 0.5).astype(np.float16).view(np.uint8)
    else:                                                       # random quants, block scales around 0.01
        blk = rng.integers(0, 256, size=(kvd // 32 * 1024, esz), dtype=np.uint8)
        blk[:, 0:2] = np.asarray(rng.uniform(0.005, 0.02, len(blk)), dtype=np.float16).view(np.uint8).reshape(-1, 2)
        blk = blk.reshape(-1)
    # one K/V cache pair per "layer": cycling through them keeps the working set above the 126 MB L2, as in a real token
    nl = max(2, min(32, int(300e6 // (2 * tok * kv_size)) + 1))
    kvs = []
    for _ in range(nl):
        kc, vc = ctx.empty(TKV, [kvd, kv_size]), ctx.empty(TKV, [kvd, kv_size])
        for c in (kc, vc):
            for off in range(0, c.nbytes, blk.nbytes): ctx.h2d(c.ptr + off, blk[: min(blk.size, c.nbytes - off)])
        kvs.append((kc, vc, [c.view([D, n_kv, Hk, 1], [esz, tok, head, tok * kv_size]) for c in (kc, vc)]))
    nm = max(8, (N + 63) // 64 * 64)
    mask = np.zeros((nm, n_kv), dtype=np.float16)
    depth = n_kv - N
    for i in range(N): mask[i, depth + i + 1:] = -np.inf
    dm = ctx.upload(B.F16, [n_kv, nm], mask)
    qv = q.view([D, N, H, 1], [4, D * H * 4, D * 4, D * H * 4 * N])
    dst = ctx.empty(B.F32, [D, H, N, 1])
    sc = 1.0 / np.sqrt(D)
    for l in range(nl): ctx.flash_attn_ext(qv, kvs[l][2][0], kvs[l][2][1], dm, None, dst, sc)
    ctx.synchronize()
    ctx.graph_begin()
    for l in range(nl): ctx.flash_attn_ext(qv, kvs[l][2][0], kvs[l][2][1], dm, None, dst, sc)
    gr = ctx.graph_end(); ctx.graph_launch(gr); ctx.synchronize()
    e0, e1 = ctx.event(), ctx.event(); ctx.record(e0)
    for _ in range(a.iters): ctx.graph_launch(gr)
    ctx.record(e1); ms = ctx.elapsed_ms(e0, e1) / a.iters / nl
    flops = 2.0 * H * N * 2 * D * n_kv                       # op_flops of test-backend-ops (:6084-6089), mask not discounted
    live = 2.0 * H * 2 * D * sum(min(n_kv, depth + i + 1) for i in range(N))
    byts = 2.0 * n_kv * tok                                   # K and V rows as stored (quantized bytes for a quantized cache)
    print(f"kv={a.kv} N={N:5d} n_kv={n_kv:6d}  {ms*1e3:9.1f} us  {flops/ms/1e9:8.1f} TFLOP/s nominal ({live/ms/1e9:7.1f} unmasked)  KV {byts/ms/1e6:7.1f} GB/s", flush=True)
    out.append(dict(kv=a.kv, N=N, n_kv=n_kv, us=ms * 1e3, tflops_nominal=flops / ms / 1e9, tflops_unmasked=live / ms / 1e9, kv_gbs=byts / ms / 1e6))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"fattn_bench_{a.kv}.json"), "w"), indent=1)
