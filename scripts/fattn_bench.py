#!/usr/bin/env python
"""Micro-benchmark of FLASH_ATTN_EXT through the C-ABI on KV-cache views (Llama-3-8B head geometry): prefill ubatches at a given
context depth (config 3 of BASELINE.json: pp2048 @ 8k -> ubatches of 512 at n_kv 8704..10240) and decode rows."""
import argparse, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B

ap = argparse.ArgumentParser(); ap.add_argument("--cases", default="512:512,512:2048,512:8704,512:10240,1:8192,1:32768")
ap.add_argument("--iters", type=int, default=20); a = ap.parse_args()
ctx = B.Context(0); rng = np.random.default_rng(0)
H, Hk, D = 32, 8, 128
out = []
for case in a.cases.split(","):
    N, n_kv = (int(v) for v in case.split(":"))
    kv_size = n_kv
    kvd = Hk * D
    q = ctx.upload(B.F32, [D * H, N], rng.standard_normal(D * H * N).astype(np.float32))
    blk = (rng.standard_normal(kvd * 1024) * 0.5).astype(np.float16)
    tok, head = kvd * 2, D * 2
    # one K/V cache pair per "layer": cycling through them keeps the working set above the 126 MB L2, as in a real token
    nl = max(2, min(32, int(300e6 // (2 * kvd * kv_size * 2)) + 1))
    kvs = []
    for _ in range(nl):
        kc, vc = ctx.empty(B.F16, [kvd, kv_size]), ctx.empty(B.F16, [kvd, kv_size])
        for c in (kc, vc):
            for off in range(0, c.nbytes, blk.nbytes): ctx.h2d(c.ptr + off, blk[: min(blk.size, (c.nbytes - off) // 2)])
        kvs.append((kc, vc, [c.view([D, n_kv, Hk, 1], [2, tok, head, tok * kv_size]) for c in (kc, vc)]))
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
    byts = 2.0 * n_kv * kvd * 2
    print(f"N={N:5d} n_kv={n_kv:6d}  {ms*1e3:9.1f} us  {flops/ms/1e9:8.1f} TFLOP/s nominal ({live/ms/1e9:7.1f} unmasked)  KV {byts/ms/1e6:7.1f} GB/s", flush=True)
    out.append(dict(N=N, n_kv=n_kv, us=ms * 1e3, tflops_nominal=flops / ms / 1e9, tflops_unmasked=live / ms / 1e9, kv_gbs=byts / ms / 1e6))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "fattn_bench.json"), "w"), indent=1)
