#!/usr/bin/env python
"""Load-path benchmark (b200_load_file): a Llama-3-8B Q4_K_M sized set of weight tensors (random GGUF blocks) written to a
file, then loaded file -> pinned ring -> device with the q6_K tensors re-laid out on the way, and the rank-by-rank sharded read.  Run it a second time with GGML_B200_LOAD_SERIAL=1 for the shape of the reference's loop
(llama-model-loader.cpp:1136-1186: one thread alternating between read() into a 1 MiB pinned buffer and the async copy, 4 buffers, one
stream; the entry named pipelined_1MiB_x4 is then exactly that).  The file is read from the page cache (written just before), so the numbers are the
pipeline's, not a disk's.  usage: load_bench.py [--gb 2.0] [--chunk-mb 4] [--buffers 4]"""
import argparse, json, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
import ggml_b200 as B

ap = argparse.ArgumentParser()
ap.add_argument("--gb", type=float, default=2.0); ap.add_argument("--chunk-mb", type=float, default=4.0); ap.add_argument("--buffers", type=int, default=8)
ap.add_argument("--no-sharded", action="store_true")
a = ap.parse_args()
ctx = B.Context(0); rng = np.random.default_rng(0)
specs = []
total = 0
while total < a.gb * 1e9:                                    # per layer: wq wk wo gate up (q4_K), wv down (q6_K on every other layer)
    more = (len(specs) // 7) % 2 == 0
    for t, K, M in ((B.Q4_K, 4096, 4096), (B.Q4_K, 4096, 1024), (B.Q6_K if more else B.Q4_K, 4096, 1024), (B.Q4_K, 4096, 4096),
                    (B.Q4_K, 4096, 14336), (B.Q4_K, 4096, 14336), (B.Q6_K if more else B.Q4_K, 14336, 4096)):
        specs.append((t, K, M)); total += K // 256 * B.TSIZE[t] * M
path = os.path.join(tempfile.gettempdir(), "b200_load_bench.bin")
pool = rng.integers(0, 256, 64 << 20, dtype=np.uint8)
offs = []
with open(path, "wb") as f:
    for t, K, M in specs:
        n = K // 256 * B.TSIZE[t] * M
        offs.append((f.tell(), n)); f.write(np.resize(pool, n).tobytes()); f.write(b"\0" * 32)
tens = [ctx.empty(t, [K, M]) for t, K, M in specs]
rp = [int(bool(B.lib().b200_repack_supported(t, K))) for t, K, M in specs]
res = {"bytes": total, "tensors": len(specs), "repacked_bytes": sum(n for (o, n), r in zip(offs, rp) if r)}
for name, chunk, nb in (("pipelined", int(a.chunk_mb * (1 << 20)), a.buffers), ("pipelined_1MiB_x4", 1 << 20, 4)):
    for t in tens: t.flags = 0
    best = None
    for rep in range(3):
        for t in tens: t.flags = 0
        s, rs, nbytes, chunks = ctx.load_file(path, [(o, n, d, 0, r) for (o, n), d, r in zip(offs, tens, rp)], chunk_bytes=chunk, n_buffers=nb)
        best = s if best is None else min(best, s)
    res[name] = {"seconds": best, "GB/s": total / best / 1e9, "read_seconds": rs, "chunk_bytes": chunk, "buffers": nb}
print(json.dumps(res)); sys.stdout.flush()
# sharded: 8 ranks' worth of row slices, one after the other on this GPU (each rank would run its own in parallel on its own GPU)
world = 8; per_rank = []
for rank in range(0 if a.no_sharded else world):
    items = []
    for (t, K, M), (o, n) in zip(specs, offs):
        lo, hi = B.split_row_range(M, world, rank, 64)
        if hi > lo:
            it = B.LoadItem(); B.lib().b200_load_row_slice(t, K, o, lo, hi, it)
            items.append((it.file_offset, it.nbytes, ctx.empty(t, [K, hi - lo]), 0, 0))
    s, rs, nbytes, chunks = ctx.load_file(path, items, chunk_bytes=int(a.chunk_mb * (1 << 20)), n_buffers=a.buffers)
    per_rank.append({"rank": rank, "bytes": nbytes, "seconds": s})
    for it in items: it[2].free()
if per_rank: res["row_sharded_8_ranks"] = {"bytes_per_rank": [p["bytes"] for p in per_rank], "seconds_per_rank_max": max(p["seconds"] for p in per_rank),
                              "whole_file_fraction_read_per_rank": max(p["bytes"] for p in per_rank) / total}
os.remove(path)
print(json.dumps(res))
