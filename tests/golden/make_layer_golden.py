#!/usr/bin/env python
"""Generates tests/golden/layer_assign_golden.json from the REFERENCE's own llama_multi_gpu_manager (oracle/_ref/avx2/layer_assign_ref,
built by `make -C oracle llama` from /root/reference/src/llama-multi-gpu.cpp + oracle/layer_assign_ref.cpp).  Run here, where the
reference exists; the fixture travels.  Sizes are MiB on the command line, bytes in the fixture."""
import json, os, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
EXE = os.path.join(ROOT, "oracle", "_ref", "avx2", "layer_assign_ref")
SCENARIOS = [  # strategy, tp, n_layers, gpu free MiB, layer MiB
    (0, 1, 10, [1000, 1000, 1000], [50] * 10),
    (0, 1, 5, [10], [1, 2, 3]),
    (1, 1, 6, [100, 300, 200], [50, 50, 50, 50]),
    (1, 1, 32, [81920] * 8, [139.5] * 32),                      # Llama-3-8B Q4_K_M layers over 8 equal GPUs
    (1, 1, 80, [40000, 80000, 80000, 20000], [480.25] * 80),      # 70B-sized layers, unequal free memory
    (1, 1, 12, [100, 100], [90, 90, 90, 5, 5, 5, 5, 5, 5, 5, 5, 5]),  # runs out: remaining 0 everywhere -> GPU 0
    (1, 1, 4, [0, 0, 0], [10, 10, 10, 10]),
    (1, 1, 0, [100, 100], []),
    (4, 2, 8, [1000, 1000, 1000, 1000], [100] * 8),
    (4, 8, 5, [1000, 1000, 1000], [33, 34, 35]),                  # tp clamps to the GPU count
    (4, 1, 6, [100, 300, 200], [50] * 6),                         # tp < 2: falls back to memory-balanced
    (4, 4, 3, [500], [10, 20, 30]),                               # one GPU: falls back too
    (2, 1, 6, [100, 300, 200], [50] * 6), (3, 2, 6, [300, 100, 200], [70] * 6), (5, 2, 6, [100, 200, 300], [60] * 6),   # other strategies: default branch
    (1, 1, 3, [], [1, 2, 3]),                                     # no GPUs: empty plan
]
out = []
for st, tp, nl, gpus, sizes in SCENARIOS:
    args = [EXE, str(st), str(tp), str(nl), ",".join(map(str, gpus)) or "-", ",".join(map(str, sizes)) or "-"]
    txt = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, check=True).stdout.splitlines()
    n = int(txt[0].split("=")[1])
    rows = [[int(x) for x in l.split()] for l in txt[1:1 + n]]
    out.append({"strategy": st, "tp": tp, "n_layers": nl, "gpu_free": [int(g * 1024 * 1024) for g in gpus], "layer_sizes": [int(s * 1024 * 1024) for s in sizes],
                "assignments": [{"layer_id": r[0], "gpu_id": r[1], "memory_required": r[2], "is_split": r[3], "split_gpu_ids": r[4:]} for r in rows]})
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "layer_assign_golden.json"), "w"), indent=0)
print(f"{len(out)} scenarios, {sum(len(o['assignments']) for o in out)} assignments")
