"""TEST INFRASTRUCTURE (oracle/): plain-Python restatement of the fork's layer distributor, llama_multi_gpu_manager::distribute_layers
(src/llama-multi-gpu.cpp:156-332) -- SURVEY.md 8(a) A9.  Pinned against the reference's own class (oracle/layer_assign_ref.cpp drives it
with fabricated devices; tests/golden/layer_assign_golden.json holds its answers).  Only tests/ may import this."""

ROUND_ROBIN, MEMORY_BALANCED, COMPUTE_BALANCED, PIPELINE, TENSOR_PARALLEL, HYBRID = range(6)      # llama-multi-gpu.h:84-91


def distribute_layers(strategy, tp_size, n_layers, layer_sizes, gpu_free):
    """-> list of (layer_id, gpu_id, memory_required, is_split, [split gpu ids])"""
    n_gpus = len(gpu_free)
    if n_gpus == 0:                                                        # :163-166
        return []
    size = lambda i: layer_sizes[i] if i < len(layer_sizes) else 0
    if strategy == ROUND_ROBIN:                                            # :194-214
        return [(i, i % n_gpus, size(i), 0, []) for i in range(n_layers)]
    if strategy == TENSOR_PARALLEL:                                        # :277-332
        tp = min(tp_size, n_gpus)
        if tp >= 2:
            return [(i, 0, size(i), 1, list(range(tp))) for i in range(n_layers)]
    # MEMORY_BALANCED, every other strategy (:170-185), and TENSOR_PARALLEL with fewer than two GPUs: greedy most-remaining (:216-275)
    allocated = [0] * n_gpus
    out = []
    for i in range(n_layers):
        best, best_left = 0, 0
        for g in range(n_gpus):
            left = gpu_free[g] - allocated[g] if gpu_free[g] > allocated[g] else 0
            if left > best_left:
                best, best_left = g, left
        out.append((i, best, size(i), 0, []))
        allocated[best] += size(i)
    return out
