"""SURVEY.md 8(a) A9, the fork's layer distributor (src/llama-multi-gpu.cpp:156-332): host-only logic, so everything here runs without a
GPU.  tests/golden/layer_assign_golden.json holds the answers of the REFERENCE's own llama_multi_gpu_manager (driven with fabricated
devices by oracle/layer_assign_ref.cpp, fixture written by tests/golden/make_layer_golden.py); the Python restatement
(oracle/layer_assign.py) and the product's b200_distribute_layers (C-ABI, csrc/b200_layers.cpp) must both reproduce them exactly."""
import json
import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "layer_assign_golden.json")))


def want(sc):
    return [(a["layer_id"], a["gpu_id"], a["memory_required"], a["is_split"], a["split_gpu_ids"]) for a in sc["assignments"]]


@pytest.mark.parametrize("i", range(len(GOLD)))
def test_restatement_matches_reference(i):
    import layer_assign as L
    sc = GOLD[i]
    assert L.distribute_layers(sc["strategy"], sc["tp"], sc["n_layers"], sc["layer_sizes"], sc["gpu_free"]) == want(sc)


@pytest.mark.parametrize("i", range(len(GOLD)))
def test_c_abi_matches_reference(i):
    import ggml_b200 as B
    sc = GOLD[i]
    got, charged = B.distribute_layers(sc["strategy"], sc["tp"], sc["n_layers"], sc["layer_sizes"], sc["gpu_free"])
    assert got == want(sc)
    # bytes charged per GPU: whole layers, or size / tp on every GPU of a split layer (llama-multi-gpu.cpp:245,310-321)
    exp = [0] * len(sc["gpu_free"])
    for lid, gpu, mem, split, ids in want(sc):
        if split:
            for g in ids:
                exp[g] += mem // len(ids)
        else:
            exp[gpu] += mem
    assert charged == exp


def test_c_abi_random_sweep_against_restatement():
    import random
    import ggml_b200 as B
    import layer_assign as L
    rnd = random.Random(7)
    for _ in range(300):
        n_gpus = rnd.randint(0, 8); nl = rnd.randint(0, 90)
        gpu_free = [rnd.choice([0, rnd.randint(1, 1 << 36)]) for _ in range(n_gpus)]
        sizes = [rnd.randint(0, 1 << 30) for _ in range(rnd.randint(0, nl + 3))]
        st, tp = rnd.randint(0, 5), rnd.randint(0, 9)
        got, _ = B.distribute_layers(st, tp, nl, sizes, gpu_free)
        assert got == L.distribute_layers(st, tp, nl, sizes, gpu_free), (st, tp, nl, gpu_free)


def test_c_abi_rejects_bad_arguments():
    import ggml_b200 as B
    assert B.lib().b200_distribute_layers(1, 1, 4, None, 4, 2, None, None, None) < 0
    assert B.lib().b200_distribute_layers(1, 1, -1, None, 0, 0, None, None, None) < 0


def test_plan_to_tensor_split_feeds_the_row_split():
    """a plan becomes llama.cpp's tensor_split; for a tensor-parallel plan that is the even row split b200_split_row_range applies"""
    import ggml_b200 as B
    import ctypes as C
    plan, _ = B.distribute_layers(4, 4, 80, [480 << 20] * 80, [80 << 30] * 8)           # TENSOR_PARALLEL over the first 4 of 8 GPUs
    ts = B.plan_tensor_split(plan, 8)
    assert ts == [0.25, 0.25, 0.25, 0.25, 0.0, 0.0, 0.0, 0.0]
    starts = [sum(ts[:g]) for g in range(8)]                    # the cumulative form ggml_backend_split_buffer_type normalises proportions to (ggml-cuda.cu:850-870)
    arr = (C.c_float * 8)(*starts)
    lo, hi = C.c_int64(), C.c_int64()
    covered = []
    for r in range(8):
        assert B.lib().b200_split_row_range(8192, arr, 8, r, 64, C.byref(lo), C.byref(hi)) == 0
        covered.append((lo.value, hi.value))
    assert covered[:4] == [(0, 2048), (2048, 4096), (4096, 6144), (6144, 8192)] and all(a == b for a, b in covered[4:])
    plan, _ = B.distribute_layers(1, 1, 32, [140 << 20] * 32, [3 << 30, 2 << 30])         # memory-balanced, unequal GPUs: the larger one fills first, then they alternate
    ts = B.plan_tensor_split(plan, 2)
    assert abs(sum(ts) - 1.0) < 1e-6 and ts[0] > ts[1] > 0
    plan, _ = B.distribute_layers(0, 1, 10, [], [1, 1, 1])                                # round-robin: 4 / 3 / 3 layers
    assert [round(t * 10) for t in B.plan_tensor_split(plan, 3)] == [4, 3, 3]
    assert B.plan_tensor_split([], 2) == [0.0, 0.0]
