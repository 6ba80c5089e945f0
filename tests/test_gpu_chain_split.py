"""GPU tests (-m gpu) for b200_decode_chain_split: the persistent decode chain row-split over the GPUs of one node, one process per GPU,
the all-gather folded into the hand-over between the phases (row owners publish q8_1 records as (word, tag) pairs into every rank's area
over NVLink peer memory; b200_chain.cu).  The decode form of the reference's split-mode row (ggml/src/ggml-cuda/ggml-cuda.cu:1451-1769).

  * one GPU: the row-split form with world = 1 must give, word for word, what the default chain gives (which tests/test_gpu_chain.py
    checks against the oracle): same row owners, same arithmetic, only the transport of the records differs;
  * two GPUs (skipped on a one-GPU box): two processes, each with its rows of every weight; every rank must end with the full q / k / v
    and its own rows of every other node, word-equal with the unsplit chain run on the same GPU from the same weights, launch after launch.
"""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

COMMON = r'''
import os, sys
import numpy as np
ROOT = sys.argv[1]
sys.path[:0] = [os.path.join(ROOT, "super-llama.cpp_b200"), os.path.join(ROOT, "oracle")]
import oracle as O, ggml_b200 as B
from decode_chain import LayerChain

def layer_weights(rng, dims, t_v, t_down):
    n_embd, n_ff, n_kv = dims
    types = {"wo": O.Q4_K, "gate": O.Q4_K, "up": O.Q4_K, "down": t_down, "q": O.Q4_K, "k": O.Q4_K, "v": t_v}
    shape = {"wo": (n_embd, n_embd), "gate": (n_embd, n_ff), "up": (n_embd, n_ff), "down": (n_ff, n_embd), "q": (n_embd, n_embd), "k": (n_embd, n_kv), "v": (n_embd, n_kv)}
    sc = 0.004 * min(1.0, (2048.0 / n_embd)) ** 0.5
    W = {n: O.random_blocks(rng, types[n], *shape[n], scale=sc) for n in types}
    nf = rng.uniform(0.5, 1.5, n_embd).astype(np.float32); na = rng.uniform(0.5, 1.5, n_embd).astype(np.float32)
    return types, shape, W, nf, na
'''

WORKER = COMMON + r'''
import torch.distributed as dist
from row_split import RowSplit, PeerExchange
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
ctx = B.Context(rank)
peer = PeerExchange(ctx, dist, world, rank, 24 << 20)
rs = RowSplit(dist, world, rank, 32)
ok = True
for dims, t_v, t_down in (((2048, 6144, 512), O.Q4_K, O.Q6_K), ((4096, 14336, 1024), O.Q6_K, O.Q4_K), ((8192, 28672, 1024), O.Q4_K, O.Q6_K)):
    rng = np.random.default_rng(dims[0] + t_down)
    types, shape, W, nf, na = layer_weights(rng, dims, t_v, t_down)
    ll_bytes = int(B.lib().b200_decode_chain_split_bytes(dims[0] + 2 * dims[2]))
    ll = peer.alloc(B.F32, [ll_bytes // 4])
    ranges, mine = {}, {}
    for n in types:
        K, M = shape[n]
        blk, lo, hi = rs.shard_weight(W[n], W[n].size // M, M)
        ranges[n] = (lo, hi); mine[n] = (types[n], blk)
    sp = dict(world=world, rank=rank, ll_ptrs=[int(peer.ex.base[r]) + (ll.ptr - peer.base) for r in range(world)], ll_bytes=ll_bytes, ranges=ranges)
    part = LayerChain(ctx, *dims, mine, nf, na, norm_outputs=False, split=sp)
    assert part.supported(), B.lib().b200_last_error().decode()
    # the unsplit chain in the same form, on this GPU, from the same weights
    l1 = ctx.empty(B.F32, [ll_bytes // 4]); ctx.memset(l1.ptr, 0, ll_bytes)
    whole = LayerChain(ctx, *dims, {n: (types[n], W[n]) for n in types}, nf, na, norm_outputs=False,
                       split=dict(world=1, rank=0, ll_ptrs=[l1.ptr], ll_bytes=ll_bytes, ranges={n: (0, shape[n][1]) for n in types}))
    assert whole.supported(), B.lib().b200_last_error().decode()
    dist.barrier()
    own = {"wo_out": "wo", "ffn_inp": "wo", "gate": "gate", "up": "up", "glu": "gate", "down_out": "down", "l_out": "down"}
    for rep in range(4):                                  # the tags count on, the areas are reused
        attn_out = rng.uniform(-1, 1, dims[0]).astype(np.float32); inp_sa = rng.uniform(-2, 2, dims[0]).astype(np.float32)
        for lc in (part, whole):
            lc.set_inputs(attn_out, inp_sa); lc.clear_outputs()
        part.run(); whole.run()
        if rank == 0 and rep == 1:
            whole.run()                                   # a launch only ONE rank makes, through other areas: the ranks' tags must not drift apart
        err = ctx.decode_chain_error()
        a, b = part.outputs(), whole.outputs()
        good = err == 0
        for n in ("q", "k", "v"):
            good = good and np.array_equal(a[n], b[n]) and bool(np.isfinite(a[n]).all())
        for n, wn in own.items():
            lo, hi = ranges[wn]
            good = good and np.array_equal(a[n][lo:hi], b[n][lo:hi])
        if not good:
            print(f"rank {rank} dims {dims} rep {rep}: error code {err:#x}; " + ", ".join(f"{n}: {int((a[n] != b[n]).sum())} differ" for n in ("q", "k", "v")), flush=True)
        ok = ok and good
        sums = [None] * world; dist.all_gather_object(sums, b"".join(a[n].tobytes() for n in ("q", "k", "v")))
        ok = ok and all(x == sums[0] for x in sums)
flag = [None] * world; dist.all_gather_object(flag, bool(ok))
dist.destroy_process_group()
sys.exit(0 if all(flag) else 1)
'''


@pytest.fixture(scope="module")
def B():
    import ggml_b200
    return ggml_b200


@pytest.fixture(scope="module")
def ctx(B):
    c = B.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("dims,t_v,t_down", [((2048, 6144, 512), O.Q4_K, O.Q4_K), ((4096, 14336, 1024), O.Q6_K, O.Q6_K), ((8192, 28672, 1024), O.Q4_K, O.Q4_K)])
def test_chain_split_form_equals_default_chain_on_one_gpu(ctx, B, dims, t_v, t_down):
    """world = 1: hand-over records validated by tags instead of an arrival counter; everything else is the default chain"""
    from decode_chain import LayerChain
    scope = {}
    exec(COMMON.replace("ROOT = sys.argv[1]", f"ROOT = {ROOT!r}"), scope)
    rng = np.random.default_rng(77 + dims[0])
    types, shape, W, nf, na = scope["layer_weights"](rng, dims, t_v, t_down)
    weights = {n: (types[n], W[n]) for n in types}
    ll_bytes = int(B.lib().b200_decode_chain_split_bytes(dims[0] + 2 * dims[2]))
    ll = ctx.empty(B.F32, [ll_bytes // 4]); ctx.memset(ll.ptr, 0, ll_bytes)
    a = LayerChain(ctx, *dims, weights, nf, na, norm_outputs=False,
                   split=dict(world=1, rank=0, ll_ptrs=[ll.ptr], ll_bytes=ll_bytes, ranges={n: (0, shape[n][1]) for n in types}))
    b = LayerChain(ctx, *dims, weights, nf, na, norm_outputs=False)
    assert a.supported() and b.supported(), B.lib().b200_last_error().decode()
    for rep in range(3):
        attn_out = rng.uniform(-1, 1, dims[0]).astype(np.float32); inp_sa = rng.uniform(-2, 2, dims[0]).astype(np.float32)
        for lc in (a, b):
            lc.set_inputs(attn_out, inp_sa); lc.clear_outputs(); lc.run()
        assert ctx.decode_chain_error() == 0
        ga, gb = a.outputs(), b.outputs()
        for n in ga:
            assert np.isfinite(ga[n]).all(), (rep, n)
            assert np.array_equal(ga[n], gb[n]), f"rep {rep} {n}: {int((ga[n] != gb[n]).sum())} words differ, nmse {O.nmse(gb[n], ga[n])}"


def test_chain_split_declines_what_it_cannot_take(ctx, B):
    """rows that do not start on a q8_1 block, a hand-over area that is too small, a ROPE tail: errors, not wrong results"""
    from decode_chain import LayerChain
    scope = {}
    exec(COMMON.replace("ROOT = sys.argv[1]", f"ROOT = {ROOT!r}"), scope)
    dims = (2048, 6144, 512)
    rng = np.random.default_rng(5)
    types, shape, W, nf, na = scope["layer_weights"](rng, dims, O.Q4_K, O.Q4_K)
    weights = {n: (types[n], W[n]) for n in types}
    ll_bytes = int(B.lib().b200_decode_chain_split_bytes(dims[0] + 2 * dims[2]))
    ll = ctx.empty(B.F32, [ll_bytes // 4]); ctx.memset(ll.ptr, 0, ll_bytes)
    full = {n: (0, shape[n][1]) for n in types}
    small = LayerChain(ctx, *dims, weights, nf, na, norm_outputs=False, split=dict(world=1, rank=0, ll_ptrs=[ll.ptr], ll_bytes=4096, ranges=full))
    assert not small.supported()
    nolist = LayerChain(ctx, *dims, weights, nf, na, norm_outputs=False, split=dict(world=2, rank=0, ll_ptrs=[ll.ptr], ll_bytes=ll_bytes, ranges=full))
    assert not nolist.supported()                              # rank 1's area is missing
    # this rank's rows of wo start at row 16: not a whole q8_1 block
    odd = dict(full); odd["wo"] = (16, shape["wo"][1])
    w2 = dict(weights); w2["wo"] = (types["wo"], np.ascontiguousarray(W["wo"].reshape(shape["wo"][1], -1)[16:]).reshape(-1))
    bad = LayerChain(ctx, *dims, w2, nf, na, norm_outputs=False, split=dict(world=1, rank=0, ll_ptrs=[ll.ptr], ll_bytes=ll_bytes, ranges=odd))
    assert not bad.supported()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def test_chain_split_two_gpus(tmp_path, B):
    if B.lib().b200_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"; script.write_text(WORKER)
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
