"""GPU test (-m gpu, ONE GPU): split-mode=row with N shards faked on device 0.  N contexts -- N streams -- in one
process, one exchange arena each, wired to each other as "peers" (PeerExchange.local_group): b200_mul_mat_rowsplit / _multi run exactly
the code the multi-GPU bench times (row partition of ggml-cuda.cu:777-803, exchange epilogue of the streaming / mma GEMV, receive
kernels, tag counters), only the peer stores stay on the device.  Every shard must end up with the complete column, bit for bit the
column the unsplit b200_mul_mat gives, and the oracle's within the q8_1 bar."""
import os
import sys

import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))

CASES = [(O.Q4_K, 4096, 1024, False), (O.Q6_K, 4096, 512, True), (O.Q4_K, 2048, 96, False), (O.Q8_0, 1024, 640, True), (O.Q4_0, 512, 256, True),
         (O.Q4_K, 4096, 14336, False)]


@pytest.fixture(scope="module")
def B():
    import ggml_b200
    return ggml_b200


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_rowsplit_n_shards_on_one_gpu(B, world):
    from row_split import RowSplit, PeerExchange
    ctxs = [B.Context(0) for _ in range(world)]
    peers = PeerExchange.local_group(ctxs, 8 << 20)
    splits = [RowSplit(None, world, r, 32) for r in range(world)]
    orc = O.Oracle()
    whole = B.Context(0)
    bufs = {}
    try:
        for rep in range(3):                                   # several rounds through the same buffers: the tag counters keep counting
            for ci, (t, K, M, repack) in enumerate(CASES):
                rng = np.random.default_rng(K + M + rep)
                w = O.random_blocks(rng, t, K, M)
                x = rng.standard_normal((1, K)).astype(np.float32)
                wd = whole.upload(t, [K, M], w)
                if repack: whole.repack(wd)
                xd = whole.upload(B.F32, [K, 1], x); dd = whole.empty(B.F32, [M, 1])
                whole.mul_mat(wd, xd, dd); want = whole.download(dd).reshape(-1).copy()
                xd.free(); dd.free()
                wd.free()
                assert O.nmse(orc.mul_mat(t, w, K, M, x, 1, act=O.ACT_Q8_1).reshape(-1), want) < 1e-9
                shards = []
                for r in range(world):
                    mine, lo, hi = splits[r].shard_weight(w, w.size // M, M)
                    dw = ctxs[r].upload(t, [K, hi - lo], mine)
                    if repack and hi > lo: ctxs[r].repack(dw)
                    elif repack: dw.flags = 1
                    if (r, ci) not in bufs:
                        bufs[(r, ci)] = (peers[r].alloc(B.F32, [K, 1]), peers[r].alloc(B.F32, [M, 1]))
                    ctxs[r].h2d(bufs[(r, ci)][0].ptr, x)
                    shards.append((dw, lo))
                for r in range(world): ctxs[r].synchronize()
                for r in range(world):                         # all launches are asynchronous: every rank's receive kernel polls while the others' GEMVs run
                    peers[r].mul_mat(shards[r][0], bufs[(r, ci)][0], bufs[(r, ci)][1], shards[r][1])
                for r in range(world): ctxs[r].synchronize()
                for r in range(world):
                    got = ctxs[r].download(bufs[(r, ci)][1]).reshape(-1)
                    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (world, r, rep, ci, int((got != want).sum()))
                    shards[r][0].free()
        # wq / wk / wv of one column in one call (two q4_K share a launch, the repacked q6_K gets its own); M = 32 leaves ranks without rows
        multi = [(O.Q4_K, 1024, False), (O.Q4_K, 32, False), (O.Q6_K, 256, True)]
        K = 2048
        mx = [p.alloc(B.F32, [K, 1]) for p in peers]
        md = [[p.alloc(B.F32, [M, 1]) for _, M, _ in multi] for p in peers]
        for rep in range(3):
            rng = np.random.default_rng(77 + rep)
            x = rng.standard_normal((1, K)).astype(np.float32)
            full = [O.random_blocks(rng, t, K, M) for t, M, _ in multi]
            per_rank = []
            for r in range(world):
                ctxs[r].h2d(mx[r].ptr, x)
                ws, los = [], []
                for (t, M, repack), w in zip(multi, full):
                    mine, lo, hi = splits[r].shard_weight(w, w.size // M, M)
                    dw = ctxs[r].upload(t, [K, hi - lo], mine)
                    if repack and hi > lo: ctxs[r].repack(dw)
                    elif repack: dw.flags = 1
                    ws.append(dw); los.append(lo)
                per_rank.append((ws, los))
            for r in range(world): ctxs[r].synchronize()
            for r in range(world): peers[r].mul_mat_multi(per_rank[r][0], mx[r], md[r], per_rank[r][1])
            for r in range(world): ctxs[r].synchronize()
            for (t, M, _), w, j in zip(multi, full, range(len(multi))):
                want = orc.mul_mat(t, w, K, M, x, 1, act=O.ACT_Q8_1).reshape(-1)
                got0 = ctxs[0].download(md[0][j]).reshape(-1)
                assert O.nmse(want, got0) < 1e-9
                for r in range(1, world):
                    assert np.array_equal(ctxs[r].download(md[r][j]).reshape(-1).view(np.uint32), got0.view(np.uint32))
            for ws, _ in per_rank:
                for dw in ws: dw.free()
        for r in range(world):
            assert ctxs[r].exchange_error(peers[r].ex) == 0
    finally:
        for c in ctxs + [whole]:
            c.close()
