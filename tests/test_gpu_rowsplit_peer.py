"""GPU test (-m gpu, needs 2 GPUs): split-mode=row with the all-gather fused into the streaming GEMV over NVLink peer memory
(b200_mul_mat_rowsplit).  Two processes, one per GPU; torch.distributed (gloo) only carries the IPC
handles and the start-up barrier.  Every rank must end up with the complete, bit-exact result in its own arena."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = r'''
import os, sys
import numpy as np, torch.distributed as dist
ROOT = sys.argv[1]
sys.path[:0] = [os.path.join(ROOT, "super-llama.cpp_b200"), os.path.join(ROOT, "oracle")]
import oracle as O, ggml_b200 as B
from row_split import RowSplit, PeerExchange
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
ctx = B.Context(rank)
orc = O.Oracle()
peer = PeerExchange(ctx, dist, world, rank, 8 << 20)
rs = RowSplit(dist, world, rank, 32)
ok = True
cases = [(O.Q4_K, 4096, 1024, False), (O.Q6_K, 4096, 512, True), (O.Q4_K, 2048, 96, False), (O.Q8_0, 1024, 640, True), (O.Q4_0, 512, 256, True)]
dsts = []
for rep in range(3):                                   # several rounds through the same buffers: counters keep counting
    for ci, (t, K, M, repack) in enumerate(cases):
        rng = np.random.default_rng(K + M + rep)
        w = O.random_blocks(rng, t, K, M)
        x = rng.standard_normal((1, K)).astype(np.float32)
        mine, lo, hi = rs.shard_weight(w, w.size // M, M)
        dw = ctx.upload(t, [K, hi - lo], mine)
        if repack: ctx.repack(dw)
        if rep == 0:
            dsts.append((peer.alloc(B.F32, [K, 1]), peer.alloc(B.F32, [M, 1])))
        dx, dd = dsts[ci]
        ctx.h2d(dx.ptr, x)
        peer.mul_mat(dw, dx, dd, lo)
        got = ctx.download(dd).reshape(1, 1, 1, M)
        want = orc.mul_mat(t, w, K, M, x, 1, act=O.ACT_Q8_1)
        same = O.nmse(want, got) < 1e-9
        if not same:
            bad = np.nonzero(got.reshape(-1) != want.reshape(-1))[0]
            print(f"rank {rank} rep {rep} case {ci}: {bad.size} of {M} rows differ, first {bad[:8]}, mine [{lo},{hi})", flush=True)
        ok = ok and same
        sums = [None] * world; dist.all_gather_object(sums, got.tobytes())      # every rank holds the identical gathered tensor
        ok = ok and all(b == sums[0] for b in sums)
        dist.barrier()                                  # the next case reuses dx on the host side
# wq/wk/wv of one column in one call: q4_K, q4_K share a launch, the q6_K one (repacked) gets its own; M = 32 leaves rank 1 without rows
multi = [(O.Q4_K, 1024, False), (O.Q4_K, 32, False), (O.Q6_K, 256, True)]
K = 2048
mx = peer.alloc(B.F32, [K, 1])
md = [peer.alloc(B.F32, [M, 1]) for _, M, _ in multi]
for rep in range(3):
    rng = np.random.default_rng(77 + rep)
    x = rng.standard_normal((1, K)).astype(np.float32)
    ctx.h2d(mx.ptr, x)
    ws, los, full = [], [], []
    for t, M, repack in multi:
        w = O.random_blocks(rng, t, K, M)
        mine, lo, hi = rs.shard_weight(w, w.size // M, M)
        dw = ctx.upload(t, [K, hi - lo], mine)
        if repack and hi > lo: ctx.repack(dw)
        elif repack: dw.flags = 1
        ws.append(dw); los.append(lo); full.append(w)
    peer.mul_mat_multi(ws, mx, md, los)
    for (t, M, _), w, dd in zip(multi, full, md):
        got = ctx.download(dd).reshape(1, 1, 1, M)
        want = orc.mul_mat(t, w, K, M, x, 1, act=O.ACT_Q8_1)
        same = O.nmse(want, got) < 1e-9
        if not same: print(f"rank {rank} multi rep {rep} type {t} M {M}: nmse {O.nmse(want, got)}", flush=True)
        ok = ok and same
        sums = [None] * world; dist.all_gather_object(sums, got.tobytes())
        ok = ok and all(b == sums[0] for b in sums)
    dist.barrier()
err = ctx.exchange_error(peer.ex)
if err: print(f'rank {rank}: exchange error flag {err}', flush=True)
ok = ok and err == 0
flag = [None] * world; dist.all_gather_object(flag, bool(ok))
dist.destroy_process_group()
sys.exit(0 if all(flag) else 1)
'''


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


@pytest.mark.parametrize("mma", ["", "1"])           # "1": the raw q4_K cases go through the mma.sync GEMV's exchange epilogue
def test_rowsplit_peer_exchange_two_gpus(tmp_path, mma):
    sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
    import ggml_b200 as B
    if B.lib().b200_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = tmp_path / "worker.py"; script.write_text(WORKER)
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        if mma:
            env["GGML_B200_GEMV_MMA"] = mma
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=600)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
