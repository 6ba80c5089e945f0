"""N>1 host logic of split-mode=row on CPU: two gloo ranks, each computing its row slice of MUL_MAT with the oracle and
exchanging slices through row_split.RowSplit.gather -- the same partition (b200_split_row_range through the C-ABI) and the
same gather code bench.py --gpus N runs over NCCL.  No GPU is used; the matmul stand-in is the CPU oracle."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = sys.argv[1]
sys.path[:0] = [os.path.join(ROOT, "super-llama.cpp_b200"), os.path.join(ROOT, "oracle")]
import oracle as O
from row_split import RowSplit
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
orc = O.Oracle()
ok = True
for t, K, M, N, rounding in ((O.Q4_K, 512, 256, 1, 32), (O.Q6_K, 256, 200, 1, 32), (O.Q8_0, 256, 67, 3, 32), (O.Q4_K, 256, 96, 5, 16), (O.Q4_0, 128, 64, 2, 64)):
    rng = np.random.default_rng(K + M)                      # same seed on every rank: replicated inputs
    w = O.random_blocks(rng, t, K, M)
    x = rng.standard_normal((N, K)).astype(np.float32)
    rs = RowSplit(dist, world, rank, rounding)
    ranges = rs.ranges(M)
    assert ranges[0][0] == 0 and ranges[-1][1] == M and all(a[1] == b[0] for a, b in zip(ranges, ranges[1:])), ranges
    assert all(lo % rounding == 0 for lo, _ in ranges), ranges
    row_bytes = w.size // M
    mine, lo, hi = rs.shard_weight(w, row_bytes, M)
    full = torch.zeros((N, M), dtype=torch.float32)
    if hi > lo:
        part = orc.mul_mat(t, mine, K, hi - lo, x, N, act=O.ACT_Q8_1).reshape(N, hi - lo)
        full[:, lo:hi] = torch.from_numpy(part)
    rs.gather(full, M, N)
    want = orc.mul_mat(t, w, K, M, x, N, act=O.ACT_Q8_1).reshape(N, M)
    ok = ok and np.array_equal(full.numpy(), want)
flag = torch.tensor([1 if ok else 0]); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
'''


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


@pytest.mark.parametrize("world", [2, 3])
def test_row_split_gather_matches_unsplit(tmp_path, world):
    import subprocess
    script = tmp_path / "worker.py"; script.write_text(WORKER)
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=240)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)


def test_split_row_range_rule():
    """partition rule of ggml-cuda.cu:777-803: cumulative fractions, boundaries rounded down to the row tile, last rank ends at nrows"""
    sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))
    import ggml_b200 as B
    for nrows, world, rounding in ((4096, 8, 32), (128256, 8, 32), (1000, 3, 64), (31, 4, 32), (14336, 7, 128)):
        rs = [B.split_row_range(nrows, world, r, rounding) for r in range(world)]
        assert rs[0][0] == 0 and rs[-1][1] == nrows
        for (a, b), (c, d) in zip(rs, rs[1:]):
            assert b == c and a <= b
        assert all(lo % rounding == 0 for lo, _ in rs)
    rs = [B.split_row_range(1024, 2, r, 32, [0.0, 0.75]) for r in range(2)]          # uneven tensor_split: device 1 starts at 75 %
    assert rs == [(0, 768), (768, 1024)]
