"""split-mode=row tensor parallelism for MUL_MAT, one process per GPU.

The reference shards a weight's rows over the GPUs of one process inside its split buffer type
(ggml/src/ggml-cuda/ggml-cuda.cu:777-803 for the partition, :1451-1769 for the copy/compute/copy loop).  Here every rank is its
own process with its own ggml-b200 context: rank r holds rows [lo, hi) of every weight (same partition rule, through the
C-ABI's b200_split_row_range), computes the matching rows of dst in place inside the full-size dst buffer, and the slices are
exchanged with ONE all-gather (torch.distributed: NCCL over NVLink on GPUs, gloo in the CPU tests).  Activations are
replicated, so the matmul itself needs no other collective.
"""
import numpy as np

import ggml_b200 as B


class RowSplit:
    def __init__(self, dist, world, rank, rounding=32, tensor_split=None):
        self.dist, self.world, self.rank, self.rounding, self.tensor_split = dist, world, rank, rounding, tensor_split

    def row_range(self, nrows, rank=None):
        """rows [lo, hi) owned by `rank` (default: this rank)"""
        if self.world == 1:
            return 0, nrows
        return B.split_row_range(nrows, self.world, self.rank if rank is None else rank, self.rounding, self.tensor_split)

    def ranges(self, nrows):
        return [self.row_range(nrows, r) for r in range(self.world)]

    def shard_weight(self, blocks, row_bytes, nrows):
        """this rank's rows of a host-side weight given as a flat uint8 array of GGUF blocks (rows are contiguous)"""
        lo, hi = self.row_range(nrows)
        return np.ascontiguousarray(blocks.reshape(-1)[lo * row_bytes: hi * row_bytes]), lo, hi

    def gather(self, full, nrows, ncols=1):
        """`full` is a torch tensor viewing dst as [ncols, nrows] (ggml [nrows, ncols]) in which this rank has filled rows
        [lo, hi) of every column; on return every rank holds all rows.  In place when the slices are equal (the model
        shapes), through padded staging otherwise."""
        if self.world == 1:
            return
        import torch
        rs = self.ranges(nrows)
        lo, hi = rs[self.rank]
        sizes = [b - a for a, b in rs]
        t = full.view(ncols, nrows)
        if ncols == 1 and len(set(sizes)) == 1:
            self.dist.all_gather_into_tensor(t.view(-1), t.view(-1)[lo:hi])
            return
        mx = max(sizes)
        mine = torch.zeros((ncols, mx), dtype=t.dtype, device=t.device)
        mine[:, : hi - lo] = t[:, lo:hi]
        parts = [torch.empty_like(mine) for _ in range(self.world)]
        self.dist.all_gather(parts, mine)
        for (a, b), part in zip(rs, parts):
            t[:, a:b] = part[:, : b - a]
