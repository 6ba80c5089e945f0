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


class PeerExchange:
    """The all-gather of split-mode=row done by the matmul kernel itself: every rank maps every other rank's exchange arena
    (CUDA IPC over NVLink), the streaming GEMV's epilogue stores its rows as (value, tag) pairs into all arenas and a receive kernel validates them; torch.distributed is only
    used to pass the 64-byte IPC handles around and for the start-up barrier."""

    CTRL_BYTES = 256

    def __init__(self, ctx, dist, world, rank, arena_bytes):
        self.ctx, self.world, self.rank = ctx, world, rank
        self.arena_bytes = (arena_bytes + 255) // 256 * 256
        total = 3 * self.arena_bytes + self.CTRL_BYTES        # tensors | control block | (value, tag) mirror of the tensor area
        self.base = ctx.malloc(total)
        B._check(B.lib().b200_memset(ctx.h, self.base, 0, total), "memset")
        ctx.synchronize()
        handles = [None] * world
        dist.all_gather_object(handles, ctx.peer_handle(self.base))
        self.ex = B.Exchange()
        self.ex.world, self.ex.rank, self.ex.ctrl_off = world, rank, self.arena_bytes
        for r in range(world):
            self.ex.base[r] = self.base if r == rank else ctx.peer_open(handles[r])
        self.cursor = 0
        dist.barrier()                                       # nobody signals before everybody has zeroed and mapped

    @classmethod
    def local_group(cls, ctxs, arena_bytes):
        """N exchanges inside ONE process: rank r is ctxs[r] (its own stream; the contexts may all sit on one device), the arenas are
        plain allocations of that process, so no IPC handles are passed.  The kernels cannot tell the difference: they only see
        base[0..world) -- which is what lets a one-GPU box exercise the exchange epilogue and the receive kernels."""
        world = len(ctxs)
        out = []
        for rank, ctx in enumerate(ctxs):
            self = cls.__new__(cls)
            self.ctx, self.world, self.rank = ctx, world, rank
            self.arena_bytes = (arena_bytes + 255) // 256 * 256
            total = 3 * self.arena_bytes + cls.CTRL_BYTES
            self.base = ctx.malloc(total)
            B._check(B.lib().b200_memset(ctx.h, self.base, 0, total), "memset")
            ctx.synchronize()
            self.cursor = 0
            out.append(self)
        for self in out:
            self.ex = B.Exchange()
            self.ex.world, self.ex.rank, self.ex.ctrl_off = world, self.rank, self.arena_bytes
            for r in range(world):
                self.ex.base[r] = out[r].base
        return out

    def alloc(self, t, ne):
        """carve a tensor out of the arena: same offset on every rank (every rank allocates in the same order)"""
        dt = B.DeviceTensor(self.ctx, t, ne, ptr=self.base + self.cursor)
        self.cursor += (dt.nbytes + 255) // 256 * 256
        assert self.cursor <= self.arena_bytes, "exchange arena too small"
        return dt

    def mul_mat(self, w_local, x, dst_full, row_lo):
        self.ctx.mul_mat_rowsplit(w_local, x, dst_full, row_lo, self.ex)

    def mul_mat_multi(self, ws_local, x, dsts_full, row_los):
        """up to 3 row-split MUL_MATs of one activation column in one launch per run of equal weight type"""
        self.ctx.mul_mat_rowsplit_multi(ws_local, x, dsts_full, row_los, self.ex)
