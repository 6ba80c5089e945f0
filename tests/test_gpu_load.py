"""GPU tests (-m gpu) of the weight load path b200_load_file / b200_load_row_slice (include/ggml-b200.h, csrc/b200_load.cpp)
through the C-ABI: what lands on the device must be the file's bytes, bit for bit -- whole tensors at unaligned file offsets through a
staging ring much smaller than a tensor, tensors re-laid out by the overlapped repack (compared with the upload + b200_repack path the
plugin's buffer hook takes), and row slices read rank by rank (the partition of b200_split_row_range, ggml-cuda.cu:777-803)."""
import os
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def B():
    import ggml_b200
    return ggml_b200


@pytest.fixture(scope="module")
def ctx(B):
    c = B.Context(0)
    yield c
    c.close()


def write_file(tmp_path, rng, specs, pad=37):
    """specs: [(type, K, M)] -> (path, [(offset, nbytes, host bytes)]); tensors separated by `pad` junk bytes so that no offset is aligned"""
    path = os.path.join(tmp_path, "weights.bin")
    out = []
    with open(path, "wb") as f:
        f.write(rng.integers(0, 256, 11, dtype=np.uint8).tobytes())
        for t, K, M in specs:
            raw = O.random_blocks(rng, t, K, M)
            out.append((f.tell(), raw.size, raw))
            f.write(raw.tobytes())
            f.write(rng.integers(0, 256, pad, dtype=np.uint8).tobytes())
    return path, out


def test_load_whole_tensors_bit_exact(ctx, B, tmp_path):
    rng = np.random.default_rng(1)
    specs = [(O.Q4_K, 4096, 96), (O.Q6_K, 2048, 64), (O.Q8_0, 1024, 33), (O.Q4_0, 512, 7), (O.F32, 100, 3)]
    path, blobs = write_file(str(tmp_path), rng, specs)
    tens = [ctx.empty(t, [K, M]) for t, K, M in specs]
    items = [(off, n, d, 0, 0) for (off, n, _), d in zip(blobs, tens)]
    secs, read_s, nbytes, chunks = ctx.load_file(path, items, chunk_bytes=64 * 1024, n_buffers=3)      # ring of 192 KiB, tensors up to 221 KiB
    assert nbytes == sum(n for _, n, _ in blobs) and chunks >= len(specs) and secs > 0
    for (off, n, raw), d in zip(blobs, tens):
        assert np.array_equal(ctx.download(d, np.uint8)[:n], raw)


def test_load_with_overlapped_repack_equals_upload_then_repack(ctx, B, tmp_path):
    rng = np.random.default_rng(2)
    specs = [(O.Q6_K, 4096, 256), (O.Q4_K, 4096, 128), (O.Q8_0, 2048, 96), (O.Q4_0, 2048, 96), (O.Q6_K, 2048, 32)]
    path, blobs = write_file(str(tmp_path), rng, specs)
    tens = [ctx.empty(t, [K, M]) for t, K, M in specs]
    want_repack = [bool(B.lib().b200_repack_supported(t, K)) for t, K, M in specs]
    assert any(want_repack) and not all(want_repack)                      # q4_K stays in GGUF layout
    ctx.load_file(path, [(off, n, d, 0, int(rp)) for (off, n, _), d, rp in zip(blobs, tens, want_repack)], chunk_bytes=128 * 1024, n_buffers=2)
    for (t, K, M), (off, n, raw), d, rp in zip(specs, blobs, tens, want_repack):
        assert d.flags == (1 if rp else 0)
        ref = ctx.upload(t, [K, M], raw)
        if rp:
            ctx.repack(ref)
        assert np.array_equal(ctx.download(d, np.uint8), ctx.download(ref, np.uint8))
        # ... and the re-laid-out tensor multiplies like the GGUF one
        x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
        dx = ctx.upload(B.F32, [K, 1], x); y0, y1 = ctx.empty(B.F32, [M, 1]), ctx.empty(B.F32, [M, 1])
        ctx.mul_mat(d, dx, y0); ctx.mul_mat(ctx.upload(t, [K, M], raw), dx, y1)
        assert O.nmse(ctx.download(y1), ctx.download(y0)) < 1e-9


@pytest.mark.parametrize("world", [2, 3, 8])
def test_load_row_slices_per_rank(ctx, B, tmp_path, world):
    """every rank reads only its rows of every tensor: the slices, put together in rank order, are the file's tensor"""
    rng = np.random.default_rng(3 + world)
    specs = [(O.Q4_K, 4096, 1024), (O.Q6_K, 2048, 512)]
    path, blobs = write_file(str(tmp_path), rng, specs)
    for (t, K, M), (off, n, raw) in zip(specs, blobs):
        rb = O.row_size(t, K)
        parts = []
        for rank in range(world):
            lo, hi = B.split_row_range(M, world, rank, 64)
            it = B.LoadItem()
            assert B.lib().b200_load_row_slice(t, K, off, lo, hi, it) == 0
            assert it.file_offset == off + lo * rb and it.nbytes == (hi - lo) * rb
            if hi == lo:
                continue
            d = ctx.empty(t, [K, hi - lo])
            ctx.load_file(path, [(it.file_offset, it.nbytes, d, 0, 0)], chunk_bytes=32 * 1024, n_buffers=4)
            parts.append(ctx.download(d, np.uint8))
        assert np.array_equal(np.concatenate(parts), raw)


def test_load_rejects_bad_requests(ctx, B, tmp_path):
    rng = np.random.default_rng(4)
    path, blobs = write_file(str(tmp_path), rng, [(O.Q4_K, 256, 4)])
    d = ctx.empty(O.Q4_K, [256, 4])
    with pytest.raises(B.B200Error):
        ctx.load_file(os.path.join(str(tmp_path), "missing.bin"), [(0, 16, d, 0, 0)])
    with pytest.raises(B.B200Error):
        ctx.load_file(path, [(0, d.nbytes + 1, d, 0, 0)])                 # does not fit the tensor
    with pytest.raises(B.B200Error):
        ctx.load_file(path, [(blobs[0][0], d.nbytes, d, 0, 1)])           # q4_K has no device layout
    with pytest.raises(B.B200Error):
        ctx.load_file(path, [(os.path.getsize(path) - 8, 64, d, 0, 0)])   # the range runs past the end of the file
