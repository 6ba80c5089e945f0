"""GPU test (-m gpu) of b200_mul_mat_noalloc (include/ggml-b200.h): the allocation-free MUL_MAT entry point whose q8_1 scratch is a static
device array (super-llama.cpp_b200/csrc/b200_noalloc.cu).  No counterpart in the reference.

Called through the C-ABI and compared with the existing entry point b200_mul_mat on the same device tensors, twice in a row (the second
call finds the first call's image in the shared array), within the tolerance the MUL_MAT tests use against the kernel's own numerics
(NMSE < 1e-9; where both entry points run the very same kernel the results are also bit-identical), and against the oracle.
The file sorts last on purpose: a failure here must not hide the other tests.
"""
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
NMSE_HARNESS = 5e-4


@pytest.fixture(scope="module")
def B():
    import ggml_b200
    return ggml_b200


@pytest.fixture(scope="module")
def ctx(B):
    c = B.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("t", [O.Q4_K, O.Q8_0, O.Q6_K])
def test_mul_mat_noalloc_matches_mul_mat_two_calls(ctx, B, t):
    orc = O.Oracle()
    rng = np.random.default_rng(4242 + t)
    K, M = 4096, 1031
    w = O.random_blocks(rng, t, K, M)
    dw = ctx.upload(t, [K, M], w)
    for call, n in enumerate((1, 4)):                      # two calls in a row, different activations and widths: the shared image is rewritten
        x = rng.uniform(-1, 1, (n, K)).astype(np.float32)
        dx = ctx.upload(B.F32, [K, n], x)
        want_d, got_d = ctx.empty(B.F32, [M, n]), ctx.empty(B.F32, [M, n])
        ctx.memset(got_d.ptr, 0xff, got_d.nbytes)
        ctx.mul_mat(dw, dx, want_d)                        # the existing entry point
        ctx.mul_mat_noalloc(dw, dx, got_d)                 # the new one
        want, got = ctx.download(want_d).reshape(n, M), ctx.download(got_d).reshape(n, M)
        assert np.isfinite(got).all(), f"call {call}"
        assert O.nmse(want, got) < 1e-9, f"call {call}: nmse {O.nmse(want, got)} vs b200_mul_mat"
        if n > 1:                                          # b200_mul_mat runs the same two kernels for 2..8 columns
            assert np.array_equal(want, got), f"call {call}"
        assert O.nmse(orc.mul_mat(t, w, K, M, x, n, act=O.ACT_CPU), got) < NMSE_HARNESS
        # same inputs again: the image left in the array by the call before must not matter
        ctx.mul_mat_noalloc(dw, dx, got_d)
        assert np.array_equal(got, ctx.download(got_d).reshape(n, M)), f"call {call} repeated"


def test_mul_mat_noalloc_allocates_nothing_and_rejects_oversize(ctx, B):
    rng = np.random.default_rng(7)
    K, M = 2048, 64
    dw = ctx.upload(O.Q4_K, [K, M], O.random_blocks(rng, O.Q4_K, K, M))
    dx = ctx.upload(B.F32, [K, 2], rng.uniform(-1, 1, (2, K)).astype(np.float32))
    dd = ctx.empty(B.F32, [M, 2])
    gen = B.lib().b200_ctx_scratch_generation(ctx.h)
    ctx.mul_mat_noalloc(dw, dx, dd)
    ctx.synchronize()
    assert B.lib().b200_ctx_scratch_generation(ctx.h) == gen   # the context's cudaMalloc'ed scratch did not move or grow
    # the q8_1 image of ne[1] * ne[2] columns must fit the static array: K * ncol * 9 / 8 <= b200_noalloc_scratch_bytes()
    cap = B.lib().b200_noalloc_scratch_bytes()
    assert cap == 1 << 20
    Kbig, nb = 32768, 8                                    # 8 x 4 columns x 32768 x 9/8 = 1.125 MiB > 1 MiB
    wbig = ctx.empty(O.Q4_K, [Kbig, 8])
    xbig = ctx.empty(B.F32, [Kbig, nb, 4])
    dbig = ctx.empty(B.F32, [8, nb, 4])
    with pytest.raises(B.B200Error, match="static scratch"):
        ctx.mul_mat_noalloc(wbig, xbig, dbig)
    with pytest.raises(B.B200Error, match="at most 8"):    # more than 8 columns is the GEMM's business
        ctx.mul_mat_noalloc(dw, ctx.empty(B.F32, [K, 9]), ctx.empty(B.F32, [M, 9]))
