"""GPU parity tests (-m gpu) for b200_flash_attn_ext_paged: decode attention over a PAGED KV cache (intended semantics of the reference's
src/llama-kv-cache-paged.cpp:137-196: a layer's cache is a set of fixed-size pages allocated on demand).  The pool holds more pages than
the sequence uses, in SHUFFLED order; the page table maps logical pages to pool pages.  Checked against the CPU oracle on the gathered
contiguous cache (NMSE < 1e-6, the kernels' own fp32 numerics) and, word for word, against b200_flash_attn_ext on the gathered cache
uploaded contiguously (same kernels, same order of operations: the page table only changes where a row is fetched from)."""
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


# (n_batch, n_kv, page_tokens, identity): 256 keys -> the row-gather kernel; > 512 keys or several tokens -> the GQA tensor-core kernel (TMA tiles);
# 1000 keys: not a whole number of 16-key tiles, the GQA kernel declines the paged form and the gather kernel takes it
@pytest.mark.parametrize("n_batch,n_kv,page,identity", [(1, 256, 256, False), (1, 4096, 256, False), (1, 8192, 256, True), (4, 1024, 64, False),
                                                        (1, 1000, 128, False), (2, 768, 16, False), (1, 32768, 256, False)])
@pytest.mark.parametrize("D", [128, 64])
def test_flash_attn_paged_shuffled_pages(ctx, B, D, n_batch, n_kv, page, identity):
    rng = np.random.default_rng(D + n_batch + n_kv + page)
    H, Hk = 32, 8
    n_pages = (n_kv + page - 1) // page
    pool_pages = n_pages if identity else n_pages + n_pages // 2 + 3
    pool_k = rng.standard_normal((pool_pages * page, Hk, D)).astype(np.float16)
    pool_v = rng.standard_normal((pool_pages * page, Hk, D)).astype(np.float16)
    table = np.arange(n_pages, dtype=np.int32) if identity else rng.permutation(pool_pages)[:n_pages].astype(np.int32)
    rows = (table.astype(np.int64)[:, None] * page + np.arange(page)[None, :]).reshape(-1)[:n_kv]
    q = rng.standard_normal((1, H, n_batch, D)).astype(np.float32)
    nm = max(8, (n_batch + 63) // 64 * 64)
    mask = np.zeros((1, 1, nm, n_kv), dtype=np.float16)
    for i in range(n_batch):                                   # causal over the last n_batch positions, plus a masked run in the middle
        mask[0, 0, i, n_kv - n_batch + i + 1:] = -np.inf
    if n_kv >= 512:
        mask[0, 0, :, 100:164] = -np.inf
    sc = 1.0 / np.sqrt(D)
    tok, head = Hk * D * 2, D * 2
    dq = ctx.upload(B.F32, [D, n_batch, H, 1], q)
    dpk = ctx.upload(B.F16, [D * Hk, pool_pages * page], pool_k)
    dpv = ctx.upload(B.F16, [D * Hk, pool_pages * page], pool_v)
    kview = dpk.view([D, n_kv, Hk, 1], [2, tok, head, tok * pool_pages * page])
    vview = dpv.view([D, n_kv, Hk, 1], [2, tok, head, tok * pool_pages * page])
    dt = ctx.upload(B.I32, [n_pages], table)
    dm = ctx.upload(B.F16, [n_kv, nm, 1, 1], mask)
    dd = ctx.empty(B.F32, [D, H, n_batch, 1])
    ctx.flash_attn_ext_paged(dq, kview, vview, dt, page, pool_pages * page, dm, None, dd, sc)
    got = ctx.download(dd).reshape(1, n_batch, H, D).copy()
    assert np.isfinite(got).all()
    # the gathered contiguous cache, as llama-kv-cache.cpp lays it out (heads interleaved inside a token row)
    gk, gv = np.ascontiguousarray(pool_k[rows]), np.ascontiguousarray(pool_v[rows])
    dck = ctx.upload(B.F16, [D * Hk, n_kv], gk); dcv = ctx.upload(B.F16, [D * Hk, n_kv], gv)
    ck = dck.view([D, n_kv, Hk, 1], [2, tok, head, tok * n_kv]); cv = dcv.view([D, n_kv, Hk, 1], [2, tok, head, tok * n_kv])
    ctx.flash_attn_ext(dq, ck, cv, dm, None, dd, sc)
    flat = ctx.download(dd).reshape(1, n_batch, H, D)
    if n_kv % 16 == 0:                                         # (else the two calls take different kernels: the comparison with the oracle below decides)
        assert np.array_equal(flat.view(np.uint32), got.view(np.uint32))
    else:
        assert O.nmse(flat, got) < 1e-6
    want = O.Oracle().flash_attn_ext(q, gk.transpose(1, 0, 2)[None], gv.transpose(1, 0, 2)[None], mask, None, sc)
    assert O.nmse(want, got) < 1e-6
    for d in (dq, dpk, dpv, dt, dm, dd, dck, dcv):
        d.free()


def test_flash_attn_paged_rejects(ctx, B):
    """bad page geometry is an error, prefill-sized batches are declined (UNSUPPORTED), nothing falls back silently"""
    H, Hk, D, n_kv = 8, 2, 128, 512
    dq = ctx.empty(B.F32, [D, 1, H, 1]); dk = ctx.empty(B.F16, [D, n_kv, Hk, 1]); dv = ctx.empty(B.F16, [D, n_kv, Hk, 1])
    dd = ctx.empty(B.F32, [D, H, 1, 1]); dt = ctx.upload(B.I32, [2], np.arange(2, dtype=np.int32))
    with pytest.raises(B.B200Error):
        ctx.flash_attn_ext_paged(dq, dk, dv, dt, 100, 512, None, None, dd, 0.1)          # not a power of two
    with pytest.raises(B.B200Error):
        ctx.flash_attn_ext_paged(dq, dk, dv, dt, 128, 512, None, None, dd, 0.1)          # 2 pages x 128 < 512 keys
    dq2 = ctx.empty(B.F32, [D, 64, H, 1]); dd2 = ctx.empty(B.F32, [D, H, 64, 1])
    with pytest.raises(B.B200Error):
        ctx.flash_attn_ext_paged(dq2, dk, dv, dt, 256, 512, None, None, dd2, 0.1)        # a prefill batch
