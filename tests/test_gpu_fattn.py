"""GPU parity tests (-m gpu) for GGML_OP_FLASH_ATTN_EXT through the C-ABI, against the CPU oracle
(oracle/ggml_oracle.c :: orc_flash_attn_ext, restating ggml-cpu/ops.cpp:8047-8272) and the reference's golden vectors.
Bars: NMSE < 5e-4 vs the ggml-cpu result (tests/test-backend-ops.cpp:6080-6082); NMSE < 1e-6 vs the oracle run in
fp32 (the kernel's own numerics), which is what catches indexing mistakes.
"""
import os
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def B():
    import ggml_b200
    return ggml_b200


@pytest.fixture(scope="module")
def ctx(B):
    c = B.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def orc():
    return O.Oracle()


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "hotpath_golden.npz"))


def make_mask(rng, nm, kv, frac=0.1):
    """~10 % of 16-wide runs set to -inf, the rest uniform, like init_tensor_kq_mask (tests/test-backend-ops.cpp:152-175)"""
    m = rng.uniform(-1, 1, (1, 1, nm, kv)).astype(np.float32)
    holes = rng.uniform(size=(nm, (kv + 15) // 16)) < frac
    m[0, 0][np.repeat(holes, 16, axis=1)[:, :kv]] = -np.inf
    return m.astype(np.float16)


@pytest.mark.parametrize("case,kw", [
    ("fa_out_plain", dict(mask=False, sinks=False)),
    ("fa_out_mask", dict(mask=True, sinks=False)),
    ("fa_out_mask_sinks", dict(mask=True, sinks=True)),
    ("fa_out_alibi", dict(mask=True, sinks=False, max_bias=8.0)),
    ("fa_out_softcap", dict(mask=True, sinks=False, softcap=10.0)),
    ("fa_out_f32prec", dict(mask=True, sinks=True)),
])
def test_flash_attn_golden(ctx, gold, case, kw):
    q, k, v = gold["fa_q"], gold["fa_k"], gold["fa_v"]
    mask = gold["fa_mask"] if kw["mask"] else None
    sinks = gold["fa_sinks"] if kw["sinks"] else None
    got = ctx.flash_attn_ext_host(q, k, v, mask, sinks, 1.0 / np.sqrt(q.shape[-1]), kw.get("max_bias", 0.0), kw.get("softcap", 0.0))
    assert np.isfinite(got).all()
    assert O.nmse(gold[case], got) < 5e-4


@pytest.mark.parametrize("hs", [64, 128, 256])
@pytest.mark.parametrize("nr2", [1, 4])
@pytest.mark.parametrize("kv", [113, 512, 1024])
@pytest.mark.parametrize("nb", [1, 3, 35])
def test_flash_attn_grid(ctx, orc, hs, nr2, kv, nb):
    """head size x GQA ratio x kv (113 is ragged) x batch, mask on (test-backend-ops :8191-8235)"""
    if hs == 256 and kv == 1024 and nb == 35:
        pytest.skip("oracle too slow for this corner")
    rng = np.random.default_rng(hs + nr2 * 7 + kv + nb)
    Hk = 2; H = Hk * nr2; S = 1
    q = rng.uniform(-1, 1, (S, H, nb, hs)).astype(np.float32)
    k = rng.uniform(-1, 1, (S, Hk, kv, hs)).astype(np.float16)
    v = rng.uniform(-1, 1, (S, Hk, kv, hs)).astype(np.float16)
    mask = make_mask(rng, nb, kv)
    sc = 1.0 / np.sqrt(hs)
    got = ctx.flash_attn_ext_host(q, k, v, mask, None, sc)
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, None, sc), got) < 1e-6
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, None, sc, faithful=True), got) < 5e-4


@pytest.mark.parametrize("opts", [dict(), dict(sinks=True), dict(max_bias=8.0), dict(softcap=10.0), dict(nomask=True),
                                  dict(sinks=True, max_bias=8.0, softcap=10.0)])
def test_flash_attn_options(ctx, orc, opts):
    rng = np.random.default_rng(len(str(opts)))
    S, H, Hk, N, KV, D = 2, 8, 2, 5, 200, 128
    q = rng.uniform(-1, 1, (S, H, N, D)).astype(np.float32)
    k = rng.uniform(-1, 1, (S, Hk, KV, D)).astype(np.float16)
    v = rng.uniform(-1, 1, (S, Hk, KV, D)).astype(np.float16)
    mask = None if opts.get("nomask") else make_mask(rng, 8, KV)         # mask rows padded beyond N
    sinks = rng.uniform(-10, 10, H).astype(np.float32) if opts.get("sinks") else None
    args = (0.0884, opts.get("max_bias", 0.0), opts.get("softcap", 0.0))
    got = ctx.flash_attn_ext_host(q, k, v, mask, sinks, *args)
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, sinks, *args), got) < 1e-6


def test_flash_attn_fully_masked_rows_and_heads_mask(ctx, orc):
    rng = np.random.default_rng(2)
    S, H, Hk, N, KV, D = 1, 4, 4, 4, 96, 64
    q = rng.uniform(-1, 1, (S, H, N, D)).astype(np.float32)
    k = rng.uniform(-1, 1, (S, Hk, KV, D)).astype(np.float16)
    v = rng.uniform(-1, 1, (S, Hk, KV, D)).astype(np.float16)
    mask = rng.uniform(-1, 1, (1, H, N, KV)).astype(np.float16)          # per-head mask (ne2 = H)
    mask[0, :, 2, :] = -np.inf                                           # row 2 fully masked -> zeros, not NaN
    got = ctx.flash_attn_ext_host(q, k, v, mask, None, 0.125)
    want = orc.flash_attn_ext(q, k, v, mask, None, 0.125)
    assert np.isfinite(got).all() and np.all(got[0, 2] == 0.0)
    assert O.nmse(want, got) < 1e-6


@pytest.mark.parametrize("n_batch", [1, 7, 64])
def test_flash_attn_kv_cache_views(ctx, B, orc, n_batch):
    """K and V as the KV cache hands them over (src/llama-kv-cache.cpp:1008-1058 + the permute in
    src/llama-graph.cpp:1556): one [n_embd_gqa, kv_size] buffer per layer, heads interleaved inside each token row,
    n_kv < kv_size, so nb1 = token stride and nb2 = head_dim * 2."""
    rng = np.random.default_rng(n_batch)
    H, Hk, D, kv_size, n_kv = 8, 2, 128, 768, 512
    cache_k = rng.uniform(-1, 1, (kv_size, Hk, D)).astype(np.float16)
    cache_v = rng.uniform(-1, 1, (kv_size, Hk, D)).astype(np.float16)
    q = rng.uniform(-1, 1, (1, H, n_batch, D)).astype(np.float32)
    mask = make_mask(rng, max(n_batch, 8), n_kv)
    dq = ctx.upload(B.F32, [D, n_batch, H, 1], q)
    dck = ctx.upload(B.F16, [D * Hk, kv_size], cache_k)
    dcv = ctx.upload(B.F16, [D * Hk, kv_size], cache_v)
    tok, head = Hk * D * 2, D * 2
    kview = dck.view([D, n_kv, Hk, 1], [2, tok, head, tok * kv_size])
    vview = dcv.view([D, n_kv, Hk, 1], [2, tok, head, tok * kv_size])
    dm = ctx.upload(B.F16, [n_kv, mask.shape[2], 1, 1], mask)
    dd = ctx.empty(B.F32, [D, H, n_batch, 1])
    ctx.flash_attn_ext(dq, kview, vview, dm, None, dd, 0.0884)
    got = ctx.download(dd).reshape(1, n_batch, H, D)
    kc = np.ascontiguousarray(cache_k[:n_kv].transpose(1, 0, 2))[None]
    vc = np.ascontiguousarray(cache_v[:n_kv].transpose(1, 0, 2))[None]
    assert O.nmse(orc.flash_attn_ext(q, kc, vc, mask, None, 0.0884), got) < 1e-6


def test_flash_attn_decode_8k_context(ctx, orc):
    """Llama-3-8B decode attention at 8k context: 32 heads / 8 kv heads, head 128, one query row"""
    rng = np.random.default_rng(8)
    H, Hk, D, KV = 32, 8, 128, 8192
    q = rng.standard_normal((1, H, 1, D)).astype(np.float32)
    k = rng.standard_normal((1, Hk, KV, D)).astype(np.float16)
    v = rng.standard_normal((1, Hk, KV, D)).astype(np.float16)
    mask = np.zeros((1, 1, 8, KV), dtype=np.float16); mask[..., 8000:] = -np.inf
    got = ctx.flash_attn_ext_host(q, k, v, mask, None, 0.0884)
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, None, 0.0884), got) < 1e-6


def causal_mask(n_batch, n_kv, n_past, pad=64):
    """the KQ mask llama.cpp builds for a prefill ubatch: query i sees keys 0..n_past+i, everything else -inf; rows padded"""
    nm = (n_batch + pad - 1) // pad * pad
    m = np.full((1, 1, nm, n_kv), -np.inf, dtype=np.float32)
    for i in range(n_batch):
        m[0, 0, i, : n_past + i + 1] = 0.0
    return m.astype(np.float16)


@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("n_batch,n_kv,n_past", [(512, 512, 0), (200, 768, 300), (64, 1024, 900), (16, 256, 17), (129, 130, 1)])
def test_flash_attn_prefill_causal(ctx, orc, D, n_batch, n_kv, n_past):
    """prefill attention (tcgen05 kernel): causal mask with whole tiles masked (skipped), partial tiles, ragged query and
    key counts, GQA; checked against the fp32 oracle and against the reference's fp16 numerics"""
    rng = np.random.default_rng(D + n_batch + n_kv)
    H, Hk = 8, 2
    q = rng.standard_normal((1, H, n_batch, D)).astype(np.float32)
    k = rng.standard_normal((1, Hk, n_kv, D)).astype(np.float16)
    v = rng.standard_normal((1, Hk, n_kv, D)).astype(np.float16)
    mask = causal_mask(n_batch, n_kv, n_past)
    sc = 1.0 / np.sqrt(D)
    got = ctx.flash_attn_ext_host(q, k, v, mask, None, sc)
    assert np.isfinite(got).all()
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, None, sc), got) < 2e-6
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, None, sc, faithful=True), got) < 5e-4


@pytest.mark.parametrize("opts", [dict(sinks=True), dict(max_bias=8.0), dict(softcap=10.0), dict(nomask=True), dict(sinks=True, max_bias=8.0, softcap=10.0)])
def test_flash_attn_prefill_options(ctx, orc, opts):
    rng = np.random.default_rng(len(str(opts)) + 99)
    S, H, Hk, N, KV, D = 2, 4, 2, 150, 333, 128
    q = rng.uniform(-1, 1, (S, H, N, D)).astype(np.float32)
    k = rng.uniform(-1, 1, (S, Hk, KV, D)).astype(np.float16)
    v = rng.uniform(-1, 1, (S, Hk, KV, D)).astype(np.float16)
    mask = None if opts.get("nomask") else make_mask(rng, 192, KV)
    sinks = rng.uniform(-10, 10, H).astype(np.float32) if opts.get("sinks") else None
    args = (0.0884, opts.get("max_bias", 0.0), opts.get("softcap", 0.0))
    got = ctx.flash_attn_ext_host(q, k, v, mask, sinks, *args)
    assert O.nmse(orc.flash_attn_ext(q, k, v, mask, sinks, *args), got) < 2e-6


def test_flash_attn_rejects_unsupported(ctx, B):
    dq = ctx.empty(B.F32, [40, 1, 1, 1]); dk = ctx.empty(B.F16, [40, 8, 1, 1]); dd = ctx.empty(B.F32, [40, 1, 1, 1])
    with pytest.raises(B.B200Error):
        ctx.flash_attn_ext(dq, dk, dk, None, None, dd, 1.0)


def test_flash_attn_gqa_decode_kernel_all_shapes():
    """every decode-sized case again with GGML_B200_FATTN_GQA=1, which also sends one-token / short-context calls (by default the
    gather kernel's) through the GQA-grouped mma.sync kernel: ragged kv, head sizes 64 / 128, sinks, softcap, ALiBi, KV-cache views"""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, GGML_B200_FATTN_GQA="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_fattn.py"), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "not prefill and not all_shapes"], env=env, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]



@pytest.mark.parametrize("n_kv,n_past", [(8704, 8192), (10240, 9728)])
def test_flash_attn_prefill_8k_depth_vs_reference(ctx, B, n_kv, n_past):
    """BASELINE.json configs[2] (pp2048 @ d8192, n_ubatch 512): the first and the last micro-batch's attention, 32 query heads over
    8 kv heads of 128, K/V as strided views of a 10240-token cache (heads interleaved inside a token row), causal mask.  Checked
    against the compiled reference itself (oracle/_ref, ggml-cpu's FLASH_ATTN_EXT with its fp16 K/V arithmetic): NMSE < 5e-4,
    the bar of tests/test-backend-ops.cpp:6080-6082."""
    if not O.have_reference():
        pytest.skip("oracle/_ref not built")
    ref = O.Reference()
    rng = np.random.default_rng(n_kv)
    H, Hk, D, kv_size, n_batch = 32, 8, 128, 10240, 512
    cache_k = rng.standard_normal((kv_size, Hk, D)).astype(np.float16)
    cache_v = rng.standard_normal((kv_size, Hk, D)).astype(np.float16)
    q = rng.standard_normal((1, H, n_batch, D)).astype(np.float32)
    mask = causal_mask(n_batch, n_kv, n_past)
    sc = 1.0 / np.sqrt(D)
    dq = ctx.upload(B.F32, [D, n_batch, H, 1], q)
    dck = ctx.upload(B.F16, [D * Hk, kv_size], cache_k)
    dcv = ctx.upload(B.F16, [D * Hk, kv_size], cache_v)
    tok, head = Hk * D * 2, D * 2
    kview = dck.view([D, n_kv, Hk, 1], [2, tok, head, tok * kv_size])
    vview = dcv.view([D, n_kv, Hk, 1], [2, tok, head, tok * kv_size])
    dm = ctx.upload(B.F16, [n_kv, mask.shape[2], 1, 1], mask)
    dd = ctx.empty(B.F32, [D, H, n_batch, 1])
    ctx.flash_attn_ext(dq, kview, vview, dm, None, dd, sc)
    got = ctx.download(dd).reshape(1, n_batch, H, D)
    assert np.isfinite(got).all()
    kc = np.ascontiguousarray(cache_k[:n_kv].transpose(1, 0, 2))[None]
    vc = np.ascontiguousarray(cache_v[:n_kv].transpose(1, 0, 2))[None]
    want = ref.flash_attn_ext(q, kc, vc, mask, None, sc)
    assert O.nmse(want, got) < 5e-4
    # the rows of one query over its own visible keys only: a second, size-independent check (softmax weights sum to 1, so the
    # output of a query lies inside the per-dimension range of the V rows it can see)
    vis = cache_v[: n_past + 1].astype(np.float32)
    assert (got[0, 0] <= vis.max(axis=0).reshape(Hk, 1, D).repeat(H // Hk, 1).reshape(H, D) + 1e-3).all()
    for d in (dq, dck, dcv, dm, dd):
        d.free()
