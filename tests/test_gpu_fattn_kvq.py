"""GPU parity tests (-m gpu) for GGML_OP_FLASH_ATTN_EXT over a QUANTIZED KV cache (q8_0 / q4_0 K and V) and for the
SET_ROWS that writes such a cache, through the C-ABI.  Checked against
  * the compiled reference (oracle/_ref: ggml-cpu's FLASH_ATTN_EXT with a quantized K / V, ops.cpp:8047-8272) where it is built: NMSE < 5e-4,
    the bar of tests/test-backend-ops.cpp:6080-6082 (measured ~1e-7: same q8_0 query rows, same integer dot products);
  * a float64 attention over the DEQUANTIZED cache (numpy, independent of both implementations): NMSE < 5e-4 (the query's own q8_0
    rounding is the difference).
The reference's unmodified test-backend-ops covers the same op with type_KV = q8_0 / q4_0 in tests/test_gpu_backend_ops.py."""
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


@pytest.fixture(scope="module")
def orc():
    return O.Oracle()


def quant_rows(orc, t, x):
    """x[..., D] f32 -> raw blocks, row by row (quantize_row_q8_0_ref / q4_0_ref: what SET_ROWS writes into the cache)"""
    return orc.quantize_weights(t, np.ascontiguousarray(x, dtype=np.float32))


def attn_f64(q, k, v, mask, scale):
    """q[S,H,N,D], k/v[S,Hk,KV,D] (dequantized f32), mask[1,1,Nm,KV] f16 or None -> [S,N,H,D]"""
    S, H, N, D = q.shape; Hk = k.shape[1]; g = H // Hk
    out = np.zeros((S, N, H, D))
    for s in range(S):
        for h in range(H):
            sc = q[s, h].astype(np.float64) @ k[s, h // g].astype(np.float64).T * scale
            if mask is not None:
                sc = sc + mask[0, 0, :N].astype(np.float64)
            m = sc.max(axis=1, keepdims=True)
            m[~np.isfinite(m)] = 0.0
            p = np.exp(sc - m)
            den = p.sum(axis=1, keepdims=True)
            den[den == 0] = 1.0
            out[s, :, h] = (p / den) @ v[s, h // g].astype(np.float64)
    return out


def make_mask(rng, nm, kv, frac=0.1):
    m = rng.uniform(-1, 1, (1, 1, nm, kv)).astype(np.float32)
    holes = rng.uniform(size=(nm, (kv + 15) // 16)) < frac
    m[0, 0][np.repeat(holes, 16, axis=1)[:, :kv]] = -np.inf
    return m.astype(np.float16)


@pytest.mark.parametrize("t", ["q8_0", "q4_0"])
@pytest.mark.parametrize("hs", [64, 128])
@pytest.mark.parametrize("nr2", [1, 4, 8])
@pytest.mark.parametrize("kv", [113, 512, 1400])
@pytest.mark.parametrize("nb", [1, 3, 35])
def test_flash_attn_quantized_kv(ctx, B, orc, t, hs, nr2, kv, nb):
    """type x head size x GQA ratio x keys (113: ragged; 1400: several KV splits) x batch (1, 3: the in-kernel dequantization; 35: the
    f16 expansion in front of the prefill kernels), mask with -inf runs"""
    T = {"q8_0": O.Q8_0, "q4_0": O.Q4_0}[t]
    rng = np.random.default_rng(hs + nr2 * 7 + kv + nb)
    Hk = 2; H = Hk * nr2; S = 1
    q = rng.uniform(-1, 1, (S, H, nb, hs)).astype(np.float32)
    kf = rng.uniform(-1, 1, (S, Hk, kv, hs)).astype(np.float32)
    vf = rng.uniform(-1, 1, (S, Hk, kv, hs)).astype(np.float32)
    kq, vq = quant_rows(orc, T, kf), quant_rows(orc, T, vf)
    mask = make_mask(rng, nb, kv)
    sc = 1.0 / np.sqrt(hs)
    got = ctx.flash_attn_ext_host(q, kq, vq, mask, None, sc, kv_type=T, kv_shape=(S, Hk, kv, hs, hs))
    assert np.isfinite(got).all()
    kd = orc.dequantize(T, kq, kf.size).reshape(kf.shape); vd = orc.dequantize(T, vq, vf.size).reshape(vf.shape)
    want = attn_f64(q, kd, vd, mask, sc)
    assert O.nmse(want, got) < 5e-4
    if O.have_reference():
        ref = O.Reference()
        assert O.nmse(ref.flash_attn_ext(q, kq, vq, mask, None, sc, kv_type=T, kv_shape=(S, Hk, kv, hs, hs)), got) < 5e-4


@pytest.mark.parametrize("tk,tv", [("q8_0", "q4_0"), ("q4_0", "q8_0")])
def test_flash_attn_mixed_kv_types(ctx, B, orc, tk, tv):
    """-ctk and -ctv are independent switches of llama.cpp"""
    TK, TV = {"q8_0": O.Q8_0, "q4_0": O.Q4_0}[tk], {"q8_0": O.Q8_0, "q4_0": O.Q4_0}[tv]
    rng = np.random.default_rng(5)
    S, Hk, H, kv, hs, nb = 1, 2, 8, 300, 128, 2
    q = rng.uniform(-1, 1, (S, H, nb, hs)).astype(np.float32)
    kf = rng.uniform(-1, 1, (S, Hk, kv, hs)).astype(np.float32); vf = rng.uniform(-1, 1, (S, Hk, kv, hs)).astype(np.float32)
    kq, vq = quant_rows(orc, TK, kf), quant_rows(orc, TV, vf)
    got = ctx.flash_attn_ext_host(q, kq, vq, None, None, 0.09, kv_type=(TK, TV), kv_shape=(S, Hk, kv, hs, hs))
    kd = orc.dequantize(TK, kq, kf.size).reshape(kf.shape); vd = orc.dequantize(TV, vq, vf.size).reshape(vf.shape)
    assert O.nmse(attn_f64(q, kd, vd, None, 0.09), got) < 5e-4


@pytest.mark.parametrize("t", ["q8_0", "q4_0"])
def test_flash_attn_quantized_cache_views_decode(ctx, B, orc, t):
    """decode over the cache layout llama.cpp hands over with -ctk / -ctv: one [n_embd_gqa, kv_size] buffer per layer, heads interleaved
    inside a token row, K / V as strided views of the first n_kv rows (src/llama-kv-cache.cpp:1008-1058); 32 query heads over 8 kv
    heads of 128 at 4096 keys, sinks and ALiBi slopes on, vs the compiled reference."""
    T = {"q8_0": O.Q8_0, "q4_0": O.Q4_0}[t]
    rng = np.random.default_rng(11)
    H, Hk, D, kv_size, n_kv = 32, 8, 128, 4352, 4096
    ck = rng.standard_normal((kv_size, Hk, D)).astype(np.float32); cv = rng.standard_normal((kv_size, Hk, D)).astype(np.float32)
    ckq, cvq = quant_rows(orc, T, ck), quant_rows(orc, T, cv)
    q = rng.standard_normal((1, H, 1, D)).astype(np.float32)
    mask = np.zeros((1, 1, 1, n_kv), np.float16); mask[..., n_kv - 37:] = -np.inf          # the padded tail of the view
    sinks = rng.uniform(-1, 1, H).astype(np.float32)
    sc = 1.0 / np.sqrt(D)
    rowb = O.row_size(T, D)
    dq = ctx.upload(B.F32, [D, 1, H, 1], q)
    dck = ctx.upload(T, [D * Hk, kv_size], ckq); dcv = ctx.upload(T, [D * Hk, kv_size], cvq)
    bs = B.TSIZE[T]
    kview = dck.view([D, n_kv, Hk, 1], [bs, rowb * Hk, rowb, rowb * Hk * kv_size])
    vview = dcv.view([D, n_kv, Hk, 1], [bs, rowb * Hk, rowb, rowb * Hk * kv_size])
    dm = ctx.upload(B.F16, [n_kv, 1, 1, 1], mask)
    ds = ctx.upload(B.F32, [H], sinks)
    dd = ctx.empty(B.F32, [D, H, 1, 1])
    ctx.flash_attn_ext(dq, kview, vview, dm, ds, dd, sc, 8.0)
    got = ctx.download(dd).reshape(1, 1, H, D)
    assert np.isfinite(got).all()
    if not O.have_reference():
        pytest.skip("oracle/_ref not built")
    ref = O.Reference()
    kc = np.ascontiguousarray(ckq.reshape(kv_size, Hk, rowb)[:n_kv].transpose(1, 0, 2)); vc = np.ascontiguousarray(cvq.reshape(kv_size, Hk, rowb)[:n_kv].transpose(1, 0, 2))
    want = ref.flash_attn_ext(q, kc, vc, mask, sinks, sc, 8.0, kv_type=T, kv_shape=(1, Hk, n_kv, D, D))
    assert O.nmse(want, got) < 5e-4
    for d in (dq, dck, dcv, dm, ds, dd):
        d.free()


@pytest.mark.parametrize("t", ["q8_0", "q4_0"])
def test_set_rows_quantized_cache(ctx, B, orc, t):
    """the cache write behind -ctk / -ctv (GGML_OP_SET_ROWS f32 -> q8_0 / q4_0) lives in the plugin's glue layer and is exercised by the
    reference's test-backend-ops (tests/test_gpu_backend_ops.py); here: what the flat API sees of it -- blocks written by the reference's
    row quantizers attend identically whether they were produced on the host (above) or are read back: bytes round-trip through upload."""
    T = {"q8_0": O.Q8_0, "q4_0": O.Q4_0}[t]
    rng = np.random.default_rng(3)
    x = rng.standard_normal((7, 256)).astype(np.float32)
    raw = quant_rows(orc, T, x)
    d = ctx.upload(T, [256, 7], raw)
    assert np.array_equal(ctx.download(d, np.uint8), raw)
    d.free()
