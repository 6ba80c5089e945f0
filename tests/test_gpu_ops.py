"""GPU parity tests (-m gpu): the CUDA hot path, called through the C-ABI (include/ggml-b200.h via ctypes),
against the CPU oracle on the same seeded inputs, and against the reference's committed golden vectors.

Bars (BASELINE.json north_star):
  * q8_1 activation quantization: BIT-EXACT with the reference's quantize_row_q8_1 (both rounding variants);
  * MUL_MAT / MUL_MAT_ID / FLASH_ATTN_EXT fp32 outputs: NMSE < 5e-4 against the ggml-cpu result
    (tests/test-backend-ops.cpp:3716-3718,3852-3854,6080-6082).  Because NMSE 5e-4 would hide an indexing bug that
    hits a few elements, each case is ALSO compared with the oracle run in the kernel's own numerics
    (q8_1 integer dots / fp32 attention) at NMSE < 1e-9 / 1e-6.
"""
import os
import numpy as np
import pytest

import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NMSE_HARNESS = 5e-4
QTYPES = [O.Q4_0, O.Q8_0, O.Q4_K, O.Q5_K, O.Q6_K]


@pytest.fixture(scope="module")
def B():
    import ggml_b200
    return ggml_b200


@pytest.fixture(scope="module")
def ctx(B):
    c = B.Context(0)          # raises without a B200: no fallback
    yield c
    c.close()


@pytest.fixture(scope="module")
def orc():
    return O.Oracle()


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "hotpath_golden.npz"))


# ---- q8_1 ------------------------------------------------------------------------------------------
def test_q8_1_bit_exact_golden(ctx, gold):
    x = gold["q8_1_x"]
    assert np.array_equal(ctx.quantize_q8_1_host(x, 0), gold["q8_1_x86"])
    assert np.array_equal(ctx.quantize_q8_1_host(x, 1), gold["q8_1_ref"])
    xn = gold["q8_1_nan_x"]                                         # NaN / inf at every lane position: same amax fold order as the reference
    assert np.array_equal(ctx.quantize_q8_1_host(xn, 0), gold["q8_1_nan_x86"])
    assert np.array_equal(ctx.quantize_q8_1_host(xn, 1), gold["q8_1_nan_ref"])


@pytest.mark.parametrize("variant", [0, 1])
def test_q8_1_bit_exact_1e6_blocks(ctx, orc, variant):
    rng = np.random.default_rng(100 + variant)
    rows, k = 2048, 16384                                         # 1,048,576 blocks
    x = (rng.standard_normal((rows, k), dtype=np.float32) * (10.0 ** rng.integers(-5, 5, (rows, 1))).astype(np.float32))
    x[::5, :64] = np.round(x[::5, :64] * 2) / 2                   # exact .5 ties
    x[1, :32] = 0.0
    x[2, :32] = 1e-40                                             # denormal amax -> id = inf on x86
    x[3, :32] = 3.0e38
    x[4, :k] = np.float32(-1.0)
    got = ctx.quantize_q8_1_host(x, variant)
    want = orc.quantize_q8_1(x, variant)
    bad = np.nonzero((got.reshape(-1, 36) != want.reshape(-1, 36)).any(1))[0]
    assert bad.size == 0, f"{bad.size} blocks differ, first {bad[:5]}"


def test_q8_1_strided_rows_and_empty(ctx, B, orc):
    rng = np.random.default_rng(7)
    x = rng.uniform(-1, 1, (6, 512)).astype(np.float32)
    dx = ctx.upload(B.F32, [512, 6], x)
    dy = ctx.empty(B.Q8_1, [256, 3])
    ctx.quantize_q8_1(dx.ptr + 256 * 4, dy.ptr, 256, 3, row_stride=2 * 512 * 4)      # second half of rows 0, 2, 4
    got = ctx.download(dy, np.uint8)
    assert np.array_equal(got, orc.quantize_q8_1(np.ascontiguousarray(x[::2, 256:]), 0))
    ctx.quantize_q8_1(dx.ptr, dy.ptr, 256, 0)                                          # zero rows: no-op
    with pytest.raises(B.B200Error):
        ctx.quantize_q8_1(dx.ptr, dy.ptr, 100, 1)                                      # k % 32 != 0


@pytest.mark.parametrize("K", [1024, 4096, 14336])
@pytest.mark.parametrize("mma", [False, True])
def test_q8_1_fused_in_gemv_bit_exact_probe(ctx, B, orc, K, mma, monkeypatch):
    """The decode GEMV quantizes its activation column itself (b200_quantize_column_x86).  Probe weights read the quantized
    column back through MUL_MAT exactly: a q8_0 row holding a single 1 (d = 1) returns d_a * q_a of one element, a q4_0 row of
    zero nibbles with d = 1 on one block returns -8 * s of that block; both must equal the reference's quantize_row_q8_1
    bit for bit, special blocks (zeros, denormal amax, ties, NaN elements) included."""
    rng = np.random.default_rng(K)
    x = (rng.standard_normal(K) * 10.0 ** rng.integers(-4, 4, K // 32).repeat(32)).astype(np.float32)
    x[:32] = 0.0
    x[32:64] = 1e-40                                               # denormal amax: id = inf, quants become the x86 "integer indefinite"
    x[64:96] = 1000.0
    x[96:128] = np.round(x[96:128] * 2) / 2 + 0.5                  # exact .5 ties
    x[128:160] = -1.0
    x[192:224] = rng.uniform(-1e-3, 1e-3, 32)                      # (keeps d * 2^31 inside fp16: a finite s)
    x[200] = np.nan                                                # its quant is the saturated x86 "integer indefinite", the int32 sum wraps
    x[224:256] = rng.uniform(-1e-3, 1e-3, 32); x[231] = np.nan; x[250] = np.nan
    q = orc.quantize_q8_1(x[None, :], 0).reshape(-1, 36)
    d = q[:, 0:2].copy().view(np.float16).astype(np.float32).reshape(-1)
    sm = q[:, 2:4].copy().view(np.float16).astype(np.float32).reshape(-1)
    qa = q[:, 4:].view(np.int8).astype(np.float32)
    one = np.float16(1.0).view(np.uint16)
    dx = ctx.upload(B.F32, [K, 1], x)
    with np.errstate(invalid="ignore", over="ignore"):
        want_dq = (d[:, None] * qa).reshape(-1)                        # element probe
        want_s = np.float32(-8.0) * sm                                  # block probe
    # q8_0 element probe: row r = weight 1 at position r
    w8 = np.zeros((K, K // 32, 34), dtype=np.uint8)
    r = np.arange(K)
    w8[r, r // 32, 0:2] = np.array([one], dtype=np.uint16).view(np.uint8)
    w8[r, r // 32, 2 + r % 32] = 1
    w4 = np.zeros((K // 32, K // 32, 18), dtype=np.uint8)              # q4_0 block probe: row b = d = 1 on block b, nibbles 0 (value -8)
    b = np.arange(K // 32)
    w4[b, b, 0:2] = np.array([one], dtype=np.uint16).view(np.uint8)
    for t, w, M, want in ((O.Q8_0, w8, K, want_dq), (O.Q4_0, w4, K // 32, want_s)):
        dw = ctx.upload(t, [K, M], w.reshape(-1)); ctx.repack(dw)
        dd = ctx.empty(B.F32, [M, 1])
        ctx.mul_mat(dw, dx, dd)
        got = ctx.download(dd).reshape(-1)
        assert np.isfinite(want).all()                                 # (an inf scale would poison every row through 0 * inf)
        assert np.array_equal(got, want), (t, np.nonzero(got != want)[0][:8], got[got != want][:4], want[got != want][:4])
    if mma:                                                            # the tensor-core GEMV has its own copy of the fused quantizer: q4_K probe
        M = 64
        w = O.random_blocks(rng, O.Q4_K, K, M)
        xs = np.where(np.isfinite(x), x, 0.0).astype(np.float32); xs[32:96] = 0.25
        dxs = ctx.upload(B.F32, [K, 1], xs)
        dw = ctx.upload(O.Q4_K, [K, M], w); dd = ctx.empty(B.F32, [M, 1])
        ctx.mul_mat(dw, dxs, dd)
        got = ctx.download(dd).reshape(1, 1, 1, M)
        assert O.nmse(orc.mul_mat(O.Q4_K, w, K, M, xs[None, :], 1, act=O.ACT_Q8_1), got) < 1e-9


# ---- MUL_MAT ------------------------------------------------------------------------------------------
def _check_mm(orc, got, t, w, K, M, x, N, want_cpu=None, **bc):
    exact = orc.mul_mat(t, w, K, M, x, N, act=O.ACT_Q8_1, **bc)
    cpu = want_cpu if want_cpu is not None else orc.mul_mat(t, w, K, M, x, N, act=O.ACT_CPU, **bc)
    assert np.isfinite(got).all()
    e1, e2 = O.nmse(exact, got), O.nmse(cpu, got)
    assert e2 < NMSE_HARNESS, f"vs ggml-cpu numerics: nmse {e2:.3e}"
    return e1


@pytest.mark.parametrize("t", QTYPES)
def test_mul_mat_golden(ctx, orc, gold, t):
    name = O.TYPE_NAMES[t]
    x = gold["mm_x"]; N, K = x.shape
    want = gold[f"mm_{name}"]; M = want.shape[-1]
    got = ctx.mul_mat_host(t, gold[f"w_{name}"], K, M, x, N)
    e = _check_mm(orc, got, t, gold[f"w_{name}"], K, M, x, N, want_cpu=want)
    assert e < 1e-9, f"vs q8_1-path oracle: nmse {e:.3e}"


GEMM_TYPES = (O.Q4_K,)          # types the tcgen05 prefill GEMM takes (bf16 operands, fp32 TMEM accumulation) when n > 8
NMSE_BF16 = 5e-5                # bf16 rounding of weights and activations vs exact-activation arithmetic


def _check_prefill(orc, got, t, w, K, M, x, N):
    assert np.isfinite(got).all()
    e_cpu = O.nmse(orc.mul_mat(t, w, K, M, x, N, act=O.ACT_CPU), got)
    assert e_cpu < NMSE_HARNESS, f"vs ggml-cpu numerics: nmse {e_cpu:.3e}"
    e = O.nmse(orc.mul_mat(t, w, K, M, x, N, act=O.ACT_F64), got)
    assert e < NMSE_BF16, f"vs exact-activation oracle: nmse {e:.3e}"


@pytest.mark.parametrize("t", GEMM_TYPES)
@pytest.mark.parametrize("M,N,K", [(128, 16, 256), (16, 9, 256), (200, 33, 512), (129, 64, 1024), (300, 260, 1024), (1024, 512, 4096),
                                   (4096, 100, 256), (256, 128, 14336)])
def test_mul_mat_prefill_gemm(ctx, orc, t, M, N, K):
    """n > 8: tcgen05 GEMM; ragged M (not a multiple of 128), ragged N (TMA zero fill), every BN tile width, long K"""
    rng = np.random.default_rng(M + 3 * N + K)
    w = O.random_blocks(rng, t, K, M)
    x = rng.standard_normal((N, K)).astype(np.float32)
    got = ctx.mul_mat_host(t, w, K, M, x, N)
    _check_prefill(orc, got, t, w, K, M, x, N)


@pytest.mark.parametrize("M,N,K", [(128, 16, 2048), (200, 33, 2048), (300, 260, 4096), (1024, 512, 4096), (256, 128, 14336)])
def test_mul_mat_prefill_gemm_q6_K_repacked(ctx, B, orc, M, N, K):
    """q6_K weights in the load-time group layout (b200_repack) through the tcgen05 GEMM: 4-D TMA box per 256-K block"""
    rng = np.random.default_rng(M + 5 * N + K)
    w = O.random_blocks(rng, O.Q6_K, K, M)
    x = rng.standard_normal((N, K)).astype(np.float32)
    dw = ctx.upload(B.Q6_K, [K, M], w); ctx.repack(dw)
    dx = ctx.upload(B.F32, [K, N], x); dd = ctx.empty(B.F32, [M, N])
    ctx.mul_mat(dw, dx, dd)
    _check_prefill(orc, ctx.download(dd).reshape(1, 1, N, M), O.Q6_K, w, K, M, x, N)


@pytest.mark.parametrize("t", GEMM_TYPES)
def test_mul_mat_prefill_gemm_strided(ctx, B, orc, t):
    """activation rows and result columns with padding between them (views into larger buffers)"""
    rng = np.random.default_rng(77)
    K, M, N = 512, 192, 40
    w = O.random_blocks(rng, t, K, M)
    xbuf = rng.standard_normal((N, K + 64)).astype(np.float32)
    dx = ctx.upload(B.F32, [K + 64, N], xbuf).view([K, N, 1, 1], [4, (K + 64) * 4, (K + 64) * 4 * N, (K + 64) * 4 * N])
    dw = ctx.upload(t, [K, M], w)
    dbuf = ctx.upload(B.F32, [M + 32, N], np.full((N, M + 32), 7.0, np.float32))
    dd = dbuf.view([M, N, 1, 1], [4, (M + 32) * 4, (M + 32) * 4 * N, (M + 32) * 4 * N])
    ctx.mul_mat(dw, dx, dd)
    out = ctx.download(dbuf).reshape(N, M + 32)
    assert (out[:, M:] == 7.0).all()
    _check_prefill(orc, out[:, :M].reshape(1, 1, N, M), t, w, K, M, np.ascontiguousarray(xbuf[:, :K]), N)


@pytest.mark.parametrize("t", QTYPES)
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 6, 7, 8, 9, 16])
def test_mul_mat_small_n(ctx, orc, t, n):
    """m=16, n=1..9, k=256 of test-backend-ops (tests/test-backend-ops.cpp:7616-7620), random bit patterns."""
    rng = np.random.default_rng(1000 * t + n)
    K, M = 256, 16
    w = O.random_blocks(rng, t, K, M)
    x = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    got = ctx.mul_mat_host(t, w, K, M, x, n)
    if n > 8 and t in GEMM_TYPES:
        return _check_prefill(orc, got, t, w, K, M, x, n)
    assert _check_mm(orc, got, t, w, K, M, x, n) < 1e-9


@pytest.mark.parametrize("t", [O.Q4_0, O.Q4_K, O.Q8_0])
@pytest.mark.parametrize("bs,nr", [((3, 2), (1, 1)), ((1, 1), (2, 2)), ((3, 1), (2, 1)), ((2, 2), (1, 2))])
def test_mul_mat_batch_broadcast(ctx, orc, t, bs, nr):
    rng = np.random.default_rng(17)
    K, M, N = 256, 16, 3
    ne02, ne03 = bs; ne12, ne13 = bs[0] * nr[0], bs[1] * nr[1]
    w = O.random_blocks(rng, t, K, M * ne02 * ne03)
    x = rng.uniform(-1, 1, (ne13, ne12, N, K)).astype(np.float32)
    got = ctx.mul_mat_host(t, w, K, M, x, N, ne02, ne03, ne12, ne13)
    assert _check_mm(orc, got, t, w, K, M, x, N, ne02=ne02, ne03=ne03, ne12=ne12, ne13=ne13) < 1e-9


@pytest.mark.parametrize("t", [O.Q4_K, O.Q8_0])
def test_mul_mat_permuted_activations(ctx, B, orc, t):
    """src1 permuted (0,2,1,3) as test-backend-ops does (:7636-7684): rows contiguous, dims 1 and 2 swapped in memory."""
    rng = np.random.default_rng(23)
    K, M, N, ne12 = 256, 16, 4, 3
    w = O.random_blocks(rng, t, K, M)
    xmem = rng.uniform(-1, 1, (N, ne12, K)).astype(np.float32)               # memory order [n][i12][k]
    dx = ctx.upload(B.F32, [K, ne12, N], xmem)
    xv = dx.view([K, N, ne12, 1], [4, K * 4 * ne12, K * 4, K * 4 * ne12 * N])  # logical [K, N, ne12]
    dw = ctx.upload(t, [K, M], w)
    dd = ctx.empty(B.F32, [M, N, ne12])
    ctx.mul_mat(dw, xv, dd)
    got = ctx.download(dd).reshape(1, ne12, N, M)
    xlog = np.ascontiguousarray(xmem.transpose(1, 0, 2))                      # [i12][n][k]
    assert _check_mm(orc, got, t, w, K, M, xlog, N, ne12=ne12) < 1e-9


@pytest.mark.parametrize("t", QTYPES)
def test_mul_mat_decode_llama8b_ffn_down(ctx, orc, t):
    """the BASELINE.json decode shape: m=4096, n=1, k=14336 (test-backend-ops perf case :8410-8416)"""
    rng = np.random.default_rng(31 + t)
    K, M = 14336, 4096
    w = O.random_blocks(rng, t, K, M)
    x = rng.standard_normal((1, K)).astype(np.float32)
    got = ctx.mul_mat_host(t, w, K, M, x, 1)
    assert _check_mm(orc, got, t, w, K, M, x, 1) < 1e-9


@pytest.mark.parametrize("t", QTYPES)
@pytest.mark.parametrize("M,K", [(1, 256), (3, 512), (5, 2304), (37, 4096), (130, 2560), (1000, 256), (4099, 1024)])
def test_mul_mat_decode_ragged_rows(ctx, orc, t, M, K):
    """one activation column (the streaming GEMV): odd row counts, a k-chunk tail (9, 10 blocks), more rows than warps"""
    rng = np.random.default_rng(M * 7 + K)
    w = O.random_blocks(rng, t, K, M)
    x = rng.standard_normal((1, K)).astype(np.float32)
    got = ctx.mul_mat_host(t, w, K, M, x, 1)
    assert _check_mm(orc, got, t, w, K, M, x, 1) < 1e-9


def test_mul_mat_decode_q4_K_tensor_core_gemv(ctx, B, orc):
    """raw q4_K weights of >= 48 MB take the mma.sync GEMV (b200_gemv_mma.cu) by default: a 4096 x 24580 matrix (ragged last
    8-row group) alone, and as one of two segments sharing the column"""
    rng = np.random.default_rng(4242)
    K, M = 4096, 24580
    w = O.random_blocks(rng, O.Q4_K, K, M)
    x = rng.standard_normal((1, K)).astype(np.float32)
    dw = ctx.upload(O.Q4_K, [K, M], w); dx = ctx.upload(B.F32, [K, 1], x); dd = ctx.empty(B.F32, [M, 1])
    ctx.mul_mat(dw, dx, dd)
    want = orc.mul_mat(O.Q4_K, w, K, M, x, 1, act=O.ACT_Q8_1)
    assert O.nmse(want, ctx.download(dd)) < 1e-9
    assert O.nmse(orc.mul_mat(O.Q4_K, w, K, M, x, 1, act=O.ACT_CPU), ctx.download(dd)) < NMSE_HARNESS
    w2 = O.random_blocks(rng, O.Q4_K, K, 1027)
    dw2 = ctx.upload(O.Q4_K, [K, 1027], w2); dd2 = ctx.empty(B.F32, [1027, 1])
    ctx.h2d(dd.ptr, np.zeros(M, np.float32))
    ctx.mul_mat_multi([dw2, dw], dx, [dd2, dd])
    assert O.nmse(want, ctx.download(dd)) < 1e-9
    assert O.nmse(orc.mul_mat(O.Q4_K, w2, K, 1027, x, 1, act=O.ACT_Q8_1), ctx.download(dd2)) < 1e-9


def test_mul_mat_decode_tensor_core_gemv_all_shapes():
    """the decode tests again with GGML_B200_GEMV_MMA=1, which sends every supported raw q4_K shape (ragged rows, odd block counts,
    fused segments, K = 14336 through the q8_1 image) to the mma.sync GEMV instead of only the large ones"""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, GGML_B200_GEMV_MMA="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_gpu_ops.py"), "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider",
                        "-k", "(decode or multi_shared or golden or small_n or shared_activation) and not all_shapes"], env=env, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.parametrize("t", [O.Q8_0, O.Q4_0])
@pytest.mark.parametrize("K,M", [(256, 5), (2304, 67), (4096, 131), (14336, 64)])
def test_repack_run8_roundtrip_and_parity(ctx, B, orc, t, K, M):
    """q8_0 / q4_0 load-time repack ([8 x d][8 x qs] per 256 weights): exactly invertible, and the streaming GEMV (n=1), the
    column-blocked GEMV (n=3), MUL_MAT_ID and the tcgen05 GEMM (n=24) all read it"""
    rng = np.random.default_rng(K + M + t)
    w = O.random_blocks(rng, t, K, M)
    dw = ctx.upload(t, [K, M], w)
    ctx.repack(dw)
    assert dw.flags == 1
    packed = ctx.download(dw, np.uint8)
    assert not np.array_equal(packed, w) and np.array_equal(np.sort(packed.reshape(M, -1), axis=1), np.sort(w.reshape(M, -1), axis=1))
    for n in (1, 3, 24):
        x = rng.standard_normal((n, K)).astype(np.float32)
        dx = ctx.upload(B.F32, [K, n], x); dd = ctx.empty(B.F32, [M, n])
        ctx.mul_mat(dw, dx, dd)
        got = ctx.download(dd).reshape(1, 1, n, M)
        if n <= 8:
            assert O.nmse(orc.mul_mat(t, w, K, M, x, n, act=O.ACT_CPU), got) < NMSE_HARNESS
            assert O.nmse(orc.mul_mat(t, w, K, M, x, n, act=O.ACT_Q8_1), got) < 1e-9
        else:
            _check_prefill(orc, got, t, w, K, M, x, n)
    ctx.unrepack(dw)
    assert dw.flags == 0 and np.array_equal(ctx.download(dw, np.uint8), w)


@pytest.mark.parametrize("t", [O.Q8_0, O.Q4_0])
def test_mul_mat_prefill_gemm_run8(ctx, B, orc, t):
    rng = np.random.default_rng(5 + t)
    K, M, N = 4096, 1024, 512
    w = O.random_blocks(rng, t, K, M)
    x = rng.standard_normal((N, K)).astype(np.float32)
    dw = ctx.upload(t, [K, M], w); ctx.repack(dw)
    dx = ctx.upload(B.F32, [K, N], x); dd = ctx.empty(B.F32, [M, N])
    ctx.mul_mat(dw, dx, dd)
    _check_prefill(orc, ctx.download(dd).reshape(1, 1, N, M), t, w, K, M, x, N)


@pytest.mark.parametrize("K,M", [(2048, 67), (4096, 131), (14336, 64)])
def test_repack_q6_K_roundtrip_and_parity(ctx, B, orc, K, M):
    """load-time repack (b200_repack): in place, exactly invertible, and every MUL_MAT path reads the repacked layout"""
    rng = np.random.default_rng(K + M)
    w = O.random_blocks(rng, O.Q6_K, K, M)
    dw = ctx.upload(B.Q6_K, [K, M], w)
    ctx.repack(dw)
    assert dw.flags == 1
    packed = ctx.download(dw, np.uint8)
    assert not np.array_equal(packed, w) and np.array_equal(np.sort(packed.reshape(M, -1), axis=1), np.sort(w.reshape(M, -1), axis=1))
    for n in (1, 3, 12):                                         # streaming GEMV, column-blocked GEMV, prefill path
        x = rng.standard_normal((n, K)).astype(np.float32)
        dx = ctx.upload(B.F32, [K, n], x); dd = ctx.empty(B.F32, [M, n])
        ctx.mul_mat(dw, dx, dd)
        got = ctx.download(dd).reshape(1, 1, n, M)
        assert O.nmse(orc.mul_mat(O.Q6_K, w, K, M, x, n, act=O.ACT_CPU), got) < NMSE_HARNESS
        if n <= 8:
            assert O.nmse(orc.mul_mat(O.Q6_K, w, K, M, x, n, act=O.ACT_Q8_1), got) < 1e-9
    ctx.unrepack(dw)
    assert np.array_equal(ctx.download(dw, np.uint8), w)
    assert B.lib().b200_repack_supported(B.Q6_K, 256) == 0 and B.lib().b200_repack_supported(B.Q4_K, 4096) == 0


@pytest.mark.parametrize("t", [O.Q4_K, O.Q6_K])
def test_mul_mat_multi_shared_src1(ctx, B, orc, t):
    """wq/wk/wv in one launch (b200_mul_mat_multi): identical to three separate MUL_MATs"""
    rng = np.random.default_rng(5 + t)
    K, Ms = 4096, (4096, 1024, 1027)
    ws = [O.random_blocks(rng, t, K, M) for M in Ms]
    x = rng.standard_normal((1, K)).astype(np.float32)
    dws = [ctx.repack(ctx.upload(t, [K, M], w)) for w, M in zip(ws, Ms)]
    dx = ctx.upload(B.F32, [K, 1], x)
    dds = [ctx.empty(B.F32, [M, 1]) for M in Ms]
    n0 = ctx.launches
    ctx.mul_mat_multi(dws, dx, dds)
    assert ctx.launches - n0 == 1                                   # one fused launch, no separate quantize
    for w, M, dd in zip(ws, Ms, dds):
        assert O.nmse(orc.mul_mat(t, w, K, M, x, 1, act=O.ACT_Q8_1), ctx.download(dd)) < 1e-9
    # a mixed-type group falls back to per-tensor calls with the same results
    wq8 = O.random_blocks(rng, O.Q8_0, K, 64)
    dq8 = ctx.upload(B.Q8_0, [K, 64], wq8); d8 = ctx.empty(B.F32, [64, 1])
    ctx.mul_mat_multi([dws[0], dq8], dx, [dds[0], d8])
    assert O.nmse(orc.mul_mat(O.Q8_0, wq8, K, 64, x, 1, act=O.ACT_Q8_1), ctx.download(d8)) < 1e-9


@pytest.mark.parametrize("K,M", [(4096, 14336), (2048, 5632), (4096, 1027), (8192, 28672)])
def test_mul_mat_gate_up_swiglu_one_launch(ctx, B, orc, K, M):
    """ffn_gate + ffn_up + GLU(SWIGLU) in one launch (b200_mul_mat_gate_up_swiglu, tensor-core q4_K GEMV with the paired work split):
    gate / up identical to b200_mul_mat_multi, glu == x / (1 + exp(-x)) * y of those values"""
    rng = np.random.default_rng(91 + M)
    wg, wu = O.random_blocks(rng, O.Q4_K, K, M), O.random_blocks(rng, O.Q4_K, K, M)
    x = rng.standard_normal((1, K)).astype(np.float32)
    dg_w, du_w = ctx.upload(O.Q4_K, [K, M], wg), ctx.upload(O.Q4_K, [K, M], wu)
    dx = ctx.upload(B.F32, [K, 1], x)
    g1, u1, l1 = ctx.empty(B.F32, [M, 1]), ctx.empty(B.F32, [M, 1]), ctx.empty(B.F32, [M, 1])
    g2, u2 = ctx.empty(B.F32, [M, 1]), ctx.empty(B.F32, [M, 1])
    n0 = ctx.launches
    ctx.mul_mat_gate_up_swiglu(dg_w, du_w, dx, g1, u1, l1)
    assert ctx.launches - n0 == 1
    ctx.mul_mat_multi([dg_w, du_w], dx, [g2, u2])
    g, u, l = ctx.download(g1).ravel(), ctx.download(u1).ravel(), ctx.download(l1).ravel()
    assert np.array_equal(g, ctx.download(g2).ravel()) and np.array_equal(u, ctx.download(u2).ravel())
    assert O.nmse(orc.mul_mat(O.Q4_K, wg, K, M, x, 1, act=O.ACT_Q8_1), g.reshape(1, M)) < 1e-9
    want = (g / (np.float32(1) + np.exp(-g, dtype=np.float32)) * u).astype(np.float32)
    assert np.allclose(l, want, rtol=2e-6, atol=1e-6 * np.abs(want).max())


@pytest.mark.parametrize("t", [O.Q4_K, O.Q6_K])
def test_mul_mat_multi_prefill_gemm(ctx, B, orc, t):
    """wq/wk/wv (and gate/up) of a prefill ubatch in ONE tcgen05 GEMM launch: grid.y walks the row tiles of all weights"""
    rng = np.random.default_rng(15 + t)
    K, N, Ms = 2048, 96, (512, 128, 200)
    ws = [O.random_blocks(rng, t, K, M) for M in Ms]
    x = rng.standard_normal((N, K)).astype(np.float32)
    dws = [ctx.repack(ctx.upload(t, [K, M], w)) for w, M in zip(ws, Ms)]
    dx = ctx.upload(B.F32, [K, N], x)
    dds = [ctx.empty(B.F32, [M, N]) for M in Ms]
    k0 = ctx.launches_by_kind()["gemm"]
    ctx.mul_mat_multi(dws, dx, dds)
    assert ctx.launches_by_kind()["gemm"] - k0 == 1
    for w, M, dd in zip(ws, Ms, dds):
        _check_prefill(orc, ctx.download(dd).reshape(1, 1, N, M), t, w, K, M, x, N)


def test_mul_mat_shared_activation_cache(ctx, B, orc):
    """wq/wk/wv-style reuse of one activation tensor, then an overwrite of it through the API: the cached q8_1 image
    must be dropped (b200_ctx_invalidate semantics, include/ggml-b200.h)"""
    rng = np.random.default_rng(99)
    K, M = 1024, 64
    w1, w2 = O.random_blocks(rng, O.Q8_0, K, M), O.random_blocks(rng, O.Q8_0, K, M)
    x1 = rng.standard_normal((1, K)).astype(np.float32); x2 = rng.standard_normal((1, K)).astype(np.float32)
    dw1, dw2 = ctx.upload(B.Q8_0, [K, M], w1), ctx.upload(B.Q8_0, [K, M], w2)
    dx = ctx.upload(B.F32, [K, 1], x1); d1, d2 = ctx.empty(B.F32, [M, 1]), ctx.empty(B.F32, [M, 1])
    q0 = ctx.launches_by_kind()["quantize_q8_1"]
    ctx.mul_mat(dw1, dx, d1); ctx.mul_mat(dw2, dx, d2)
    assert ctx.launches_by_kind()["quantize_q8_1"] == q0 + 1          # quantized once for both
    assert O.nmse(orc.mul_mat(O.Q8_0, w2, K, M, x1, 1, act=O.ACT_Q8_1), ctx.download(d2)) < 1e-9
    ctx.h2d(dx.ptr, x2)                                               # overwrite the activations
    ctx.mul_mat(dw1, dx, d1)
    assert O.nmse(orc.mul_mat(O.Q8_0, w1, K, M, x2, 1, act=O.ACT_Q8_1), ctx.download(d1)) < 1e-9
    ctx.mul_mat(dw2, dx, dx.view([M, 1], [4, M * 4, M * 4, M * 4]))   # dst aliases the cached source: must invalidate
    ctx.mul_mat(dw1, dx, d1)
    x3 = ctx.download(dx).reshape(1, K)
    assert O.nmse(orc.mul_mat(O.Q8_0, w1, K, M, x3, 1, act=O.ACT_Q8_1), ctx.download(d1)) < 1e-9


@pytest.mark.parametrize("t", [O.Q4_0, O.Q4_K, O.Q8_0, O.Q6_K])
@pytest.mark.parametrize("n", [17, 64, 512])
def test_mul_mat_prefill(ctx, orc, t, n):
    rng = np.random.default_rng(41 + t + n)
    K, M = 2048, 384
    w = O.random_blocks(rng, t, K, M)
    x = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    got = ctx.mul_mat_host(t, w, K, M, x, n)
    cpu = orc.mul_mat(t, w, K, M, x, n, act=O.ACT_CPU)
    f64 = orc.mul_mat(t, w, K, M, x, n, act=O.ACT_F64)
    assert O.nmse(cpu, got) < NMSE_HARNESS
    assert O.nmse(f64, got) < 1e-4        # whichever activation format the prefill kernel uses, it is at least q8-accurate


def test_mul_mat_float_weights(ctx, orc):
    rng = np.random.default_rng(5)
    K, M, N = 320, 24, 5
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    for t in (O.F32, O.F16, O.BF16):
        w = O.random_blocks(rng, t, K, M, scale=1.0)
        got = ctx.mul_mat_host(t, w, K, M, x, N)
        assert O.nmse(orc.mul_mat(t, w, K, M, x, N, act=O.ACT_F64), got) < 1e-10


def test_mul_mat_rejects_bad_shapes(ctx, B):
    dw = ctx.empty(B.Q4_K, [256, 16]); dx = ctx.empty(B.F32, [512, 1]); dd = ctx.empty(B.F32, [16, 1])
    with pytest.raises(B.B200Error):
        ctx.mul_mat(dw, dx, dd)


# ---- MUL_MAT_ID ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("t", [O.Q4_0, O.Q4_K, O.Q8_0])
@pytest.mark.parametrize("ne11", [1, 2])
def test_mul_mat_id_golden(ctx, orc, gold, t, ne11):
    name = O.TYPE_NAMES[t]
    want = gold[f"id_out{ne11}_{name}"]; T, U, M = want.shape
    b = gold[f"id_b{ne11}_{name}"]; K = b.shape[-1]
    as_, ids = gold[f"id_as_{name}"], gold[f"id_ids_{name}"]
    got = ctx.mul_mat_id_host(t, as_, K, M, 8, b, ne11, T, ids, U)
    assert O.nmse(want, got) < NMSE_HARNESS
    assert O.nmse(orc.mul_mat_id(t, as_, K, M, 8, b, ne11, T, ids, U, act=O.ACT_Q8_1), got) < 1e-9


@pytest.mark.parametrize("t", [O.Q4_0, O.Q4_K, O.Q8_0])
@pytest.mark.parametrize("n_mats,n_used,bcast,n", [(4, 1, False, 1), (4, 2, True, 4), (8, 4, False, 5), (8, 2, True, 17),
                                                   (8, 4, True, 32), (8, 2, False, 129), (4, 4, False, 1)])
def test_mul_mat_id_shapes(ctx, orc, t, n_mats, n_used, bcast, n):
    """n_mats/n_used/broadcast/n grid of test-backend-ops (:7781-7795), m=512, k=256"""
    rng = np.random.default_rng(n_mats * 100 + n_used * 10 + n)
    K, M = 256, 512
    as_ = O.random_blocks(rng, t, K, M * n_mats)
    ids = np.stack([rng.permutation(n_mats)[:n_used] for _ in range(n)]).astype(np.int32)
    ne11 = 1 if bcast else n_used
    b = rng.uniform(-1, 1, (n, ne11, K)).astype(np.float32)
    got = ctx.mul_mat_id_host(t, as_, K, M, n_mats, b, ne11, n, ids, n_used)
    assert O.nmse(orc.mul_mat_id(t, as_, K, M, n_mats, b, ne11, n, ids, n_used, act=O.ACT_CPU), got) < NMSE_HARNESS
    if n <= 8:
        assert O.nmse(orc.mul_mat_id(t, as_, K, M, n_mats, b, ne11, n, ids, n_used, act=O.ACT_Q8_1), got) < 1e-9


@pytest.mark.parametrize("t,repack", [(O.Q4_K, False), (O.Q6_K, True), (O.Q8_0, True), (O.Q4_0, True)])
@pytest.mark.parametrize("bcast", [True, False])
def test_mul_mat_id_prefill_grouped_gemm(ctx, B, orc, t, repack, bcast):
    """Mixtral-style prefill: 8 experts, 2 used, 200 tokens -> device-side bucketing + ONE grouped tcgen05 GEMM launch; some
    experts get few rows, one gets none"""
    rng = np.random.default_rng(11 + t)
    K, M, E, U, n = (2048 if t == O.Q6_K else 1024), 300, 8, 2, 200
    as_ = O.random_blocks(rng, t, K, M * E)
    ids = np.stack([rng.permutation(E - 1)[:U] for _ in range(n)]).astype(np.int32)      # expert 7 is never routed to
    ids[:3] = [[0, 6], [6, 0], [0, 6]]
    ne11 = 1 if bcast else U
    b = rng.standard_normal((n, ne11, K)).astype(np.float32)
    da = ctx.upload(t, [K, M, E], as_)
    if repack:
        ctx.repack(da); assert da.flags == 1
    db = ctx.upload(B.F32, [K, ne11, n], b); di = ctx.upload(B.I32, [U, n], ids); dd = ctx.empty(B.F32, [M, U, n])
    ctx.mul_mat_id(da, db, di, dd)
    got = ctx.download(dd).reshape(n, U, M)
    assert np.isfinite(got).all()
    assert O.nmse(orc.mul_mat_id(t, as_, K, M, E, b, ne11, n, ids, U, act=O.ACT_CPU), got) < NMSE_HARNESS
    assert O.nmse(orc.mul_mat_id(t, as_, K, M, E, b, ne11, n, ids, U, act=O.ACT_F64), got) < NMSE_BF16


def test_mul_mat_id_mixtral_decode(ctx, orc):
    rng = np.random.default_rng(77)
    K, M, E, U = 4096, 1024, 8, 2                      # a 1024-row slice of Mixtral's [4096, 14336, 8] gate tensor
    as_ = O.random_blocks(rng, O.Q4_K, K, M * E)
    ids = np.int32([[5, 2]])
    b = rng.standard_normal((1, 1, K)).astype(np.float32)
    got = ctx.mul_mat_id_host(O.Q4_K, as_, K, M, E, b, 1, 1, ids, U)
    assert O.nmse(orc.mul_mat_id(O.Q4_K, as_, K, M, E, b, 1, 1, ids, U, act=O.ACT_Q8_1), got) < 1e-9
