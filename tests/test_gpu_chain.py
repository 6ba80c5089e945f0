"""GPU parity tests (-m gpu) of the persistent decode chain, b200_decode_chain (include/ggml-b200.h, csrc/b200_chain.cu), through the C-ABI.

Oracle: the same node sequence computed node by node on the host -- MUL_MAT by oracle/ggml_oracle.c (orc_mul_mat, restating
ggml_compute_forward_mul_mat, ggml-cpu.c:1226-1418), ADD / RMS_NORM / MUL / SWIGLU in numpy f32 as ggml-cpu/ops.cpp computes them
(rms_norm 3560-3640: scale = 1 / sqrt(mean(x^2) + eps); glu: silu(gate) * up).
Bars: NMSE < 5e-4 against ggml-cpu's numerics (q8_0 / q8_K activations), the reference's bar; < 1e-6 against the same pipeline in the
kernel's own numerics (q8_1 integer dots) -- not 1e-9, because the norm's summation order may move a q8_1 quant by one step.
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


@pytest.fixture(scope="module")
def orc():
    return O.Oracle()


def rms_norm(x, eps):
    x = x.astype(np.float32)
    mean = np.float32(np.sum(x.astype(np.float64) ** 2) / x.size)
    return (x * np.float32(1.0 / np.sqrt(mean + np.float32(eps)))).astype(np.float32)


def swiglu(g, u):
    g = g.astype(np.float32); u = u.astype(np.float32)
    return (g / (np.float32(1.0) + np.exp(-g)) * u).astype(np.float32)


def host_layer(orc, W, types, dims, attn_out, inp_sa, norm_ffn, norm_attn, eps, act):
    """the chain's nodes one by one on the host; W[name] = GGUF bytes"""
    n_embd, n_ff, n_kv = dims
    mm = lambda name, K, M, x: orc.mul_mat(types[name], W[name], K, M, x[None, :], 1, act=act).reshape(-1)
    r = {}
    r["wo_out"] = mm("wo", n_embd, n_embd, attn_out)
    r["ffn_inp"] = (r["wo_out"] + inp_sa).astype(np.float32)
    r["ffn_normed"] = rms_norm(r["ffn_inp"], eps)
    r["ffn_x"] = (r["ffn_normed"] * norm_ffn).astype(np.float32)
    r["gate"] = mm("gate", n_embd, n_ff, r["ffn_x"]); r["up"] = mm("up", n_embd, n_ff, r["ffn_x"])
    r["glu"] = swiglu(r["gate"], r["up"])
    r["down_out"] = mm("down", n_ff, n_embd, r["glu"])
    r["l_out"] = (r["down_out"] + r["ffn_inp"]).astype(np.float32)
    r["attn_normed"] = rms_norm(r["l_out"], eps)
    r["attn_x"] = (r["attn_normed"] * norm_attn).astype(np.float32)
    r["q"] = mm("q", n_embd, n_embd, r["attn_x"]); r["k"] = mm("k", n_embd, n_kv, r["attn_x"]); r["v"] = mm("v", n_embd, n_kv, r["attn_x"])
    return r


def make_layer(ctx, rng, dims, t_v, t_down, norm_outputs=True):
    from decode_chain import LayerChain
    n_embd, n_ff, n_kv = dims
    types = {"wo": O.Q4_K, "gate": O.Q4_K, "up": O.Q4_K, "down": t_down, "q": O.Q4_K, "k": O.Q4_K, "v": t_v}
    shape = {"wo": (n_embd, n_embd), "gate": (n_embd, n_ff), "up": (n_embd, n_ff), "down": (n_ff, n_embd), "q": (n_embd, n_embd), "k": (n_embd, n_kv), "v": (n_embd, n_kv)}
    sc = 0.004 * min(1.0, (2048.0 / n_embd)) ** 0.5                                   # activations stay O(1) through the chain, whatever its width
    W = {n: O.random_blocks(rng, types[n], *shape[n], scale=sc) for n in types}       # (a glu column of ~1e4 overflows the fp16 sum field of q8_1: NaN)
    norm_ffn = rng.uniform(0.5, 1.5, n_embd).astype(np.float32); norm_attn = rng.uniform(0.5, 1.5, n_embd).astype(np.float32)
    lc = LayerChain(ctx, n_embd, n_ff, n_kv, {n: (types[n], W[n]) for n in types}, norm_ffn, norm_attn, eps=1e-5, norm_outputs=norm_outputs)
    return lc, W, types, norm_ffn, norm_attn


@pytest.mark.parametrize("t_v,t_down", [(O.Q4_K, O.Q4_K), (O.Q6_K, O.Q6_K)])
def test_decode_chain_vs_oracle(ctx, B, orc, t_v, t_down):
    rng = np.random.default_rng(900 + t_v)
    dims = (2048, 6144, 512)                                  # n_embd, n_ff (3 K chunks), kv width: every phase spans > 148 units or fewer
    lc, W, types, norm_ffn, norm_attn = make_layer(ctx, rng, dims, t_v, t_down)
    assert ctx.decode_chain_supported(lc.program)
    for rep in range(2):                                      # the second launch finds the re-armed counters and other inputs
        attn_out = rng.uniform(-1, 1, dims[0]).astype(np.float32); inp_sa = rng.uniform(-2, 2, dims[0]).astype(np.float32)
        lc.set_inputs(attn_out, inp_sa)
        lc.clear_outputs()
        lc.run()
        assert ctx.decode_chain_error() == 0
        got = lc.outputs()
        exact = host_layer(orc, W, types, dims, attn_out, inp_sa, norm_ffn, norm_attn, 1e-5, O.ACT_Q8_1)
        for name in lc.OUTPUTS:
            assert np.isfinite(got[name]).all(), (rep, name)
            e = O.nmse(exact[name], got[name])
            assert e < 1e-6, f"rep {rep} {name}: nmse {e} vs the q8_1 pipeline"
        # the reference's bar, per MUL_MAT node on the chain's own input column (ggml-cpu numerics: q8_K activations)
        n_embd, n_ff, n_kv = dims
        for name, xin, K, M in (("wo", attn_out, n_embd, n_embd), ("gate", got["ffn_x"], n_embd, n_ff), ("up", got["ffn_x"], n_embd, n_ff),
                                ("down", got["glu"], n_ff, n_embd), ("q", got["attn_x"], n_embd, n_embd), ("k", got["attn_x"], n_embd, n_kv), ("v", got["attn_x"], n_embd, n_kv)):
            want = orc.mul_mat(types[name], W[name], K, M, xin[None, :], 1, act=O.ACT_CPU).reshape(-1)
            out = {"wo": "wo_out", "down": "down_out"}.get(name, name)
            assert O.nmse(want, got[out]) < NMSE_HARNESS, f"rep {rep} {name} vs ggml-cpu numerics"
    first = O.nmse(exact["wo_out"], got["wo_out"])
    assert first < 1e-9, first                                # phase 0 has no norm in front: integer-exact dots


@pytest.mark.parametrize("dims,t_v,t_down", [((2048, 6144, 512), O.Q4_K, O.Q4_K), ((2048, 6144, 512), O.Q6_K, O.Q6_K), ((4096, 14336, 1024), O.Q4_K, O.Q6_K),
                                             ((8192, 28672, 1024), O.Q4_K, O.Q4_K)])
def test_decode_chain_owner_side_handover(ctx, B, orc, dims, t_v, t_down):
    """the shape of the real llama graph: the RMS_NORM / MUL node values are dead, so the kernel takes the owner-side hand-over (b200_chain.cu:
    the CTA that owns 32 result rows adds the residual, multiplies with the norm weight, quantizes and publishes q8_1 blocks + sums of
    squares; the readers only rescale).  Same bars as above; the node values that do exist (results, residual sums, glu) are checked too.
    64 / 128 / 256 row groups on 148 CTAs; 8B and 70B widths."""
    rng = np.random.default_rng(4000 + dims[0] + t_down)
    lc, W, types, norm_ffn, norm_attn = make_layer(ctx, rng, dims, t_v, t_down, norm_outputs=False)
    assert ctx.decode_chain_supported(lc.program)
    for rep in range(2):
        attn_out = rng.uniform(-1, 1, dims[0]).astype(np.float32); inp_sa = rng.uniform(-2, 2, dims[0]).astype(np.float32)
        lc.set_inputs(attn_out, inp_sa)
        lc.clear_outputs()
        lc.run()
        assert ctx.decode_chain_error() == 0
        got = lc.outputs()
        for name in got:
            assert np.isfinite(got[name]).all(), (rep, name)
        assert np.array_equal(got["ffn_inp"], (got["wo_out"] + inp_sa).astype(np.float32))
        assert np.array_equal(got["l_out"], (got["down_out"] + got["ffn_inp"]).astype(np.float32))
        assert np.allclose(got["glu"], swiglu(got["gate"], got["up"]), rtol=1e-5, atol=1e-6)
        # every MUL_MAT node in isolation, on the chain's OWN input column (the dead norm columns recomputed from the residual sums the chain
        # did store), in the kernel's numerics (q8_1 integer dots): < 1e-6 -- a q8_1 quant that lands on the other side of a rounding boundary
        # (the owners quantize h * w and rescale d, the oracle quantizes (h * scale) * w) moves a result by ~1e-7.  Through the whole
        # pipeline such a step is amplified (the next activation is re-quantized), so the end-to-end bar is the reference's own.
        n_embd, n_ff, n_kv = dims
        ffn_x = (rms_norm(got["ffn_inp"], 1e-5) * norm_ffn).astype(np.float32); attn_x = (rms_norm(got["l_out"], 1e-5) * norm_attn).astype(np.float32)
        nodes = (("wo", attn_out, n_embd, n_embd), ("gate", ffn_x, n_embd, n_ff), ("up", ffn_x, n_embd, n_ff), ("down", got["glu"], n_ff, n_embd),
                 ("q", attn_x, n_embd, n_embd), ("k", attn_x, n_embd, n_kv), ("v", attn_x, n_embd, n_kv))
        for name, xin, K, M in nodes:
            out = {"wo": "wo_out", "down": "down_out"}.get(name, name)
            want = orc.mul_mat(types[name], W[name], K, M, xin[None, :], 1, act=O.ACT_Q8_1).reshape(-1)
            e = O.nmse(want, got[out])
            assert e < (1e-9 if name in ("wo", "down") else 1e-6), f"rep {rep} {name}: nmse {e} vs the q8_1 oracle on the chain's own column"
        exact = host_layer(orc, W, types, dims, attn_out, inp_sa, norm_ffn, norm_attn, 1e-5, O.ACT_Q8_1)
        for name in got:
            assert O.nmse(exact[name], got[name]) < NMSE_HARNESS, f"rep {rep} {name} vs the q8_1 pipeline end to end"
        if n_embd != 2048:
            continue        # (with these random blocks the wide layers' glu column is ~1e-5 and heavy-tailed: q8_K and q8_1 images of it differ by more than the bar whatever multiplies them)
        for name, xin, K, M in nodes:
            want = orc.mul_mat(types[name], W[name], K, M, xin[None, :], 1, act=O.ACT_CPU).reshape(-1)
            out = {"wo": "wo_out", "down": "down_out"}.get(name, name)
            assert O.nmse(want, got[out]) < NMSE_HARNESS, f"rep {rep} {name} vs ggml-cpu numerics"


@pytest.mark.parametrize("in_place", [False, True])
@pytest.mark.parametrize("n_dims", [128, 64])
def test_decode_chain_rope_and_kv_cache_tail(ctx, B, in_place, n_dims):
    """the chain's last phase with what a llama layer does to wq / wk / wv's rows next (b200_chain_rope): ROPE(Q), ROPE(K) -> f16 K-cache row,
    V -> f16 V-cache row, done by the row owners.  Expected values: the same chain WITHOUT the tail, then ggml_compute_forward_rope_f32's
    adjacent-pair rotation (ggml-cpu/ops.cpp:5524-5790) and the f16 rounding of SET_ROWS in numpy."""
    rng = np.random.default_rng(31 + n_dims)
    dims = (4096, 14336, 1024); hd = 128
    base = 500000.0
    rope = dict(head_dim=hd, n_dims=n_dims, theta_scale=float(np.float32(base) ** np.float32(-2.0 / n_dims)), freq_scale=1.0, attn_factor=1.0, kv_size=64, in_place=in_place)
    from decode_chain import LayerChain
    types = {"wo": O.Q4_K, "gate": O.Q4_K, "up": O.Q4_K, "down": O.Q6_K, "q": O.Q4_K, "k": O.Q4_K, "v": O.Q6_K}
    shape = {"wo": (4096, 4096), "gate": (4096, 14336), "up": (4096, 14336), "down": (14336, 4096), "q": (4096, 4096), "k": (4096, 1024), "v": (4096, 1024)}
    W = {n: (types[n], O.random_blocks(rng, types[n], *shape[n], scale=0.0028)) for n in types}
    nf = rng.uniform(0.5, 1.5, 4096).astype(np.float32); na = rng.uniform(0.5, 1.5, 4096).astype(np.float32)
    plain = LayerChain(ctx, *dims, W, nf, na, norm_outputs=False)
    tail = LayerChain(ctx, *dims, W, nf, na, norm_outputs=False, rope=rope)
    assert ctx.decode_chain_supported(tail.program)
    attn_out = rng.uniform(-1, 1, 4096).astype(np.float32); inp_sa = rng.uniform(-2, 2, 4096).astype(np.float32)
    for pos, crow in ((0, 0), (37, 5), (4099, 63)):
        for lc in (plain, tail):
            lc.set_inputs(attn_out, inp_sa); lc.clear_outputs()
        ctx.memset(tail.k_cache.ptr, 0xff, tail.k_cache.nbytes); ctx.memset(tail.v_cache.ptr, 0xff, tail.v_cache.nbytes)
        tail.set_position(pos, crow)
        plain.run(); tail.run()
        assert ctx.decode_chain_error() == 0
        want = plain.outputs()
        q, k, v = want["q"], want["k"], want["v"]
        def rot(x):
            x = x.reshape(-1, hd).astype(np.float32).copy()
            ip = np.arange(n_dims // 2, dtype=np.float32)
            theta = (np.float32(pos) * np.power(np.float32(rope["theta_scale"]), ip)).astype(np.float32)
            cs, sn = np.cos(theta.astype(np.float64)).astype(np.float32), np.sin(theta.astype(np.float64)).astype(np.float32)
            x0, x1 = x[:, 0:n_dims:2].copy(), x[:, 1:n_dims:2].copy()
            x[:, 0:n_dims:2] = x0 * cs - x1 * sn; x[:, 1:n_dims:2] = x0 * sn + x1 * cs
            return x.reshape(-1)
        got_q, got_k = ctx.download(tail.q_rope), ctx.download(tail.k_rope)
        # the angle goes through powf / sincosf on the device: a relative error of 1e-6 in an angle of 4e3 rad is 4e-3 rad
        atol = 1e-5 + 4e-6 * pos
        assert np.allclose(got_q, rot(q), rtol=1e-5, atol=atol * np.abs(q).max()), (pos, np.abs(got_q - rot(q)).max())
        assert np.allclose(got_k, rot(k), rtol=1e-5, atol=atol * np.abs(k).max())
        assert np.array_equal(ctx.download(tail.v), v)
        if not in_place:
            assert np.array_equal(ctx.download(tail.q), q) and np.array_equal(ctx.download(tail.k), k)      # the matmul nodes' own values
        kc = ctx.download(tail.k_cache, np.float16).reshape(rope["kv_size"], 1024); vc = ctx.download(tail.v_cache, np.float16).reshape(rope["kv_size"], 1024)
        assert np.array_equal(kc[crow], got_k.astype(np.float16)) and np.array_equal(vc[crow], v.astype(np.float16))
        other = np.delete(np.arange(rope["kv_size"]), crow)
        assert (kc[other].view(np.uint16) == 0xffff).all() and (vc[other].view(np.uint16) == 0xffff).all()   # no other cache row touched
        for name in ("wo_out", "ffn_inp", "gate", "up", "glu", "down_out", "l_out"):
            assert np.array_equal(ctx.download(getattr(tail, name)), want[name]), name


def test_decode_chain_llama8b_shapes_vs_node_by_node(ctx, B):
    """Llama-3-8B shapes with the Q4_K_M mix (q6_K attn_v / ffn_down), compared with the separate launches of the flat API"""
    rng = np.random.default_rng(77)
    dims = (4096, 14336, 1024)
    lc, W, types, norm_ffn, norm_attn = make_layer(ctx, rng, dims, O.Q6_K, O.Q6_K)
    attn_out = rng.uniform(-1, 1, dims[0]).astype(np.float32); inp_sa = rng.uniform(-2, 2, dims[0]).astype(np.float32)
    lc.set_inputs(attn_out, inp_sa)
    lc.clear_outputs()
    for _ in range(3):
        lc.run()
    assert ctx.decode_chain_error() == 0
    got = lc.outputs()
    # node by node: the device matmuls on the chain's own intermediate columns (each checks one phase in isolation)
    def mm(name, x_t, M):
        d = ctx.empty(B.F32, [M])
        ctx.note_write(x_t)
        ctx.mul_mat(lc.w[name], x_t, d)
        return ctx.download(d)
    for name, x_t, M in (("wo", lc.attn_out, 4096), ("gate", lc.ffn_x, 14336), ("up", lc.ffn_x, 14336), ("down", lc.glu, 4096),
                         ("q", lc.attn_x, 4096), ("k", lc.attn_x, 1024), ("v", lc.attn_x, 1024)):
        out = {"wo": "wo_out", "down": "down_out"}.get(name, name)
        e = O.nmse(mm(name, x_t, M), got[out])
        assert e < 1e-9, f"{name}: nmse {e} vs b200_mul_mat on the same column"
    assert np.array_equal(got["ffn_inp"], (got["wo_out"] + inp_sa).astype(np.float32))
    assert np.allclose(got["ffn_x"], rms_norm(got["ffn_inp"], 1e-5) * norm_ffn, rtol=2e-6, atol=1e-7)
    assert np.allclose(got["glu"], swiglu(got["gate"], got["up"]), rtol=1e-5, atol=1e-6)
    assert np.array_equal(got["l_out"], (got["down_out"] + got["ffn_inp"]).astype(np.float32))
    assert np.allclose(got["attn_x"], rms_norm(got["l_out"], 1e-5) * norm_attn, rtol=2e-6, atol=1e-7)


def test_decode_chain_declines_what_it_cannot_take(ctx, B):
    rng = np.random.default_rng(5)
    K, M = 2048, 64
    w = ctx.upload(O.Q8_0, [K, M], O.random_blocks(rng, O.Q8_0, K, M))            # type outside the chain
    x, d = ctx.empty(B.F32, [K]), ctx.empty(B.F32, [M])
    prog = ctx.chain_program([dict(w=[w], dst=[d], in0=x)])
    assert not ctx.decode_chain_supported(prog)
    with pytest.raises(B.B200Error):
        ctx.decode_chain(prog)
    w2 = ctx.upload(O.Q4_K, [K, 6], O.random_blocks(rng, O.Q4_K, K, 6))           # M % 4 != 0
    assert not ctx.decode_chain_supported(ctx.chain_program([dict(w=[w2], dst=[ctx.empty(B.F32, [6])], in0=x)]))
