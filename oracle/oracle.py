"""oracle.py -- TEST INFRASTRUCTURE ONLY: ctypes bindings for the CPU oracle.

Two libraries, both built by oracle/Makefile:
  * _ref/liboracle.so          our plain-C restatement (oracle/ggml_oracle.c); class `Oracle`
  * _ref/<isa>/ref_driver.so   flat driver over the UNMODIFIED reference compiled in place from
                               /root/reference (libggml-ref.so);  class `Reference`
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product (super-llama.cpp_b200/) never does.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K, I32, BF16 = 0, 1, 2, 8, 9, 12, 13, 14, 15, 26, 30
TYPE_NAMES = {F32: "f32", F16: "f16", Q4_0: "q4_0", Q8_0: "q8_0", Q8_1: "q8_1", Q4_K: "q4_K", Q5_K: "q5_K",
              Q6_K: "q6_K", Q8_K: "q8_K", BF16: "bf16"}
BLCK = {F32: 1, F16: 1, BF16: 1, Q4_0: 32, Q8_0: 32, Q8_1: 32, Q4_K: 256, Q5_K: 256, Q6_K: 256, Q8_K: 256}
TSIZE = {F32: 4, F16: 2, BF16: 2, Q4_0: 18, Q8_0: 34, Q8_1: 36, Q4_K: 144, Q5_K: 176, Q6_K: 210, Q8_K: 292}
ACT_CPU, ACT_Q8_1, ACT_F64 = 0, 1, 2
Q8_1_X86, Q8_1_REF = 0, 1


def row_size(t, k):
    return (k // BLCK[t]) * TSIZE[t]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def nmse(ref, out):
    """test-backend-ops' metric (tests/test-backend-ops.cpp:274-287): sum((a-b)^2)/sum(a^2)."""
    ref = np.asarray(ref, dtype=np.float64).ravel()
    out = np.asarray(out, dtype=np.float64).ravel()
    den = float(np.sum(ref * ref))
    num = float(np.sum((ref - out) ** 2))
    return num / den if den > 0 else num


def random_blocks(rng, t, k, nrows, scale=0.05):
    """Random but well-formed GGUF weight blocks of type t for a [k, nrows] matrix (uint8 array).
    Every quant/scale bit pattern is exercised; the fp16 super-scales are kept small and finite."""
    nb = nrows * (k // BLCK[t])
    raw = rng.integers(0, 256, size=(nb, TSIZE[t]), dtype=np.uint8)

    def put_half(col, vals):
        raw[:, col:col + 2] = np.asarray(vals, dtype=np.float16).view(np.uint8).reshape(nb, 2)

    if t in (Q4_0, Q8_0):
        put_half(0, rng.uniform(-1, 1, nb) * scale)
    elif t in (Q4_K, Q5_K):
        put_half(0, rng.uniform(0.1, 1, nb) * scale / 32)
        put_half(2, rng.uniform(0.1, 1, nb) * scale / 32)
    elif t == Q6_K:
        put_half(208, rng.uniform(-1, 1, nb) * scale / 64)
    elif t == F16:
        return (rng.uniform(-1, 1, nrows * k) * scale).astype(np.float16).view(np.uint8)
    elif t == BF16:
        f = (rng.uniform(-1, 1, nrows * k) * scale).astype(np.float32)
        return (f.view(np.uint32) >> 16).astype(np.uint16).view(np.uint8)
    elif t == F32:
        return (rng.uniform(-1, 1, nrows * k) * scale).astype(np.float32).view(np.uint8)
    else:
        raise ValueError(t)
    return raw.reshape(-1)


class Oracle:
    """Our plain-C restatement of the reference algorithms (oracle/ggml_oracle.c)."""

    def __init__(self, path=None):
        path = path or os.path.join(HERE, "_ref", "liboracle.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle oracle`")
        L = self.lib = C.CDLL(path)
        i64, vp, f, i = C.c_int64, C.c_void_p, C.c_float, C.c_int
        L.orc_quantize_row_q8_1.argtypes = [vp, vp, i64, i]
        L.orc_quantize_row_q8_0.argtypes = [vp, vp, i64]
        L.orc_quantize_row_q8_K.argtypes = [vp, vp, i64]
        L.orc_quantize_row_q4_0_ref.argtypes = [vp, vp, i64]
        L.orc_quantize_row_q8_0_ref.argtypes = [vp, vp, i64]
        L.orc_dequantize_row.argtypes = [i, vp, vp, i64]
        L.orc_fp32_to_fp16_row.argtypes = [vp, vp, i64]
        L.orc_fp16_to_fp32_row.argtypes = [vp, vp, i64]
        L.orc_mul_mat.argtypes = [i, vp, i64, i64, i64, i64, vp, i64, i64, i64, vp, i]
        L.orc_mul_mat_id.argtypes = [i, vp, i64, i64, i64, vp, i64, i64, vp, i64, vp, i]
        L.orc_flash_attn_ext.argtypes = [vp] * 6 + [i64] * 11 + [f, f, f, i]
        L.orc_split_row_range.argtypes = [i64, vp, i, i, i64, C.POINTER(i64), C.POINTER(i64)]
        L.orc_split_row_range.restype = None

    def quantize_q8_1(self, x, variant=Q8_1_X86):
        x = _f32(x)
        k = x.shape[-1]
        y = np.empty(x.size // 32 * 36, dtype=np.uint8)
        self.lib.orc_quantize_row_q8_1(_p(x), _p(y), x.size, variant)   # rows are independent blocks
        assert k % 32 == 0
        return y

    def quantize(self, t, x):
        x = _f32(x)
        y = np.empty(row_size(t, x.size), dtype=np.uint8)
        fn = {Q8_0: self.lib.orc_quantize_row_q8_0, Q8_K: self.lib.orc_quantize_row_q8_K}[t]
        fn(_p(x), _p(y), x.size)
        return y

    def quantize_weights(self, t, x):
        """deterministic weight quantizers restated for the two legacy types only (q4_0, q8_0)."""
        x = _f32(x)
        y = np.empty(row_size(t, x.size), dtype=np.uint8)
        fn = {Q4_0: self.lib.orc_quantize_row_q4_0_ref, Q8_0: self.lib.orc_quantize_row_q8_0_ref}[t]
        fn(_p(x), _p(y), x.size)
        return y

    def dequantize(self, t, blocks, n):
        blocks = np.ascontiguousarray(blocks)
        y = np.empty(n, dtype=np.float32)
        assert self.lib.orc_dequantize_row(t, _p(blocks), _p(y), n) == 0
        return y

    def mul_mat(self, t, w, K, M, x, N, ne02=1, ne03=1, ne12=1, ne13=1, act=ACT_CPU):
        w = np.ascontiguousarray(w); x = _f32(x)
        dst = np.empty(M * N * ne12 * ne13, dtype=np.float32)
        rc = self.lib.orc_mul_mat(t, _p(w), K, M, ne02, ne03, _p(x), N, ne12, ne13, _p(dst), act)
        assert rc == 0, rc
        return dst.reshape(ne13, ne12, N, M)

    def mul_mat_id(self, t, as_, K, M, n_expert, b, ne11, n_tokens, ids, n_used, act=ACT_CPU):
        as_ = np.ascontiguousarray(as_); b = _f32(b); ids = np.ascontiguousarray(ids, dtype=np.int32)
        dst = np.empty(M * n_used * n_tokens, dtype=np.float32)
        rc = self.lib.orc_mul_mat_id(t, _p(as_), K, M, n_expert, _p(b), ne11, n_tokens, _p(ids), n_used, _p(dst), act)
        assert rc == 0, rc
        return dst.reshape(n_tokens, n_used, M)

    def flash_attn_ext(self, q, k, v, mask=None, sinks=None, scale=1.0, max_bias=0.0, softcap=0.0, faithful=False):
        """q[S,H,N,DK] f32, k[Sk,Hk,KV,DK] f16, v[Sk,Hk,KV,DV] f16, mask[Sm,Hm,Nm,KV] f16 -> dst[S,N,H,DV]."""
        q = _f32(q); k = np.ascontiguousarray(k, dtype=np.float16); v = np.ascontiguousarray(v, dtype=np.float16)
        S, H, N, DK = q.shape; Sk, Hk, KV, _ = k.shape; DV = v.shape[-1]
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.float16); Sm, Hm, Nm, _ = mask.shape
        else:
            Sm = Hm = Nm = 1
        if sinks is not None:
            sinks = _f32(sinks)
        dst = np.empty((S, N, H, DV), dtype=np.float32)
        rc = self.lib.orc_flash_attn_ext(_p(q), _p(k), _p(v), _p(mask), _p(sinks), _p(dst), DK, DV, N, H, S, KV, Hk, Sk,
                                         Nm, Hm, Sm, scale, max_bias, softcap, 1 if faithful else 0)
        assert rc == 0, rc
        return dst

    def split_row_range(self, nrows, world, rank, rounding, tensor_split=None):
        lo, hi = C.c_int64(), C.c_int64()
        ts = None if tensor_split is None else _f32(tensor_split)
        self.lib.orc_split_row_range(nrows, _p(ts), world, rank, rounding, C.byref(lo), C.byref(hi))
        return lo.value, hi.value


class Reference:
    """The UNMODIFIED reference (ggml-cpu) through oracle/ref_driver.c.  Available wherever oracle/_ref/<isa>/
    was built (this container; it travels to the GPU box with the snapshot)."""

    def __init__(self, isa="avx2"):
        d = os.path.join(HERE, "_ref", isa)
        path = os.path.join(d, "ref_driver.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle ref` where /root/reference exists")
        self.dir = d
        C.CDLL(os.path.join(d, "libggml-ref.so"), mode=C.RTLD_GLOBAL)
        L = self.lib = C.CDLL(path)
        i64, vp, f, i, dbl = C.c_int64, C.c_void_p, C.c_float, C.c_int, C.c_double
        L.refd_from_float.argtypes = [i, vp, vp, i64]
        L.refd_quantize_row_q8_1_ref.argtypes = [vp, vp, i64]
        L.refd_quantize_chunk.argtypes = [i, vp, vp, i64, i64]; L.refd_quantize_chunk.restype = C.c_size_t
        L.refd_dequantize_row.argtypes = [i, vp, vp, i64]
        L.refd_mul_mat.argtypes = [i, vp, i64, i64, i64, i64, vp, i64, i64, i64, vp, i, i]; L.refd_mul_mat.restype = dbl
        L.refd_mul_mat_id.argtypes = [i, vp, i64, i64, i64, vp, i64, i64, vp, i64, vp, i, i]; L.refd_mul_mat_id.restype = dbl
        L.refd_flash_attn_ext.argtypes = [vp, vp, vp, i, vp, vp, vp] + [i64] * 11 + [f, f, f, i, i, i]
        L.refd_flash_attn_ext.restype = dbl
        L.refd_mm_create.argtypes = [i, i64, i64, i64, i]; L.refd_mm_create.restype = vp
        L.refd_fa_create.argtypes = [i64] * 7 + [f, i]; L.refd_fa_create.restype = vp
        L.refd_op_input.argtypes = [vp, i, C.POINTER(C.c_size_t)]; L.refd_op_input.restype = vp
        L.refd_op_output.argtypes = [vp, C.POINTER(C.c_size_t)]; L.refd_op_output.restype = vp
        L.refd_op_run.argtypes = [vp]; L.refd_op_run.restype = dbl
        L.refd_op_free.argtypes = [vp]; L.refd_op_free.restype = None
        self.last_us = None
        self.threads = os.cpu_count() or 1

    def from_float(self, t, x):
        x = _f32(x)
        y = np.empty(row_size(t, x.size), dtype=np.uint8)
        assert self.lib.refd_from_float(t, _p(x), _p(y), x.size) == 0
        return y

    def quantize_q8_1(self, x, variant=Q8_1_X86):
        x = _f32(x)
        if variant == Q8_1_X86:
            return self.from_float(Q8_1, x)
        y = np.empty(x.size // 32 * 36, dtype=np.uint8)
        self.lib.refd_quantize_row_q8_1_ref(_p(x), _p(y), x.size)
        return y

    def quantize_weights(self, t, x, k):
        x = _f32(x)
        nrows = x.size // k
        y = np.empty(row_size(t, k) * nrows, dtype=np.uint8)
        n = self.lib.refd_quantize_chunk(t, _p(x), _p(y), nrows, k)
        assert n == y.size, (n, y.size)
        return y

    def dequantize(self, t, blocks, n):
        blocks = np.ascontiguousarray(blocks)
        y = np.empty(n, dtype=np.float32)
        assert self.lib.refd_dequantize_row(t, _p(blocks), _p(y), n) == 0
        return y

    def mul_mat(self, t, w, K, M, x, N, ne02=1, ne03=1, ne12=1, ne13=1, threads=None, reps=1, want_out=True):
        w = np.ascontiguousarray(w); x = _f32(x)
        dst = np.empty(M * N * ne12 * ne13, dtype=np.float32) if want_out else None
        us = self.lib.refd_mul_mat(t, _p(w), K, M, ne02, ne03, _p(x), N, ne12, ne13, _p(dst), threads or self.threads, reps)
        assert us >= 0
        self.last_us = us
        return dst.reshape(ne13, ne12, N, M) if want_out else None

    def mul_mat_id(self, t, as_, K, M, n_expert, b, ne11, n_tokens, ids, n_used, threads=None, reps=1):
        as_ = np.ascontiguousarray(as_); b = _f32(b); ids = np.ascontiguousarray(ids, dtype=np.int32)
        dst = np.empty(M * n_used * n_tokens, dtype=np.float32)
        us = self.lib.refd_mul_mat_id(t, _p(as_), K, M, n_expert, _p(b), ne11, n_tokens, _p(ids), n_used, _p(dst),
                                      threads or self.threads, reps)
        assert us >= 0
        self.last_us = us
        return dst.reshape(n_tokens, n_used, M)

    def flash_attn_ext(self, q, k, v, mask=None, sinks=None, scale=1.0, max_bias=0.0, softcap=0.0, prec=0,
                       threads=None, reps=1, kv_type=None, kv_shape=None):
        """kv_type (one ggml type for K and V, as the reference's own test builds them) + kv_shape = (Sk, Hk, KV, DK, DV): k / v are raw blocks"""
        q = _f32(q)
        S, H, N, DK = q.shape
        if kv_type is None:
            k = np.ascontiguousarray(k, dtype=np.float16); v = np.ascontiguousarray(v, dtype=np.float16)
            Sk, Hk, KV, _ = k.shape; DV = v.shape[-1]
        else:
            k = np.ascontiguousarray(k, dtype=np.uint8); v = np.ascontiguousarray(v, dtype=np.uint8)
            Sk, Hk, KV, _, DV = kv_shape
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.float16); Sm, Hm, Nm, _ = mask.shape
        else:
            Sm = Hm = Nm = 1
        if sinks is not None:
            sinks = _f32(sinks)
        dst = np.empty((S, N, H, DV), dtype=np.float32)
        us = self.lib.refd_flash_attn_ext(_p(q), _p(k), _p(v), F16 if kv_type is None else kv_type, _p(mask), _p(sinks), _p(dst), DK, DV, N, H, S, KV, Hk, Sk,
                                          Nm, Hm, Sm, scale, max_bias, softcap, prec, threads or self.threads, reps)
        assert us >= 0
        self.last_us = us
        return dst


class RefOp:
    """A persistent one-node ggml-cpu graph (weights set once, re-run many times): the timed CPU baseline."""

    def __init__(self, ref, handle):
        if not handle:
            raise MemoryError("reference op allocation failed")
        self.ref, self.h = ref, handle

    def input(self, i):
        n = C.c_size_t()
        p = self.ref.lib.refd_op_input(self.h, i, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,)) if p else None

    def output(self):
        n = C.c_size_t()
        p = self.ref.lib.refd_op_output(self.h, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n.value // 4,))

    def run(self):
        us = self.ref.lib.refd_op_run(self.h)
        assert us >= 0
        return us

    def free(self):
        if self.h:
            self.ref.lib.refd_op_free(self.h)
            self.h = None


def ref_mm_op(ref, t, K, M, N, threads=None):
    return RefOp(ref, ref.lib.refd_mm_create(t, K, M, N, threads or ref.threads))


def ref_fa_op(ref, DK, DV, N, H, KV, Hk, Nm, scale, threads=None):
    return RefOp(ref, ref.lib.refd_fa_create(DK, DV, N, H, KV, Hk, Nm, scale, threads or ref.threads))


def have_reference(isa="avx2"):
    return os.path.exists(os.path.join(HERE, "_ref", isa, "ref_driver.so"))
