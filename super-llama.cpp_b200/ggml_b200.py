"""ggml_b200.py -- Python host side of the B200 backend: ctypes over the C-ABI in include/ggml-b200.h.

This mirrors the reference's operator interface for the hot path (ggml_mul_mat / ggml_mul_mat_id /
ggml_flash_attn_ext, ggml/include/ggml.h:1403-1420,2314-2344; quantize_row_q8_1): same argument meaning,
ggml shapes (ne[0] fastest) and byte strides.  It is used by the parity tests, smoke() and bench.py; llama.cpp
itself reaches the same library through the ggml plugin entry points.

There is NO fallback of any kind: if libggml-b200.so is missing, or no sm_100 device is present, every entry
point raises.  Nothing here imports oracle/.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libggml-b200.so")

F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, I32, BF16 = 0, 1, 2, 8, 9, 12, 13, 14, 26, 30
BLCK = {F32: 1, F16: 1, BF16: 1, I32: 1, Q4_0: 32, Q8_0: 32, Q8_1: 32, Q4_K: 256, Q5_K: 256, Q6_K: 256}
TSIZE = {F32: 4, F16: 2, BF16: 2, I32: 4, Q4_0: 18, Q8_0: 34, Q8_1: 36, Q4_K: 144, Q5_K: 176, Q6_K: 210}
Q8_1_X86, Q8_1_REF = 0, 1


class B200Error(RuntimeError):
    pass


class _Tensor(C.Structure):          # struct b200_tensor
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("flags", C.c_int32),
                ("ne", C.c_int64 * 4), ("nb", C.c_size_t * 4)]


class Exchange(C.Structure):
    """b200_exchange: the arenas of all ranks as mapped in this process"""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("base", C.c_void_p * 8), ("ctrl_off", C.c_size_t)]


class LoadItem(C.Structure):          # b200_load_item
    _fields_ = [("file_offset", C.c_int64), ("nbytes", C.c_size_t), ("dst", C.c_void_p), ("dst_offset", C.c_size_t), ("repack", C.c_int32)]


class LoadStats(C.Structure):         # b200_load_stats
    _fields_ = [("seconds", C.c_double), ("read_seconds", C.c_double), ("bytes", C.c_uint64), ("chunks", C.c_uint64)]


class KvPages(C.Structure):
    """b200_kv_pages: page table of a paged KV cache (logical page -> physical page of the pool)"""
    _fields_ = [("table", C.c_void_p), ("n_pages", C.c_int32), ("page_tokens", C.c_int32), ("pool_tokens", C.c_int64)]


class ChainRope(C.Structure):
    """b200_chain_rope: ROPE / KV-cache write of one result of the chain's last phase, done by the CTA that owns the rows"""
    _fields_ = [("kind", C.c_int32), ("skip_dst", C.c_int32), ("head_dim", C.c_int32), ("n_dims", C.c_int32), ("freq_scale", C.c_float), ("theta_scale", C.c_float), ("attn_factor", C.c_float),
                ("pos", C.c_void_p), ("freq_factors", C.c_void_p), ("out", C.c_void_p), ("cache_idx", C.c_void_p), ("cache", C.c_void_p), ("cache_nb1", C.c_size_t)]


T_NONE, T_ROPE, T_ROPE_CACHE, T_CACHE = 0, 1, 2, 3


class ChainPhase(C.Structure):
    """b200_chain_phase: one MUL_MAT group of the persistent decode chain with its prologue / epilogue"""
    _fields_ = [("prologue", C.c_int32), ("epilogue", C.c_int32), ("n_w", C.c_int32), ("eps", C.c_float),
                ("w", C.POINTER(_Tensor) * 3), ("dst", C.POINTER(_Tensor) * 3),
                ("in0", C.POINTER(_Tensor)), ("in1", C.POINTER(_Tensor)), ("normw", C.POINTER(_Tensor)),
                ("out_add", C.POINTER(_Tensor)), ("out_norm", C.POINTER(_Tensor)), ("out_mul", C.POINTER(_Tensor)), ("glu", C.POINTER(_Tensor)),
                ("tail", C.POINTER(ChainRope) * 3)]


class ChainSplit(C.Structure):
    """b200_chain_split: the decode chain row-split over the GPUs of one node (one process per GPU); ll[r] = rank r's hand-over area"""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("ll", C.c_void_p * 8), ("ll_bytes", C.c_uint64), ("row_lo", (C.c_int64 * 3) * 6)]


X_PLAIN, X_ADD_NORM_MUL, X_NORM_MUL = 0, 1, 2
Y_PLAIN, Y_SWIGLU = 0, 1

_lib = None


def lib():
    """dlopen libggml-b200.so (once) and declare the C-ABI.  Raises if the extension is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(f"{LIB_PATH} is missing: build it with `make -C super-llama.cpp_b200` "
                        "(or __graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i, i64, sz, f = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
    tp = C.POINTER(_Tensor)
    sig = {
        "b200_device_count": (i, []), "b200_last_error": (C.c_char_p, []),
        "b200_ctx_create": (vp, [i]), "b200_ctx_destroy": (None, [vp]), "b200_ctx_device": (i, [vp]),
        "b200_ctx_stream": (vp, [vp]), "b200_ctx_synchronize": (i, [vp]), "b200_ctx_launch_count": (C.c_uint64, [vp]),
        "b200_ctx_invalidate": (i, [vp]), "b200_ctx_scratch_generation": (C.c_uint64, [vp]), "b200_ctx_note_write": (i, [vp, vp, C.c_size_t]),
        "b200_malloc": (vp, [vp, sz]), "b200_free": (i, [vp, vp]), "b200_host_alloc": (vp, [sz]), "b200_host_free": (i, [vp]),
        "b200_memcpy_h2d": (i, [vp, vp, vp, sz]), "b200_memcpy_d2h": (i, [vp, vp, vp, sz]), "b200_memset": (i, [vp, vp, i, sz]),
        "b200_quantize_q8_1": (i, [vp, vp, sz, vp, i64, i64, i]),
        "b200_mul_mat": (i, [vp, tp, tp, tp]), "b200_mul_mat_noalloc": (i, [vp, tp, tp, tp]), "b200_noalloc_scratch_bytes": (sz, []),
        "b200_mul_mat_multi": (i, [vp, C.POINTER(tp), i, tp, C.POINTER(tp)]),
        "b200_decode_chain": (i, [vp, C.POINTER(ChainPhase), i]), "b200_decode_chain_supported": (i, [vp, C.POINTER(ChainPhase), i]), "b200_decode_chain_error": (i, [vp]),
        "b200_decode_chain_split": (i, [vp, C.POINTER(ChainPhase), i, C.POINTER(ChainSplit)]), "b200_decode_chain_split_supported": (i, [vp, C.POINTER(ChainPhase), i, C.POINTER(ChainSplit)]),
        "b200_decode_chain_split_bytes": (C.c_size_t, [C.c_int64]), "b200_abi_sizeof": (C.c_size_t, [i]),
        "b200_mul_mat_gate_up_swiglu": (i, [vp, tp, tp, tp, tp, tp, tp]),
        "b200_repack_supported": (i, [i, i64]), "b200_repack": (i, [vp, i, vp, i64, i64, sz]), "b200_unrepack": (i, [vp, i, vp, i64, i64, sz]), "b200_mul_mat_id": (i, [vp, tp, tp, tp, tp]),
        "b200_flash_attn_ext": (i, [vp, tp, tp, tp, tp, tp, tp, f, f, f]),
        "b200_flash_attn_ext_paged": (i, [vp, tp, tp, tp, C.POINTER(KvPages), tp, tp, tp, f, f, f]),
        "b200_ctx_launch_count_kind": (C.c_uint64, [vp, i]), "b200_ctx_profile": (i, [vp, i]),
        "b200_ctx_profile_get": (i, [vp, i, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
        "b200_event_create": (vp, []), "b200_event_destroy": (i, [vp]), "b200_event_record": (i, [vp, vp]),
        "b200_event_elapsed_ms": (i, [vp, vp, C.POINTER(f)]),
        "b200_debug_trace": (i, [vp, vp, i]),
        "b200_graph_begin": (i, [vp]), "b200_graph_end": (vp, [vp]), "b200_graph_launch": (i, [vp, vp]), "b200_graph_destroy": (i, [vp]),
        "b200_split_row_range": (i, [i64, vp, i, i, i64, C.POINTER(i64), C.POINTER(i64)]),
        "b200_distribute_layers": (i, [i, i, i, vp, i, i, vp, vp, vp]),
        "b200_plan_tensor_split": (i, [vp, i, i, vp]),
        "b200_peer_handle": (i, [vp, vp, vp]), "b200_peer_open": (i, [vp, vp, C.POINTER(vp)]), "b200_peer_close": (i, [vp, vp]),
        "b200_mul_mat_rowsplit": (i, [vp, tp, tp, tp, i64, C.POINTER(Exchange)]),
        "b200_mul_mat_rowsplit_multi": (i, [vp, C.POINTER(tp), i, tp, C.POINTER(tp), C.POINTER(i64), C.POINTER(Exchange)]),
        "b200_exchange_error": (i, [vp, C.POINTER(Exchange)]),
        "b200_load_file": (i, [vp, C.c_char_p, C.POINTER(LoadItem), i, sz, i, C.POINTER(LoadStats)]),
        "b200_load_row_slice": (i, [i, i64, i64, i64, i64, C.POINTER(LoadItem)]),
        "ggml_backend_init": (vp, []), "ggml_backend_score": (i, []), "ggml_backend_b200_reg": (vp, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _check(rc, what):
    if rc != 0:
        raise B200Error(f"{what} failed ({rc}): {lib().b200_last_error().decode()}")


def row_size(t, ne0):
    return (ne0 // BLCK[t]) * TSIZE[t]


def contiguous_strides(t, ne):
    nb = [TSIZE[t], row_size(t, ne[0]), 0, 0]
    nb[2] = nb[1] * ne[1]
    nb[3] = nb[2] * ne[2]
    return nb


class DeviceTensor:
    """A ggml-style strided view over device memory (struct b200_tensor)."""

    def __init__(self, ctx, t, ne, nb=None, ptr=None, owner=None):
        ne = list(ne) + [1] * (4 - len(ne))
        self.ctx, self.type, self.ne = ctx, t, ne
        self.nb = list(nb) if nb is not None else contiguous_strides(t, ne)
        self.owner = owner
        self.flags = 0               # B200_TENSOR_REPACKED after Context.repack()
        if ptr is None:
            self.nbytes = self.span()
            self.ptr = ctx.malloc(self.nbytes)
            self.owner = self
            self._owned = True
        else:
            self.ptr, self._owned, self.nbytes = ptr, False, self.span()

    def span(self):
        """bytes from data to the end of the last element (ggml_nbytes)."""
        if any(n <= 0 for n in self.ne):
            return 0
        n = self.ne[0] * self.nb[0] // BLCK[self.type] if BLCK[self.type] > 1 else TSIZE[self.type] + (self.ne[0] - 1) * self.nb[0]
        for d in range(1, 4):
            n += (self.ne[d] - 1) * self.nb[d]
        return n

    def c(self):
        t = _Tensor()
        t.data, t.type, t.flags = self.ptr, self.type, self.flags
        for d in range(4):
            t.ne[d], t.nb[d] = self.ne[d], self.nb[d]
        return t

    def view(self, ne, nb, offset=0):
        v = DeviceTensor(self.ctx, self.type, ne, nb, ptr=self.ptr + offset, owner=self.owner or self)
        v.flags = self.flags         # a whole-row view of a repacked weight is still repacked
        return v

    def free(self):
        if self._owned and self.ptr:
            self.ctx.free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Context:
    """One device + one stream (struct b200_ctx).  All operator calls are asynchronous on that stream."""

    def __init__(self, device=0):
        L = lib()
        n = L.b200_device_count()
        if n <= 0:
            raise B200Error("no sm_100 (B200) device visible: the B200 backend has no CPU fallback")
        self.h = L.b200_ctx_create(device)
        if not self.h:
            raise B200Error(L.b200_last_error().decode())
        self.device = device

    def close(self):
        if self.h:
            lib().b200_ctx_destroy(self.h)
            self.h = None

    # ---- memory ---------------------------------------------------------------------------------
    def malloc(self, nbytes):
        p = lib().b200_malloc(self.h, max(int(nbytes), 1))
        if not p:
            raise B200Error(lib().b200_last_error().decode())
        return p

    def free(self, p):
        if self.h:
            lib().b200_free(self.h, p)

    def invalidate(self):
        """drop the cached q8_1 activation image (after activations were written by something other than this API)"""
        lib().b200_ctx_invalidate(self.h)

    def note_write(self, t):
        """the device tensor t was (re)written outside this API: drop activation images cached from it"""
        lib().b200_ctx_note_write(self.h, t.ptr, t.nbytes)

    def synchronize(self):
        _check(lib().b200_ctx_synchronize(self.h), "synchronize")

    @property
    def stream(self):
        return lib().b200_ctx_stream(self.h)

    @property
    def launches(self):
        return int(lib().b200_ctx_launch_count(self.h))

    # ---- measurement: events on the ctx stream, per-kernel event profile, CUDA graphs --------------
    KINDS = ("quantize_q8_1", "gemv", "gemv_id", "gemm", "fattn", "fattn_combine", "matmul_float", "glue", "mmid_sort", "convert")

    def event(self):
        e = lib().b200_event_create()
        if not e:
            raise B200Error("event create failed")
        return e

    def record(self, ev):
        _check(lib().b200_event_record(self.h, ev), "event_record")

    @staticmethod
    def elapsed_ms(e0, e1):
        ms = C.c_float()
        _check(lib().b200_event_elapsed_ms(e0, e1, C.byref(ms)), "event_elapsed")
        return ms.value

    def profile(self, enable):
        _check(lib().b200_ctx_profile(self.h, 1 if enable else 0), "profile")

    def profile_get(self):
        """{kind: (total_ms, launches)} for the launches since profile(True)"""
        out = {}
        for k, name in enumerate(self.KINDS):
            ms, n = C.c_double(), C.c_uint64()
            _check(lib().b200_ctx_profile_get(self.h, k, C.byref(ms), C.byref(n)), "profile_get")
            if n.value:
                out[name] = (ms.value, n.value)
        return out

    def launches_by_kind(self):
        return {name: int(lib().b200_ctx_launch_count_kind(self.h, k)) for k, name in enumerate(self.KINDS)}

    def graph_begin(self):
        _check(lib().b200_graph_begin(self.h), "graph_begin")

    def graph_end(self):
        g = lib().b200_graph_end(self.h)
        if not g:
            raise B200Error(lib().b200_last_error().decode())
        return g

    def graph_launch(self, g):
        _check(lib().b200_graph_launch(self.h, g), "graph_launch")

    def h2d(self, dptr, arr, nbytes=None):
        arr = np.ascontiguousarray(arr)
        _check(lib().b200_memcpy_h2d(self.h, dptr, arr.ctypes.data, arr.nbytes if nbytes is None else nbytes), "h2d")

    def d2h(self, arr, dptr, nbytes=None):
        assert arr.flags["C_CONTIGUOUS"]
        _check(lib().b200_memcpy_d2h(self.h, arr.ctypes.data, dptr, arr.nbytes if nbytes is None else nbytes), "d2h")

    def memset(self, dptr, value, nbytes):
        _check(lib().b200_memset(self.h, dptr, value, nbytes), "memset")

    def upload(self, t, ne, host, nb=None):
        """contiguous host bytes -> new device tensor of ggml type t and shape ne."""
        host = np.ascontiguousarray(host)
        dt = DeviceTensor(self, t, ne, nb)
        assert host.nbytes <= dt.nbytes, (host.nbytes, dt.nbytes)
        self.h2d(dt.ptr, host)
        self.synchronize()
        return dt

    def empty(self, t, ne, nb=None):
        return DeviceTensor(self, t, ne, nb)

    def download(self, dt, dtype=np.float32):
        out = np.empty(dt.nbytes // np.dtype(dtype).itemsize, dtype=dtype)
        self.d2h(out, dt.ptr, dt.nbytes)
        self.synchronize()
        return out

    # ---- operators (device tensors, asynchronous) ---------------------------------------------------
    def quantize_q8_1(self, x, y_ptr, k, nrows, row_stride=None, variant=Q8_1_X86):
        _check(lib().b200_quantize_q8_1(self.h, x, k * 4 if row_stride is None else row_stride, y_ptr, k, nrows, variant), "quantize_q8_1")

    def mul_mat(self, w, x, dst):
        cw, cx, cd = w.c(), x.c(), dst.c()
        _check(lib().b200_mul_mat(self.h, C.byref(cw), C.byref(cx), C.byref(cd)), "mul_mat")

    def mul_mat_noalloc(self, w, x, dst):
        """b200_mul_mat_noalloc: small-batch quantized MUL_MAT whose q8_1 scratch is a static device array (no allocation of any kind)"""
        cw, cx, cd = w.c(), x.c(), dst.c()
        _check(lib().b200_mul_mat_noalloc(self.h, C.byref(cw), C.byref(cx), C.byref(cd)), "mul_mat_noalloc")

    def mul_mat_multi(self, ws, x, dsts):
        """n MUL_MATs sharing x (wq/wk/wv, gate/up): one launch when the streaming GEMV can take them"""
        n = len(ws)
        cw = [w.c() for w in ws]; cd = [d.c() for d in dsts]; cx = x.c()
        tp = C.POINTER(_Tensor)
        aw = (tp * n)(*[C.pointer(t) for t in cw]); ad = (tp * n)(*[C.pointer(t) for t in cd])
        _check(lib().b200_mul_mat_multi(self.h, aw, n, C.byref(cx), ad), "mul_mat_multi")

    def chain_program(self, phases):
        """phases: dicts {w: [...], dst: [...], in0, in1, normw, out_add, out_norm, out_mul, glu, prologue, epilogue, eps} of device tensors
        -> (b200_chain_phase array, keep-alive list) for decode_chain()"""
        arr = (ChainPhase * len(phases))()
        keep = []
        def ptr(t):
            if t is None:
                return None
            c = t.c(); keep.append(c)
            return C.pointer(c)
        for a, ph in zip(arr, phases):
            a.prologue, a.epilogue, a.n_w, a.eps = ph.get("prologue", X_PLAIN), ph.get("epilogue", Y_PLAIN), len(ph["w"]), ph.get("eps", 0.0)
            for k, (w, d) in enumerate(zip(ph["w"], ph["dst"])):
                a.w[k], a.dst[k] = ptr(w), ptr(d)
            for name in ("in0", "in1", "normw", "out_add", "out_norm", "out_mul", "glu"):
                pt = ptr(ph.get(name))
                if pt is not None:
                    setattr(a, name, pt)
            for k, tl in enumerate(ph.get("tail", ())):          # dicts of b200_chain_rope fields (device pointers as ints), or None
                if tl is not None:
                    r = ChainRope(**tl); keep.append(r)
                    a.tail[k] = C.pointer(r)
        return arr, keep

    def decode_chain(self, program, split=None):
        """b200_decode_chain: the dependent MUL_MAT groups of one token between two attention operators as ONE persistent kernel.
        split (ChainSplit): b200_decode_chain_split -- this rank's rows of every weight, the all-gather inside the hand-over"""
        arr, _keep = program
        if split is not None:
            _check(lib().b200_decode_chain_split(self.h, arr, len(arr), C.byref(split)), "decode_chain_split")
        else:
            _check(lib().b200_decode_chain(self.h, arr, len(arr)), "decode_chain")

    def chain_split(self, world, rank, ll_ptrs, ll_bytes, row_lo):
        """row_lo: per phase, the first row this rank holds of each result"""
        sp = ChainSplit()
        sp.world, sp.rank, sp.ll_bytes = world, rank, ll_bytes
        for r, pt in enumerate(ll_ptrs):
            sp.ll[r] = pt
        for p, los in enumerate(row_lo):
            for k, lo in enumerate(los):
                sp.row_lo[p][k] = lo
        return sp

    def decode_chain_split_supported(self, program, split):
        arr, _keep = program
        return bool(lib().b200_decode_chain_split_supported(self.h, arr, len(arr), C.byref(split)))

    def decode_chain_error(self):
        """0, or the code of the first bounded wait that ran out inside a chain kernel since the last call (synchronizes)"""
        return lib().b200_decode_chain_error(self.h)

    def decode_chain_supported(self, program):
        arr, _keep = program
        return bool(lib().b200_decode_chain_supported(self.h, arr, len(arr)))

    def mul_mat_gate_up_swiglu(self, wg, wu, x, gate, up, glu):
        """gate = wg x, up = wu x, glu = swiglu(gate, up) in one launch of the tensor-core q4_K GEMV (raises when it does not serve the shape)"""
        cs = [t.c() for t in (wg, wu, x, gate, up, glu)]
        _check(lib().b200_mul_mat_gate_up_swiglu(self.h, *[C.byref(t) for t in cs]), "mul_mat_gate_up_swiglu")

    def repack(self, w):
        """load-time repack of a weight tensor in place (no-op for types consumed in GGUF layout); returns w"""
        if w.flags == 0 and lib().b200_repack_supported(w.type, w.ne[0]):
            _check(lib().b200_repack(self.h, w.type, w.ptr, w.ne[0], w.ne[1] * w.ne[2] * w.ne[3], w.nb[1]), "repack")
            w.flags = 1
        return w

    def unrepack(self, w):
        if w.flags == 1:
            _check(lib().b200_unrepack(self.h, w.type, w.ptr, w.ne[0], w.ne[1] * w.ne[2] * w.ne[3], w.nb[1]), "unrepack")
            w.flags = 0
        return w

    # ---- split-mode=row over NVLink peer memory (include/ggml-b200.h: b200_exchange) -----------------------------------------
    def peer_handle(self, ptr):
        buf = C.create_string_buffer(64)
        _check(lib().b200_peer_handle(self.h, C.c_void_p(ptr), C.cast(buf, C.c_void_p)), "peer_handle")
        return buf.raw

    def peer_open(self, handle):
        out = C.c_void_p()
        _check(lib().b200_peer_open(self.h, C.cast(C.c_char_p(handle), C.c_void_p), C.byref(out)), "peer_open")
        return out.value

    def mul_mat_rowsplit(self, w, x, dst, row_lo, ex):
        cw, cx, cd = w.c(), x.c(), dst.c()
        _check(lib().b200_mul_mat_rowsplit(self.h, C.byref(cw), C.byref(cx), C.byref(cd), C.c_int64(row_lo), C.byref(ex)), "mul_mat_rowsplit")

    def mul_mat_rowsplit_multi(self, ws, x, dsts, row_los, ex):
        cws, cds, cx = [w.c() for w in ws], [d.c() for d in dsts], x.c()
        n = len(ws)
        TP = C.POINTER(_Tensor)
        wa = (TP * n)(*[C.pointer(c) for c in cws]); da = (TP * n)(*[C.pointer(c) for c in cds])
        lo = (C.c_int64 * n)(*row_los)
        _check(lib().b200_mul_mat_rowsplit_multi(self.h, wa, n, C.byref(cx), da, lo, C.byref(ex)), "mul_mat_rowsplit_multi")

    def exchange_error(self, ex):
        return lib().b200_exchange_error(self.h, C.byref(ex))

    def mul_mat_id(self, as_, b, ids, dst):
        ca, cb, ci, cd = as_.c(), b.c(), ids.c(), dst.c()
        _check(lib().b200_mul_mat_id(self.h, C.byref(ca), C.byref(cb), C.byref(ci), C.byref(cd)), "mul_mat_id")

    def load_file(self, path, items, chunk_bytes=0, n_buffers=0):
        """b200_load_file: items = [(file_offset, nbytes, device tensor, dst_offset, repack)]; file -> pinned ring -> device with the repack of
        a finished tensor overlapping the next one's transfer.  Returns (seconds, seconds inside pread, bytes, chunks)."""
        arr = (LoadItem * max(1, len(items)))()
        keep = []
        for a, (off, n, t, doff, rp) in zip(arr, items):
            c = t.c(); keep.append(c)
            a.file_offset, a.nbytes, a.dst, a.dst_offset, a.repack = off, n, C.addressof(c), doff, int(rp)
        st = LoadStats()
        _check(lib().b200_load_file(self.h, os.fsencode(path), arr, len(items), chunk_bytes, n_buffers, C.byref(st)), "load_file")
        for (off, n, t, doff, rp), c in zip(items, keep):
            t.flags = c.flags                       # B200_TENSOR_REPACKED where the loader re-laid the tensor out
        return st.seconds, st.read_seconds, st.bytes, st.chunks

    def flash_attn_ext_paged(self, q, k, v, table, page_tokens, pool_tokens, mask, sinks, dst, scale, max_bias=0.0, logit_softcap=0.0):
        """k / v describe the page pool (ne[1] = keys attended); table: device int32 tensor, logical page -> physical page"""
        cq, ck, cv, cd = q.c(), k.c(), v.c(), dst.c()
        cm = mask.c() if mask is not None else None
        cs = sinks.c() if sinks is not None else None
        pg = KvPages(table.ptr, int(table.ne[0]), page_tokens, pool_tokens)
        _check(lib().b200_flash_attn_ext_paged(self.h, C.byref(cq), C.byref(ck), C.byref(cv), C.byref(pg),
                                               C.byref(cm) if cm is not None else None, C.byref(cs) if cs is not None else None,
                                               C.byref(cd), scale, max_bias, logit_softcap), "flash_attn_ext_paged")

    def flash_attn_ext(self, q, k, v, mask, sinks, dst, scale, max_bias=0.0, logit_softcap=0.0):
        cq, ck, cv, cd = q.c(), k.c(), v.c(), dst.c()
        cm = mask.c() if mask is not None else None
        cs = sinks.c() if sinks is not None else None
        _check(lib().b200_flash_attn_ext(self.h, C.byref(cq), C.byref(ck), C.byref(cv),
                                         C.byref(cm) if cm is not None else None, C.byref(cs) if cs is not None else None,
                                         C.byref(cd), scale, max_bias, logit_softcap), "flash_attn_ext")

    # ---- host-buffer convenience calls: what a user with numpy arrays calls (H2D + op + D2H) ----------
    def quantize_q8_1_host(self, x, variant=Q8_1_X86):
        x = np.ascontiguousarray(x, dtype=np.float32)
        k = x.shape[-1]
        nrows = x.size // k
        dx = self.upload(F32, [k, nrows], x)
        dy = self.empty(Q8_1, [k, nrows])
        self.quantize_q8_1(dx.ptr, dy.ptr, k, nrows, variant=variant)
        return self.download(dy, np.uint8)

    def mul_mat_host(self, t, w, K, M, x, N, ne02=1, ne03=1, ne12=1, ne13=1, repack=False):
        dw = self.upload(t, [K, M, ne02, ne03], w)
        if repack:
            self.repack(dw)
        dx = self.upload(F32, [K, N, ne12, ne13], np.ascontiguousarray(x, dtype=np.float32))
        dd = self.empty(F32, [M, N, ne12, ne13])
        self.mul_mat(dw, dx, dd)
        return self.download(dd).reshape(ne13, ne12, N, M)

    def mul_mat_id_host(self, t, as_, K, M, n_expert, b, ne11, n_tokens, ids, n_used):
        da = self.upload(t, [K, M, n_expert], as_)
        db = self.upload(F32, [K, ne11, n_tokens], np.ascontiguousarray(b, dtype=np.float32))
        di = self.upload(I32, [n_used, n_tokens], np.ascontiguousarray(ids, dtype=np.int32))
        dd = self.empty(F32, [M, n_used, n_tokens])
        self.mul_mat_id(da, db, di, dd)
        return self.download(dd).reshape(n_tokens, n_used, M)

    def flash_attn_ext_host(self, q, k, v, mask=None, sinks=None, scale=1.0, max_bias=0.0, softcap=0.0, kv_type=None, kv_shape=None):
        """q[S,H,N,DK] f32, k[Sk,Hk,KV,DK] f16, v[Sk,Hk,KV,DV] f16, mask[Sm,Hm,Nm,KV] f16 -> dst[S,N,H,DV].
        Quantized cache: kv_type = (type of K, type of V) or one type, k / v the raw GGUF blocks (uint8), kv_shape = (Sk, Hk, KV, DK, DV)."""
        q = np.ascontiguousarray(q, dtype=np.float32)
        S, H, N, DK = q.shape
        if kv_type is None:
            k = np.ascontiguousarray(k, dtype=np.float16); v = np.ascontiguousarray(v, dtype=np.float16)
            Sk, Hk, KV, _ = k.shape; DV = v.shape[-1]
            tk = tv = F16
        else:
            tk, tv = kv_type if isinstance(kv_type, (tuple, list)) else (kv_type, kv_type)
            Sk, Hk, KV, _, DV = kv_shape
            k = np.ascontiguousarray(k, dtype=np.uint8); v = np.ascontiguousarray(v, dtype=np.uint8)
        dq = self.upload(F32, [DK, N, H, S], q)
        dk = self.upload(tk, [DK, KV, Hk, Sk], k)
        dv = self.upload(tv, [DV, KV, Hk, Sk], v)
        dm = ds = None
        if mask is not None:
            mask = np.ascontiguousarray(mask, dtype=np.float16); Sm, Hm, Nm, _ = mask.shape
            dm = self.upload(F16, [KV, Nm, Hm, Sm], mask)
        if sinks is not None:
            ds = self.upload(F32, [H], np.ascontiguousarray(sinks, dtype=np.float32))
        dd = self.empty(F32, [DV, H, N, S])
        self.flash_attn_ext(dq, dk, dv, dm, ds, dd, scale, max_bias, softcap)
        return self.download(dd).reshape(S, N, H, DV)


def split_row_range(nrows, world, rank, rounding, tensor_split=None):
    """rows [lo, hi) of an nrows-row weight owned by `rank` (split-mode=row; ggml-cuda.cu:777-803)."""
    lo, hi = C.c_int64(), C.c_int64()
    ts = None if tensor_split is None else np.ascontiguousarray(tensor_split, dtype=np.float32)
    _check(lib().b200_split_row_range(nrows, ts.ctypes.data if ts is not None else None, world, rank, rounding,
                                      C.byref(lo), C.byref(hi)), "split_row_range")
    return lo.value, hi.value


class _LayerAssignment(C.Structure):
    _fields_ = [("layer_id", C.c_int32), ("gpu_id", C.c_int32), ("memory_required", C.c_uint64), ("is_split", C.c_int32), ("n_split", C.c_int32),
                ("split_gpu_ids", C.c_int32 * 8)]


def distribute_layers(strategy, tp_size, n_layers, layer_sizes, gpu_free):
    """host-only layer plan (b200_distribute_layers; the fork's llama_multi_gpu_manager::distribute_layers): returns
    ([(layer_id, gpu_id, memory_required, is_split, [split gpu ids])], bytes charged per GPU)"""
    ls = (C.c_uint64 * max(1, len(layer_sizes)))(*layer_sizes); gf = (C.c_uint64 * max(1, len(gpu_free)))(*gpu_free)
    out = (_LayerAssignment * max(1, n_layers))(); alloc = (C.c_uint64 * max(1, len(gpu_free)))()
    n = lib().b200_distribute_layers(strategy, tp_size, n_layers, ls, len(layer_sizes), len(gpu_free), gf, out, alloc)
    if n < 0:
        _check(n, "distribute_layers")
    return [(a.layer_id, a.gpu_id, a.memory_required, a.is_split, list(a.split_gpu_ids[:a.n_split])) for a in out[:n]], list(alloc[:len(gpu_free)])


def plan_tensor_split(plan, n_gpus):
    """llama_model_params::tensor_split for a plan returned by distribute_layers (b200_plan_tensor_split)"""
    arr = (_LayerAssignment * max(1, len(plan)))()
    for a, (lid, gpu, mem, split, ids) in zip(arr, plan):
        a.layer_id, a.gpu_id, a.memory_required, a.is_split, a.n_split = lid, gpu, mem, split, len(ids)
        for k in range(8):
            a.split_gpu_ids[k] = ids[k] if k < len(ids) else -1
    ts = (C.c_float * n_gpus)()
    _check(lib().b200_plan_tensor_split(arr, len(plan), n_gpus, ts), "plan_tensor_split")
    return list(ts)


def abi_mirrors():
    """which -> ctypes mirror, in b200_abi_sizeof's numbering (tests/test_struct_mirrors.py compares the sizes without a GPU)"""
    return {0: _Tensor, 1: ChainRope, 2: ChainPhase, 3: ChainSplit, 4: KvPages, 5: LoadItem, 6: LoadStats, 7: Exchange, 8: _LayerAssignment}
