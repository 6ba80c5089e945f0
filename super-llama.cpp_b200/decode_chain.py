"""decode_chain.py -- host-side description of one llama layer's decode chain for b200_decode_chain (include/ggml-b200.h).

The chain is everything between two attention operators of a llama decode graph (reference src/models/llama.cpp:40-130):

    attn_out --wo--> + inpSA = ffn_inp --rms_norm * ffn_norm--> ffn_gate / ffn_up --swiglu--> ffn_down --> + ffn_inp = l_out
             --rms_norm * attn_norm(next layer)--> wq / wk / wv

LayerChain owns the device tensors of one such layer and builds the phase program once; tests, scripts/chain_bench.py and bench.py share
it.  Nothing here computes on the host; expected values come from the caller's oracle.
"""
import numpy as np

import ggml_b200 as B


class LayerChain:
    """Weights and activation columns of one layer; `weights` maps name -> (type, host bytes) for wo, gate, up, down, q, k, v."""

    NAMES = ("wo", "gate", "up", "down", "q", "k", "v")

    def __init__(self, ctx, n_embd, n_ff, n_kv, weights, norm_ffn, norm_attn, eps=1e-5, repack=True, tail="qkv", norm_outputs=True, rope=None, split=None):
        """norm_outputs=False: the RMS_NORM and MUL node values are dead, as in the real llama graph where only the matmuls read them (the
        plugin passes NULL for them, b200_backend.cpp try_chain) -- the shape the kernel's owner-side hand-over takes.
        split = dict(world, rank, ll_ptrs, ll_bytes, ranges={name: (lo, hi)}[, alloc]): the row-split form (b200_decode_chain_split); `weights` then
        hold this rank's rows [lo, hi) of every matrix, the columns stay full size (alloc: where the peers can reach... only the hand-over
        areas ll_ptrs have to be; the columns are private)"""
        self.ctx, self.n_embd, self.n_ff, self.n_kv, self.eps = ctx, n_embd, n_ff, n_kv, eps
        shape = {"wo": (n_embd, n_embd), "gate": (n_embd, n_ff), "up": (n_embd, n_ff), "down": (n_ff, n_embd),
                 "q": (n_embd, n_embd), "k": (n_embd, n_kv), "v": (n_embd, n_kv)}
        self.split = None
        self.w = {}
        for name in self.NAMES:
            if name not in weights:
                continue
            t, host = weights[name]
            K, M = shape[name]
            if split is not None:
                lo, hi = split["ranges"][name]; M = hi - lo
            dw = ctx.upload(t, [K, M], host)
            if repack:
                ctx.repack(dw)                      # q6_K -> group-SoA (what the plugin's load-time hook does); q4_K stays raw
            self.w[name] = dw
        f32 = B.F32
        self.norm_ffn = ctx.upload(f32, [n_embd], np.ascontiguousarray(norm_ffn, dtype=np.float32))
        self.norm_attn = ctx.upload(f32, [n_embd], np.ascontiguousarray(norm_attn, dtype=np.float32))
        col = lambda n: ctx.empty(f32, [n])
        self.attn_out, self.inp_sa = col(n_embd), col(n_embd)                       # inputs
        self.wo_out, self.ffn_inp, self.ffn_normed, self.ffn_x = col(n_embd), col(n_embd), col(n_embd), col(n_embd)
        self.gate, self.up, self.glu = col(n_ff), col(n_ff), col(n_ff)
        self.down_out, self.l_out, self.attn_normed, self.attn_x = col(n_embd), col(n_embd), col(n_embd), col(n_embd)
        self.q, self.k, self.v = col(n_embd), col(n_kv), col(n_kv)
        ph = [
            dict(w=[self.w["wo"]], dst=[self.wo_out], in0=self.attn_out),
            dict(w=[self.w["gate"], self.w["up"]], dst=[self.gate, self.up], glu=self.glu, epilogue=B.Y_SWIGLU, prologue=B.X_ADD_NORM_MUL, eps=eps,
                 in0=self.wo_out, in1=self.inp_sa, normw=self.norm_ffn, out_add=self.ffn_inp,
                 out_norm=self.ffn_normed if norm_outputs else None, out_mul=self.ffn_x if norm_outputs else None),
            dict(w=[self.w["down"]], dst=[self.down_out], in0=self.glu),
        ]
        if tail == "qkv":
            ph.append(dict(w=[self.w["q"], self.w["k"], self.w["v"]], dst=[self.q, self.k, self.v], prologue=B.X_ADD_NORM_MUL, eps=eps,
                           in0=self.down_out, in1=self.ffn_inp, normw=self.norm_attn, out_add=self.l_out,
                           out_norm=self.attn_normed if norm_outputs else None, out_mul=self.attn_x if norm_outputs else None))
            if rope is not None:
                # what the llama graph does to wq / wk / wv's rows next (b200_chain_rope): ROPE(Q), ROPE(K) + the f16 K row of the cache, the f16 V row.
                # rope = dict(head_dim, n_dims, theta_scale, freq_scale, attn_factor, kv_size[, in_place])
                hd, kvs = rope["head_dim"], rope["kv_size"]
                self.rope = rope
                self.pos = ctx.malloc(4); self.kidx = ctx.malloc(8); self.vidx = ctx.malloc(8)
                self.q_rope = self.q if rope.get("in_place") else col(n_embd)
                self.k_rope = self.k if rope.get("in_place") else col(n_kv)
                self.k_cache = ctx.empty(B.F16, [n_kv, kvs]); self.v_cache = ctx.empty(B.F16, [n_kv, kvs])
                common = dict(head_dim=hd, n_dims=rope["n_dims"], freq_scale=rope["freq_scale"], theta_scale=rope["theta_scale"], attn_factor=rope["attn_factor"], pos=self.pos)
                ph[-1]["tail"] = [dict(kind=B.T_ROPE, out=self.q_rope.ptr, **common),
                                  dict(kind=B.T_ROPE_CACHE, out=self.k_rope.ptr, cache_idx=self.kidx, cache=self.k_cache.ptr, cache_nb1=n_kv * 2, **common),
                                  dict(kind=B.T_CACHE, cache_idx=self.vidx, cache=self.v_cache.ptr, cache_nb1=n_kv * 2)]
        self.phases = ph
        self.norm_outputs = norm_outputs
        self.program = ctx.chain_program(ph)
        if split is not None:
            lo = lambda n: split["ranges"][n][0]
            row_lo = [[lo("wo")], [lo("gate"), lo("up")], [lo("down")]] + ([[lo("q"), lo("k"), lo("v")]] if tail == "qkv" else [])
            self.split = ctx.chain_split(split["world"], split["rank"], split["ll_ptrs"], split["ll_bytes"], row_lo)

    def supported(self):
        return self.ctx.decode_chain_split_supported(self.program, self.split) if self.split is not None else self.ctx.decode_chain_supported(self.program)

    def weight_bytes(self):
        return sum(w.nbytes for w in self.w.values())

    def set_inputs(self, attn_out, inp_sa):
        self.ctx.h2d(self.attn_out.ptr, np.ascontiguousarray(attn_out, dtype=np.float32))
        self.ctx.h2d(self.inp_sa.ptr, np.ascontiguousarray(inp_sa, dtype=np.float32))
        self.ctx.note_write(self.attn_out); self.ctx.note_write(self.inp_sa)

    def set_position(self, pos, cache_row):
        self.ctx.h2d(self.pos, np.int32([pos])); self.ctx.h2d(self.kidx, np.int64([cache_row])); self.ctx.h2d(self.vidx, np.int64([cache_row]))

    def run(self):
        self.ctx.decode_chain(self.program, self.split)

    OUTPUTS = ("wo_out", "ffn_inp", "ffn_normed", "ffn_x", "gate", "up", "glu", "down_out", "l_out", "attn_normed", "attn_x", "q", "k", "v")

    DEAD = ("ffn_normed", "ffn_x", "attn_normed", "attn_x")

    def outputs(self):
        return {n: self.ctx.download(getattr(self, n)) for n in self.OUTPUTS if self.norm_outputs or n not in self.DEAD}

    def clear_outputs(self, byte=0xff):
        for n in self.OUTPUTS:
            t = getattr(self, n)
            self.ctx.memset(t.ptr, byte, t.nbytes)
