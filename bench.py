#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config: llama-bench style tg128 (and pp512) tok/s of
Llama-3-8B Q4_K_M (random-init GGUF block weights, exact Q4_K_M tensor-type mix) on N B200s.

One "step" = one pass of the hot path for one decode token: every GGML_OP_MUL_MAT node of the llama graph
(wq, wk, wv, wo, ffn_gate, ffn_up, ffn_down per layer + the output head) and the per-layer GGML_OP_FLASH_ATTN_EXT
over the f16 KV-cache views, issued through the C-ABI of libggml-b200.so (include/ggml-b200.h) and replayed as a
CUDA graph.  `value` is timed with CUDA events with all inputs resident in HBM; `e2e` adds, every step, the
host->device copy of the token's embedding row from pinned memory and the device->host read of the logits.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--depth D] [--no-pp]

Multi-GPU (torchrun, one rank per GPU): split-mode=row tensor parallelism -- every rank owns rows
[row_low, row_high) of each weight (b200_split_row_range, reference ggml-cuda.cu:777-803).  The all-gather lives inside
the kernels: the decode chain's hand-over between its phases (b200_decode_chain_split: row owners publish q8_1 records
as (word, tag) pairs into every rank's area over NVLink peer memory), the GEMV epilogue for layer 0's wq / wk / wv;
torch.distributed carries the IPC handles, the barriers and the max-over-ranks of the timings (B200_BENCH_NCCL=1: one
NCCL all-gather per matmul, the baseline).  Scaling is "strong" (one model, one token stream).
--impl reference times the reference's own ggml-cpu implementation of the same ops on the host cores
(oracle/_ref, compiled in place from /root/reference; the oracle port if that is absent).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "super-llama.cpp_b200"))

F32, F16, Q4_0, Q8_0, Q4_K, Q6_K, I32 = 0, 1, 2, 8, 12, 14, 26
BLCK = {Q4_K: 256, Q6_K: 256, Q4_0: 32, Q8_0: 32}
TSIZE = {Q4_K: 144, Q6_K: 210, Q4_0: 18, Q8_0: 34}

# Llama-3-8B: n_embd, n_ff, n_head, n_head_kv, head_dim, n_layer, n_vocab
LLAMA3_8B = dict(name="llama3-8b", n_embd=4096, n_ff=14336, n_head=32, n_head_kv=8, head_dim=128, n_layer=32, n_vocab=128256)
# BASELINE.json configs[3]: the 8-GPU row-split scale point (not the default bench line)
LLAMA3_70B = dict(name="llama3-70b", n_embd=8192, n_ff=28672, n_head=64, n_head_kv=8, head_dim=128, n_layer=80, n_vocab=128256)
# BASELINE.json configs[0] (the reference's CPU-runnable case) and configs[4] (MUL_MAT_ID: 8 experts, 2 used per token)
TINYLLAMA = dict(name="tinyllama-1.1b", n_embd=2048, n_ff=5632, n_head=32, n_head_kv=4, head_dim=64, n_layer=22, n_vocab=32000)
MIXTRAL = dict(name="mixtral-8x7b", n_embd=4096, n_ff=14336, n_head=32, n_head_kv=8, head_dim=128, n_layer=32, n_vocab=32000, n_expert=8, n_expert_used=2)
MODELS = {"8b": LLAMA3_8B, "70b": LLAMA3_70B, "tinyllama": TINYLLAMA, "mixtral": MIXTRAL}


def use_more_bits(i, n):       # src/llama-quant.cpp:186-188
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def q4_k_m_layer_types(cfg, quant="q4_k_m"):
    """tensor types llama-quantize picks for Q4_K_M (src/llama-quant.cpp:207-227,302-303,358-364); Q8_0 / Q4_0 files are uniform
    (Q4_0 keeps a q6_K output head, llama-quant.cpp:207-227)"""
    n = cfg["n_layer"]
    if quant in ("q8_0", "q4_0"):
        t = Q8_0 if quant == "q8_0" else Q4_0
        return [dict(wq=t, wk=t, wv=t, wo=t, gate=t, up=t, down=t) for _ in range(n)], (Q8_0 if quant == "q8_0" else Q6_K)
    out = []
    for i in range(n):
        more = use_more_bits(i, n)
        out.append(dict(wq=Q4_K, wk=Q4_K, wv=Q6_K if more else Q4_K, wo=Q4_K, gate=Q4_K, up=Q4_K, down=Q6_K if more else Q4_K))
    return out, Q6_K   # output.weight -> q6_K


def weight_shapes(cfg):
    e, f, kv = cfg["n_embd"], cfg["n_ff"], cfg["n_head_kv"] * cfg["head_dim"]
    return dict(wq=(e, e), wk=(e, kv), wv=(e, kv), wo=(e, e), gate=(e, f), up=(e, f), down=(f, e))   # (K, M)


def pretty(cfg):
    return {"llama3-8b": "Llama-3-8B", "llama3-70b": "Llama-3-70B", "tinyllama-1.1b": "TinyLlama-1.1B", "mixtral-8x7b": "Mixtral-8x7B"}[cfg["name"]]


def wbytes(t, K, M):
    return K // BLCK[t] * TSIZE[t] * M


def random_blocks(rng, t, nblocks):
    raw = rng.integers(0, 256, size=(nblocks, TSIZE[t]), dtype=np.uint8)
    if t == Q4_K:
        raw[:, 0:2] = (rng.uniform(0.1, 1, nblocks) * 4e-4).astype(np.float16).view(np.uint8).reshape(nblocks, 2)
        raw[:, 2:4] = (rng.uniform(0.1, 1, nblocks) * 4e-4).astype(np.float16).view(np.uint8).reshape(nblocks, 2)
    elif t in (Q4_0, Q8_0):
        raw[:, 0:2] = (rng.uniform(-1, 1, nblocks) * (2e-3 if t == Q4_0 else 2e-4)).astype(np.float16).view(np.uint8).reshape(nblocks, 2)
    else:
        raw[:, 208:210] = (rng.uniform(-1, 1, nblocks) * 2e-4).astype(np.float16).view(np.uint8).reshape(nblocks, 2)
    return raw.reshape(-1)


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)"""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        pw = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None, "reasons": reasons,
                "power_w_max": max(pw) if pw else None, "samples": len(self.rows)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops", 1590.0), d.get("bf16_tflops_sustained", 1400.0), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


# ------------------------------------------------------------------------------------------------------------
class HotPathModel:
    """The hot-path nodes of one llama decode/prefill step over device-resident random-init Q4_K_M weights.
    rank/world: split-mode=row shard (rows [lo, hi) of every weight); world == 1 is the whole model."""

    def __init__(self, B, ctx, cfg, n_tokens, depth, rank=0, world=1, seed=1234, alloc=None, dist=None, quant="q4_k_m", chain=True, peer=None):
        self.B, self.ctx, self.cfg, self.N = B, ctx, cfg, n_tokens
        self.quant = quant
        self.moe = cfg.get("n_expert", 0)
        self.replicas = 1
        if self.moe and world > 1:
            # MUL_MAT_ID does not shard (the reference asserts against split buffers, ggml-cuda.cu:2274): every rank runs its own replica
            # of the whole model on its own token stream ("replicas / layer split first"); no data-path collective
            self.replicas, world, rank, dist = world, 1, 0, None
        self.rank, self.world = rank, world
        from row_split import RowSplit                      # partition rule + gather shared with tests/test_row_split_gloo.py
        self.rs = RowSplit(dist, world, rank, 32)
        # row split: every rank draws the SAME stream, builds the same virtual full-size weights and keeps rows [lo, hi) of each; the activations
        # are replicated.  (Replicas of a MoE model are independent jobs: their own streams.)
        rng = np.random.default_rng(seed + (rank if self.replicas > 1 else 0))
        types, out_type = q4_k_m_layer_types(cfg, quant)
        shapes = weight_shapes(cfg)
        used = {t for lt in types for t in lt.values()} | {out_type}
        pool = {t: random_blocks(rng, t, (48 << 20) // TSIZE[t]) for t in used}
        self.weight_bytes = 0          # bytes of weights one token reads (MoE: the experts it routes to)

        def make_weight(t, K, M, ne=1, whole=False):
            lo, hi = (0, M) if whole else self.rs.row_range(M)
            ml = hi - lo
            w = ctx.empty(t, [K, ml] if ne == 1 else [K, ml, ne])
            nb, blk = w.nbytes, TSIZE[t]
            full = wbytes(t, K, M) * ne           # the unsplit weight; this rank keeps bytes [a, b) of it
            a = wbytes(t, K, lo) if ne == 1 else 0
            b = a + nb
            off = 0
            while off < full:                     # tile the random pool at random block-aligned offsets
                start = int(rng.integers(0, pool[t].size // blk // 2)) * blk
                n = min(full - off, pool[t].size - start)
                x0, x1 = max(off, a), min(off + n, b)
                if x0 < x1:
                    ctx.h2d(w.ptr + (x0 - a), pool[t][start + (x0 - off): start + (x1 - off)])
                off += n
            self.weight_bytes += nb if ne == 1 else nb * cfg["n_expert_used"] // ne
            ctx.repack(w)                         # load-time repack (q6_K -> 16-byte aligned group layout), as set_tensor does
            return w, lo, hi

        # B200_BENCH_WHOLE_QKV=1 (measurement only): under the row-split chain wq / wk / wv stay WHOLE on every rank (14 MB of 131 per 8B layer), so the
        # one exchange of a layer that is a pure gather -- attention is replicated and needs q / k / v in full -- disappears.  Measured: 695 vs 684
        # tok/s on two GPUs, 756.8 vs 756.7 on eight: the gather hides behind the launch of the attention kernel; not the default.
        self.whole_qkv = world > 1 and peer is not None and bool(chain) and n_tokens == 1 and not self.moe and bool(os.environ.get("B200_BENCH_WHOLE_QKV")) and \
                         not os.environ.get("B200_BENCH_NO_CHAIN") and not os.environ.get("B200_BENCH_NO_SPLIT_CHAIN")
        self.layers = []
        for lt in types:
            self.layers.append({name: make_weight(lt[name], *shapes[name], ne=(self.moe if self.moe and name in ("gate", "up", "down") else 1),
                                                  whole=self.whole_qkv and name in ("wq", "wk", "wv")) for name in shapes})
        self.output = make_weight(out_type, cfg["n_embd"], cfg["n_vocab"])
        ctx.synchronize()

        e, f, hd, nh, nkv = cfg["n_embd"], cfg["n_ff"], cfg["head_dim"], cfg["n_head"], cfg["n_head_kv"]
        kvd = nkv * hd
        N = n_tokens
        self.n_kv = ((depth + N + 255) // 256) * 256                      # src/llama-kv-cache.cpp:992-1006
        self.kv_size = self.n_kv
        alloc = alloc or (lambda t, ne: ctx.empty(t, ne))
        # full-size activation buffers (what every rank sees after the all-gather)
        self.x = alloc(F32, [e, N]); self.h = alloc(F32, [e, N]); self.ff = alloc(F32, [f, N])
        self.q = alloc(F32, [e, N]); self.k = alloc(F32, [kvd, N]); self.v = alloc(F32, [kvd, N])
        self.attn = alloc(F32, [e, N]); self.o = alloc(F32, [e, N]); self.g = alloc(F32, [f, N]); self.u = alloc(F32, [f, N])
        self.d = alloc(F32, [e, N])
        if self.moe:
            nu = cfg["n_expert_used"]
            self.g3 = ctx.empty(F32, [f, nu, N]); self.u3 = ctx.empty(F32, [f, nu, N]); self.ff3 = ctx.empty(F32, [f, nu, N]); self.d3 = ctx.empty(F32, [e, nu, N])
            ctx.h2d(self.ff3.ptr, (rng.standard_normal(f * nu * N) * 0.3).astype(np.float32))
            self.h3 = self.h.view([e, 1, N, 1], [4, e * 4, e * 4, e * 4 * N])
            ids = np.stack([rng.permutation(self.moe)[:nu] for _ in range(N)]).astype(np.int32)      # top-k routing: distinct experts per token
            self.ids = ctx.upload(I32, [nu, N], ids)
        self.logits = alloc(F32, [cfg["n_vocab"], 1])
        for t, scale in ((self.x, 1.0), (self.h, 1.0), (self.ff, 0.3), (self.attn, 0.5)):
            ctx.h2d(t.ptr, (rng.standard_normal(t.ne[0] * t.ne[1]) * scale).astype(np.float32))
        # KV cache: one [n_embd_k_gqa, kv_size] f16 buffer per layer for K and for V (heads interleaved per token row)
        kvh = (rng.standard_normal(kvd * min(self.kv_size, 4096)) * 0.5).astype(np.float16)
        self.kc, self.vc = [], []
        for _ in range(cfg["n_layer"]):
            for lst in (self.kc, self.vc):
                c = ctx.empty(F16, [kvd, self.kv_size])
                for off in range(0, c.nbytes, kvh.nbytes):
                    ctx.h2d(c.ptr + off, kvh[: min(kvh.size, (c.nbytes - off) // 2)])
                lst.append(c)
        tok, head = kvd * 2, hd * 2
        self.kviews = [c.view([hd, self.n_kv, nkv, 1], [2, tok, head, tok * self.kv_size]) for c in self.kc]
        self.vviews = [c.view([hd, self.n_kv, nkv, 1], [2, tok, head, tok * self.kv_size]) for c in self.vc]
        nm = max(8, ((N + 63) // 64) * 64)                                 # mask rows padded (GGML_KQ_MASK_PAD)
        mask = np.zeros((nm, self.n_kv), dtype=np.float16)
        for i in range(N):                                                  # causal: token i of the batch sees depth + i + 1 keys
            mask[i, depth + i + 1:] = -np.inf
        self.mask = ctx.upload(F16, [self.n_kv, nm], mask)
        self.qview = self.q.view([hd, N, nh, 1], [4, e * 4, hd * 4, e * 4 * N])          # [head_dim, n_tokens, n_head]
        self.attn4 = self.attn.view([hd, nh, N, 1], [4, hd * 4, e * 4, e * 4 * N])       # dst [DV, H, N]
        self.scale = 1.0 / float(np.sqrt(hd))
        self.kv_bytes_per_layer = 2 * self.n_kv * kvd * 2
        ctx.synchronize()
        self.gather = None          # set by the multi-GPU driver: gather(full_tensor, lo, hi) after a sharded matmul
        self.chains = None
        self.peer = peer
        self.chain_splits = None
        # chain: True = the default form on one GPU, the row-split form (b200_decode_chain_split: the all-gather inside the kernel's hand-over)
        # over the peer exchange's arenas on several; "split" = the row-split form's arithmetic on ONE GPU (what the split check compares with)
        split_form = (world > 1 and peer is not None) or chain == "split"
        if chain and N == 1 and (world == 1 or peer is not None) and not self.moe and not os.environ.get("B200_BENCH_NO_CHAIN") and \
           not (world > 1 and os.environ.get("B200_BENCH_NO_SPLIT_CHAIN")):
            self._build_chains(rng, alloc, split_form)

    def reset_state(self):
        """the residual stream the chains carry from step to step, back to its first value (so that two models can be compared after any number of steps)"""
        if self.chains is not None:
            self.ctx.h2d(self.resid[0].ptr, self.resid0_host)

    def _build_chains(self, rng, alloc=None, split_form=False):
        """One b200_decode_chain program per layer: everything between two attention operators of the llama graph (src/models/llama.cpp:40-130)
        -- wo, +residual, ffn_norm, ffn_gate / ffn_up, SwiGLU, ffn_down, +ffn_inp, attn_norm of the next layer, its wq / wk / wv (the last
        layer: output_norm and the output head) -- with the data really flowing from node to node.  None when the kernel declines the model
        (other weight types): then the matmuls go out one launch per group as before."""
        B, ctx, cfg = self.B, self.ctx, self.cfg
        e, f = cfg["n_embd"], cfg["n_ff"]
        col = lambda n: ctx.empty(F32, [n])
        normw = lambda: ctx.upload(F32, [e], rng.uniform(0.5, 1.5, e).astype(np.float32))
        self.resid = [col(e), col(e)]                       # l_out of layer il - 1 / il (ping-pong); resid[0] starts as the token's embedding row
        self.resid0_host = rng.standard_normal(e).astype(np.float32)
        ctx.h2d(self.resid[0].ptr, self.resid0_host)
        ll_ptrs = ll_bytes = None
        if split_form:
            # the hand-over areas: inside the exchange arena (same offset on every rank, mapped by every rank) / a plain zeroed block on one GPU
            ll_bytes = int(B.lib().b200_decode_chain_split_bytes(cfg["n_vocab"]))
            ll = (alloc or (lambda t, ne: ctx.empty(t, ne)))(F32, [ll_bytes // 4])
            B._check(B.lib().b200_memset(ctx.h, ll.ptr, 0, ll_bytes), "memset")
            ll_ptrs = [ll.ptr] if self.world == 1 else [int(self.peer.ex.base[r]) + (ll.ptr - self.peer.base) for r in range(self.world)]
            self._ll = ll
        splits = []
        self.ffn_inp = col(e)
        q1, k1, v1 = (t.view([t.ne[0], 1, 1, 1], [4, t.ne[0] * 4, t.ne[0] * 4, t.ne[0] * 4]) for t in (self.q, self.k, self.v))
        cols = {n: getattr(self, n).view([getattr(self, n).ne[0], 1, 1, 1], [4] + [getattr(self, n).ne[0] * 4] * 3) for n in ("attn", "o", "g", "u", "ff", "d", "logits")}
        chains = []
        for il, L in enumerate(self.layers):
            last = il == len(self.layers) - 1
            tail_w = [self.output[0]] if last else [self.layers[il + 1][n][0] for n in ("wq", "wk", "wv")]
            tail_d = [cols["logits"]] if last else [q1, k1, v1]
            ph = [dict(w=[L["wo"][0]], dst=[cols["o"]], in0=cols["attn"]),
                  dict(w=[L["gate"][0], L["up"][0]], dst=[cols["g"], cols["u"]], glu=cols["ff"], epilogue=B.Y_SWIGLU, prologue=B.X_ADD_NORM_MUL, eps=1e-5,
                       in0=cols["o"], in1=self.resid[il % 2], normw=normw(), out_add=self.ffn_inp),
                  dict(w=[L["down"][0]], dst=[cols["d"]], in0=cols["ff"]),
                  dict(w=tail_w, dst=tail_d, prologue=B.X_ADD_NORM_MUL, eps=1e-5, in0=cols["d"], in1=self.ffn_inp, normw=normw(), out_add=self.resid[(il + 1) % 2])]
            prog = ctx.chain_program(ph)
            if split_form:
                tail_lo = [self.output[1]] if last else [self.layers[il + 1][n][1] for n in ("wq", "wk", "wv")]
                sp = ctx.chain_split(self.world, self.rank, ll_ptrs, ll_bytes, [[L["wo"][1]], [L["gate"][1], L["up"][1]], [L["down"][1]], tail_lo])
                if not ctx.decode_chain_split_supported(prog, sp):
                    if self.rank == 0:
                        print("bench: row-split decode chain declined:", B.lib().b200_last_error().decode(), file=sys.stderr)
                    return
                splits.append(sp)
            elif not ctx.decode_chain_supported(prog):
                return
            chains.append(prog)
        self.chains = chains
        self.chain_splits = splits if split_form else None
        self._keep = (q1, k1, v1, cols)

    def _mm(self, w, x, dst):
        wt, lo, hi = w
        if self.world == 1:
            self.ctx.mul_mat(wt, x, dst)
        elif getattr(self, "peer", None) is not None and dst.ne[1] == 1:
            self.peer.mul_mat(wt, x, dst, lo)              # GEMV + all-gather in one kernel (peer stores), then the arrival wait
        else:
            M, N = dst.ne[0], dst.ne[1]
            local = dst.view([hi - lo, N, 1, 1], [4, M * 4, M * 4 * N, M * 4 * N], offset=lo * 4)   # rows [lo, hi) of every column
            self.ctx.mul_mat(wt, x, local)
            self.gather(dst, M, N)

    def _mm_group(self, ws, x, dsts):
        """MUL_MATs that share src1: one fused launch per run of equal weight types (the plugin fuses the same nodes)"""
        if self.world > 1 and all(w[1] == 0 and w[2] == d.ne[0] for w, d in zip(ws, dsts)):
            pass                                             # whole on every rank: computed locally, like on one GPU (below)
        elif self.world > 1:
            if getattr(self, "peer", None) is not None and dsts[0].ne[1] == 1:
                self.peer.mul_mat_multi([w[0] for w in ws], x, dsts, [w[1] for w in ws])   # one launch per run of equal types, exchange included
                return
            for w, d in zip(ws, dsts):
                self._mm(w, x, d)
            return
        i = 0
        while i < len(ws):
            j = i + 1
            while j < len(ws) and ws[j][0].type == ws[i][0].type:
                j += 1
            if j - i > 1:
                self.ctx.mul_mat_multi([w[0] for w in ws[i:j]], x, dsts[i:j])
            else:
                self.ctx.mul_mat(ws[i][0], x, dsts[i])
            i = j

    def step(self, attention=True):
        """issue every MUL_MAT and FLASH_ATTN_EXT node of one llama graph evaluation on the ctx stream"""
        c = self.ctx
        if self.chains is not None:
            # the persistent decode chain: wq / wk / wv of layer 0, then per layer FLASH_ATTN_EXT + ONE kernel for everything up to the next
            # layer's wq / wk / wv (the last one ends with the output head)
            L0 = self.layers[0]
            c.note_write(self.x)
            self._mm_group([L0["wq"], L0["wk"], L0["wv"]], self.x, [self.q, self.k, self.v])
            for il in range(len(self.layers)):
                if attention:
                    c.flash_attn_ext(self.qview, self.kviews[il], self.vviews[il], self.mask, None, self.attn4, self.scale)
                c.decode_chain(self.chains[il], self.chain_splits[il] if self.chain_splits else None)
            return
        fuse = True      # wq/wk/wv and gate/up share src1: one launch per run of equal weight types (GEMV at n=1, GEMM at n=512)
        # In the real graph every matmul input is rewritten by the glue op in front of it (RMS_NORM, attention, SwiGLU), so its cached
        # q8_1 / bf16 image is stale each time.  The synthetic step reuses fixed input buffers: drop the cache at every point where the
        # real graph produces a new activation, so that each activation is converted exactly as often as in the real graph
        # (once per group of matmuls that share src1).
        fresh = c.note_write
        for il, L in enumerate(self.layers):
            fresh(self.x)
            if fuse:
                self._mm_group([L["wq"], L["wk"], L["wv"]], self.x, [self.q, self.k, self.v])
            else:
                self._mm(L["wq"], self.x, self.q); self._mm(L["wk"], self.x, self.k); self._mm(L["wv"], self.x, self.v)
            if attention:
                c.flash_attn_ext(self.qview, self.kviews[il], self.vviews[il], self.mask, None, self.attn4, self.scale)
            fresh(self.attn)
            self._mm(L["wo"], self.attn, self.o)
            fresh(self.h)
            if self.moe:                                  # build_moe_ffn: MUL_MAT_ID x 3 (src/llama-graph.cpp), router and glue not on the hot path
                c.mul_mat_id(L["gate"][0], self.h3, self.ids, self.g3); c.mul_mat_id(L["up"][0], self.h3, self.ids, self.u3)
                fresh(self.ff3)
                c.mul_mat_id(L["down"][0], self.ff3, self.ids, self.d3)
            elif fuse:
                self._mm_group([L["gate"], L["up"]], self.h, [self.g, self.u])
                fresh(self.ff)
                self._mm(L["down"], self.ff, self.d)
            else:
                self._mm(L["gate"], self.h, self.g); self._mm(L["up"], self.h, self.u)
                self._mm(L["down"], self.ff, self.d)
        fresh(self.x)
        last = self.x if self.N == 1 else self.x.view([self.cfg["n_embd"], 1, 1, 1], self.x.nb, offset=(self.N - 1) * self.cfg["n_embd"] * 4)
        self._mm(self.output, last, self.logits)          # output head runs for the last row only (src/models/llama.cpp:97-98)

    # ---- algorithmic work --------------------------------------------------------------
    def matmul_flops(self):
        shapes = weight_shapes(self.cfg)
        per_layer = sum(2 * K * M * (self.cfg["n_expert_used"] if self.moe and name in ("gate", "up", "down") else 1) for name, (K, M) in shapes.items())
        return self.N * per_layer * self.cfg["n_layer"] + 2 * self.cfg["n_embd"] * self.cfg["n_vocab"]

    def attn_flops(self):
        return 2 * self.cfg["n_head"] * self.N * 2 * self.cfg["head_dim"] * self.n_kv * self.cfg["n_layer"]


def traffic_per_launch(algorithmic_bytes_per_launch):
    """(DRAM bytes per launch of the dominant kernel, where the figure comes from).  It is NOT a live counter: ncu cannot run inside the
    timed region, so the ratio dram__bytes_read.sum + dram__bytes_write.sum over algorithmic bytes of the newest committed `ncu --set full`
    capture (profiles/*_traffic.json) is applied to this run's algorithmic bytes per launch.  (None, why) when no capture is committed."""
    try:
        import re
        cands = sorted((f for f in os.listdir(os.path.join(ROOT, "profiles")) if f.endswith("_traffic.json")), key=lambda f: [int(x) for x in re.findall(r"\d+", f)[:2]])
        d = json.load(open(os.path.join(ROOT, "profiles", cands[-1])))
        return algorithmic_bytes_per_launch * float(d["traffic_over_algorithmic"]), f"profiles/{cands[-1]} ratio {float(d['traffic_over_algorithmic']):.4f} (ncu --set full of {d.get('kernel', 'the decode GEMV')}) x this run's algorithmic bytes; not measured live"
    except Exception as ex:
        return None, f"no committed capture ({type(ex).__name__})"


# ------------------------------------------------------------------------------------------------------------
def cpu_baseline(cfg, depth, budget_s=20.0, steps=5, warmup=3, quant="q4_k_m"):
    """The reference's ggml-cpu implementation of the same ops, timed on this box's host cores (oracle/_ref compiled
    in place from /root/reference; the plain-C oracle port if that is absent).  Bounded sample: ONE transformer layer
    of each tensor-type variant (7 MUL_MAT + FLASH_ATTN_EXT at n=1) + the output head, extrapolated to n_layer layers.
    Returns (tok/s, kind, cores, sample description)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as O
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(7)
    shapes = weight_shapes(cfg)
    if cfg.get("n_expert"):
        raise RuntimeError("no CPU baseline for the MUL_MAT_ID workload in this harness")
    types, out_type = q4_k_m_layer_types(cfg, quant)
    n_kv = ((depth + 1 + 255) // 256) * 256
    isa = "avx2"
    try:
        flags = open("/proc/cpuinfo").read()
        if all(n in flags for n in ("avx512f", "avx512_vnni", "avx512_bf16", "avx512vbmi", "avx_vnni")) and O.have_reference("avx512"):
            isa = "avx512"                                  # the same reference sources compiled for the host's AVX-512 VNNI/BF16 (oracle/Makefile ISA=avx512)
    except OSError:
        pass
    if O.have_reference(isa):
        ref = O.Reference(isa)
        ops = {}

        def mm(t, K, M):
            key = (t, K, M)
            if key not in ops:
                op = O.ref_mm_op(ref, t, K, M, 1, cores)
                w = op.input(0)
                blk = random_blocks(rng, t, min(w.size // TSIZE[t], 1 << 16))
                for off in range(0, w.size, blk.size):
                    n = min(blk.size, w.size - off); w[off:off + n] = blk[:n]
                op.input(1).view(np.float32)[:] = rng.standard_normal(K).astype(np.float32)
                ops[key] = op
            return ops[key]

        fa = O.ref_fa_op(ref, cfg["head_dim"], cfg["head_dim"], 1, cfg["n_head"], n_kv, cfg["n_head_kv"], 8, 0.0884, cores)
        fa.input(0).view(np.float32)[:] = rng.standard_normal(fa.input(0).size // 4).astype(np.float32)
        for i in (1, 2):
            fa.input(i).view(np.float16)[:] = (rng.standard_normal(fa.input(i).size // 2) * 0.5).astype(np.float16)
        variants = {}
        for lt in types:
            variants.setdefault(tuple(sorted(lt.items())), 0)
            variants[tuple(sorted(lt.items()))] += 1
        total_us = 0.0
        t_start = time.time()
        head = mm(out_type, cfg["n_embd"], cfg["n_vocab"])
        groups = [([mm(t, *shapes[name]) for name, t in var] + [fa], count) for var, count in variants.items()] + [([head], 1)]
        done = steps
        for layer_ops, count in groups:
            for _ in range(max(warmup, 1)):                         # warm-up: thread pool + page faults
                for op in layer_ops:
                    op.run()
            acc, n = 0.0, 0
            for _ in range(steps):                                  # a step = this sample run once; the mean over the steps is reported
                acc += sum(op.run() for op in layer_ops); n += 1
                if time.time() - t_start > budget_s:
                    break
            done = min(done, n)
            total_us += acc / n * count
        for op in list(ops.values()) + [fa]:
            op.free()
        return 1e6 / total_us, "reference", cores, (f"one layer of each of the {len(variants)} {quant.upper()} tensor-type variants (7 MUL_MAT + FLASH_ATTN_EXT, n=1, "
                                                    f"n_kv={n_kv}) + output head on ggml-cpu (oracle/_ref/{isa}, OpenMP, {cores} threads), mean of {done} timed steps after {max(warmup, 1)} warm-up, x{cfg['n_layer']} layers")
    # port: the scalar oracle on one core, one q4_K ffn_down GEMV, scaled by bytes
    orc = O.Oracle()
    K, M = shapes["down"]
    M = 512
    w = O.random_blocks(rng, O.Q4_K, K, M)
    x = rng.standard_normal((1, K)).astype(np.float32)
    t0 = time.time(); orc.mul_mat(O.Q4_K, w, K, M, x, 1); dt = time.time() - t0
    total_bytes = sum(wbytes(lt[n], *shapes[n]) for lt in types for n in shapes) + wbytes(out_type, cfg["n_embd"], cfg["n_vocab"])
    return 1.0 / (dt * total_bytes / w.nbytes), "port", 1, f"scalar oracle, q4_K {M}x{K} GEMV on 1 core scaled by weight bytes"


def metric_name(cfg, quant):
    return f"tg128 tok/s, {pretty(cfg)} {quant.upper()} (llama-bench tg128: one decode token per step)"


def bench_config(cfg, args):
    """what both arms run -- kept identical so the driver can compare the lines"""
    n_kv = ((args.depth + 1 + 255) // 256) * 256
    return {"workload": f"{cfg['name']}-{args.quant} tg128: one decode token per step (llama-bench -n 128 -fa 1), random-init GGUF weights",
            "tensor_types": "Q4_K_M mix: q4_K, q6_K on attn_v/ffn_down of half the layers, q6_K output" if args.quant == "q4_k_m" else args.quant + " weights",
            "depth": args.depth, "n_kv": n_kv}


def run_reference(args):
    """the reference's own CPU implementation on this box's host cores: its llama-bench binary with -ngl 0 (oracle/_ref, compiled in place; the
    AVX-512 build when the host runs it), else a sampled layer of ggml-cpu graphs through oracle/ref_driver.c"""
    cfg = MODELS[args.model]
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    args.lb_reps = 2
    lb = llama_bench_through_plugin(args, budget_s=400, ngl=0) if not cfg.get("n_expert") or True else {"unavailable": "-"}
    extra = {}
    if "tg128" in lb:
        tps, kind, cores = lb["tg128"]["value"], "reference", lb["threads"]
        sample = f"{lb['command']}: 128 generated tokens x 2 repetitions after llama-bench's own warm-up run, {lb['build']}, {cores} threads"
        extra = {"llama_bench": lb}
    else:
        tps, kind, cores, sample = cpu_baseline(cfg, args.depth, budget_s=150.0, steps=args.steps, warmup=args.warmup, quant=args.quant)
        extra = {"llama_bench": lb}
    line = {"impl": "reference", "metric": metric_name(cfg, args.quant), "value": tps, "unit": "tok/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / tps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8 dot / f32 accumulate (q8_0, q8_K activations)", "data": "synthetic",
            "config": bench_config(cfg, args),
            "cpu_baseline": {"value": tps, "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": tps, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.time() - t0}
    line.update(extra)
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import ggml_b200 as B
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = torch = None
    if world > 1:
        # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = B.Context(local)
    cfg = MODELS[args.model]
    hbm_gbs, tf_burst, tf_sust, peak_src = measured_peaks()

    keep = []
    peer = None
    replicas = world if (cfg.get("n_expert") and world > 1) else 1       # MUL_MAT_ID does not shard: N independent replicas, no data-path collective
    use_nccl = bool(os.environ.get("B200_BENCH_NCCL"))     # baseline: torch.distributed all-gather after every sharded matmul
    if world > 1 and not use_nccl and replicas == 1:
        from row_split import PeerExchange                   # the product: all-gather fused into the GEMV epilogue over NVLink peer memory
        peer = PeerExchange(ctx, dist, world, rank, 64 << 20)

    def alloc(t, ne):
        if world == 1 or replicas > 1:
            return ctx.empty(t, ne)
        if peer is not None:
            return peer.alloc(t, ne)
        buf = torch.empty(int(np.prod(ne)) * (4 if t == F32 else 2), dtype=torch.uint8, device=f"cuda:{local}")
        keep.append(buf)
        dt = B.DeviceTensor(ctx, t, ne, ptr=buf.data_ptr())
        dt.torch = buf
        return dt

    tg = HotPathModel(B, ctx, cfg, 1, args.depth, rank, world, alloc=alloc, dist=dist, quant=args.quant, peer=peer)
    if world > 1 and peer is None and replicas == 1:
        ext = torch.cuda.ExternalStream(ctx.stream, device=f"cuda:{local}")

        def gather(full, M, N):
            # the one exchange step of split-mode=row: all-gather of the row slices, issued on the context's own stream
            with torch.cuda.stream(ext):
                tg.rs.gather(full.torch.view(torch.float32), M, N)
        tg.gather = gather

    def timed(model, steps, warmup, use_graph):
        if dist:
            dist.barrier()                                  # the ranks meet inside the kernels (bounded waits): nobody starts seconds ahead of the others
        for _ in range(max(warmup, 1)):
            model.step()
        ctx.synchronize()
        graph = None
        if use_graph:
            ctx.graph_begin(); model.step(); graph = ctx.graph_end()
            for _ in range(2):
                ctx.graph_launch(graph)
            ctx.synchronize()
        if dist:
            dist.barrier()
        e0, e1 = ctx.event(), ctx.event()
        n0 = ctx.launches
        ctx.synchronize()
        ctx.record(e0)
        for _ in range(steps):
            if graph:
                ctx.graph_launch(graph)
            else:
                model.step()
        ctx.record(e1)
        ctx.synchronize()
        ms = ctx.elapsed_ms(e0, e1)
        if dist:
            t = torch.tensor([ms], device=f"cuda:{local}"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        return ms / steps, graph, ctx.launches - n0

    use_graph = (world == 1 or peer is not None or replicas > 1) and not args.no_graph     # the peer exchange is all device-side: it captures
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches_before = ctx.launches
    ms_tg, graph, eager_launches = timed(tg, args.steps, args.warmup, use_graph)
    clocks = sampler.stop() if rank == 0 else None
    launches_per_step = (ctx.launches - launches_before) // (max(args.warmup, 1) + (1 if use_graph else args.steps)) if use_graph else eager_launches // args.steps

    # ---- e2e: same step through the public API with host buffers: H2D of the embedding row, D2H of the logits, sync per token
    L = B.lib()
    e = cfg["n_embd"]
    hx = L.b200_host_alloc(e * 4); hl = L.b200_host_alloc(cfg["n_vocab"] * 4)
    np.ctypeslib.as_array((__import__("ctypes").c_float * e).from_address(hx))[:] = np.random.default_rng(1).standard_normal(e).astype(np.float32)
    e2e_steps = args.steps
    ctx.synchronize()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        L.b200_memcpy_h2d(ctx.h, tg.x.ptr, hx, e * 4)
        if graph:
            ctx.graph_launch(graph)
        else:
            tg.step()
        L.b200_memcpy_d2h(ctx.h, hl, tg.logits.ptr, cfg["n_vocab"] * 4)
        ctx.synchronize()                                   # llama-bench synchronizes after every generated token
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    if dist:
        t = torch.tensor([e2e_ms], device=f"cuda:{local}"); dist.all_reduce(t, op=dist.ReduceOp.MAX); e2e_ms = float(t.item())

    # ---- roofline leg: the dominant kernel (the streaming dequant-GEMV) timed live with CUDA events on the launching stream:
    # the same step with the attention nodes left out, replayed as a graph (bracketing every launch with its own event pair
    # costs ~5 us per launch, which would swamp 2-10 us kernels; the per-kind event profile is reported separately)
    def mm_only():
        tg.step(attention=False)
    gemv_ms_step = None
    if dist:
        ctx.synchronize(); dist.barrier()
    if use_graph:                                           # (every rank replays its own graph; with the peer exchange they meet inside the kernels)
        mm_only(); ctx.synchronize()
        l0 = ctx.launches_by_kind()
        ctx.graph_begin(); mm_only(); g2 = ctx.graph_end()
        l1 = ctx.launches_by_kind()
        gemv_n = l1["gemv"] - l0["gemv"] + l1["gemv_id"] - l0["gemv_id"]
        other_n = sum(l1.values()) - sum(l0.values()) - gemv_n
        ctx.graph_launch(g2); ctx.synchronize()
        e0, e1 = ctx.event(), ctx.event()
        ctx.record(e0)
        for _ in range(10):
            ctx.graph_launch(g2)
        ctx.record(e1)
        gemv_ms_step = ctx.elapsed_ms(e0, e1) / 10
        if dist:
            t = torch.tensor([gemv_ms_step], device=f"cuda:{local}"); dist.all_reduce(t, op=dist.ReduceOp.MAX); gemv_ms_step = float(t.item())
    roof_how = "CUDA events around 10 replays of a graph holding only the MUL_MAT kernels of one token (max over ranks)" if use_graph else \
               "sum of the per-launch CUDA-event brackets of the MUL_MAT kernels, 3 eager steps (no graph in this mode: launch gaps excluded)"
    if dist:
        ctx.synchronize(); dist.barrier()
    ctx.profile(True)
    for _ in range(3):
        tg.step()
    prof = ctx.profile_get()
    ctx.profile(False)
    if gemv_ms_step is None:
        gemv_ms_step = (prof.get("gemv", (0.0, 0))[0] + prof.get("gemv_id", (0.0, 0))[0]) / 3
        gemv_n, other_n = (prof.get("gemv", (0.0, 0))[1] + prof.get("gemv_id", (0.0, 0))[1]) // 3, 0
    shapes = weight_shapes(cfg)
    act_bytes = sum(K * 4 + M * 4 for lt in tg.layers for (K, M) in [shapes[n] for n in shapes]) // max(tg.world, 1)   # (MoE: x2 on the expert tensors, < 0.1 % of the bytes)
    gemv_bytes_step = tg.weight_bytes + act_bytes
    achieved = gemv_bytes_step / (gemv_ms_step * 1e-3) / 1e9 if gemv_ms_step > 0 else 0.0

    # ---- row split only: the gathered results of one step must be the unsplit ones.  Every rank runs the step once more; all ranks must hold
    # identical columns; rank 0 builds the same model unsplit (same stream of random blocks), runs the same step and compares every matmul
    # result of the last layer and the logits word for word.
    split_check = None
    if world > 1 and replicas == 1:
        import hashlib
        names = ("q", "k", "v", "o", "g", "u", "d", "logits")
        chained = tg.chains is not None
        # (with the chain, gate / up leave the kernel as their SwiGLU product: every rank holds only its own rows of g and u)
        # and wo's / ffn_down's results are handed on as quantized blocks by their row owners: the same holds for o and d)
        Ll = tg.layers[-1]
        part = {"g": Ll["gate"][1:], "u": Ll["up"][1:], "o": Ll["wo"][1:], "d": Ll["down"][1:]} if chained else {}
        ctx.synchronize(); dist.barrier()
        tg.reset_state()
        tg.step(); ctx.synchronize()
        mine = {n: ctx.download(getattr(tg, n)).reshape(-1).copy() for n in names}
        digest = hashlib.sha1(b"".join(mine[n].tobytes() for n in names if n not in part)).hexdigest()
        digests = [None] * world
        dist.all_gather_object(digests, digest)
        ok = [len(set(digests)) == 1]
        if rank == 0:
            # the unsplit model: separate launches / the chain in the same hand-over form on one GPU (then every word must be equal: a row's
            # dot product is summed in the same order whichever rank or CTA owns it)
            one = HotPathModel(B, ctx, cfg, 1, args.depth, 0, 1, quant=args.quant, chain="split" if chained else False)
            for n in ("x",) if chained else ("x", "h", "ff", "attn"):   # the same inputs, as they are NOW (the e2e leg above uploaded its own embedding row into x)
                ctx.h2d(getattr(one, n).ptr, ctx.download(getattr(tg, n)).reshape(-1))
            # (a model whose unsplit form the chain kernel declines -- 70B's output head is 868 stages per CTA on one GPU, the table holds 640 --
            # has no unsplit counterpart in the same arithmetic: then only the ranks are compared with each other; the same widths are covered
            # word for word by tests/test_gpu_chain_split.py)
            no_ref = chained and one.chains is None
            one.reset_state()
            one.step(); ctx.synchronize()
            whole = {n: ctx.download(getattr(one, n)).reshape(-1).copy() for n in names}
            del one
            if no_ref:
                whole = {n: mine[n].copy() for n in names}
            for n, (lo, hi) in part.items():
                mine[n], whole[n] = mine[n][lo:hi], whole[n][lo:hi]
            diff = {n: int((whole[n].view(np.uint32) != mine[n].view(np.uint32)).sum()) for n in names}
            err = {n: float(((whole[n].astype(np.float64) - mine[n]) ** 2).sum() / max((whole[n].astype(np.float64) ** 2).sum(), 1e-300)) for n in names}
            # (the unsplit model may pick another GEMV kernel for a shape than the sharded one does: a different summation order, a few last-bit
            # differences; anything beyond that is a wrong or missing row)
            ok[0] = ok[0] and max(err.values()) < 1e-9 and all(np.isfinite(mine[n]).all() for n in names)
            split_check = {"ranks_identical": len(set(digests)) == 1, "nmse_vs_unsplit_max": None if no_ref else max(err.values()), "words_differing_from_unsplit": None if no_ref else diff,
                           "unsplit_reference": "none: the chain kernel declines this model unsplit on one GPU (ranks compared with each other only)" if no_ref else
                                                ("the chain in the same hand-over form on one GPU" if chained else "separate launches on one GPU"),
                           "exchange_error": int(ctx.exchange_error(peer.ex)) if peer is not None else 0}
        dist.broadcast_object_list(ok, src=0)
        if not ok[0]:
            if rank == 0:
                print(json.dumps({"error": "row-split results differ from the unsplit model", "split_check": split_check}), file=sys.stderr, flush=True)
            dist.barrier(); dist.destroy_process_group()
            sys.exit(1)

    result = None
    if rank == 0:
        tps = replicas * 1e3 / ms_tg
        total_bytes = tg.weight_bytes * (world if replicas == 1 else 1) + tg.kv_bytes_per_layer * cfg["n_layer"]
        chain = tg.chains is not None
        kernel_name = ("k_decode_chain: one persistent kernel per layer for wo, +residual, ffn_norm, ffn_gate / ffn_up, SwiGLU, ffn_down, +ffn_inp, attn_norm, wq / wk / wv "
                       "(TMA tensor requests into one shared-memory ring, dp4a dots, q8_1 quantization and the glue in the phase prologues)") if chain else \
                      "decode dequant-GEMV with fused q8_1 quantization: k_gemv_q4k_mma (raw q4_K, TMA boxes + int8 mma.sync) and k_gemv_stream (q6_K / long-row q4_K, TMA-streamed dp4a)"
        traffic, traffic_src = traffic_per_launch(gemv_bytes_step / max(gemv_n, 1))
        result = {
            # BASELINE.json's metric is llama-bench tg128 / pp512.  `value` is the device-resident form of tg128 (the hot path of one token as a CUDA
            # graph, inputs in HBM); `e2e` is the reference's llama-bench itself through the plugin (host token ids in, host logits out) when its
            # binary is present, else the flat C-ABI with host buffers; llama_bench_tg128 / llama_bench_pp512 carry the tool's own figures
            "metric": metric_name(cfg, args.quant), "value": tps, "unit": "tok/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_tg, "higher_is_better": True, "scaling": "weak" if replicas > 1 else "strong", "vs_baseline": None,
            "dtype": "int8 dot (dp4a / mma.sync) with f32 accumulate over q8_1 activations; f16 KV, f32 softmax", "data": "synthetic",
            "config": dict(bench_config(cfg, args), **({} if world == 1 else {"parallelism": (f"{replicas} independent replicas (MUL_MAT_ID does not shard, ggml-cuda.cu:2274), no collective" if replicas > 1 else
                                       f"row-split tp{world}, all-gather " + ("inside the persistent chain kernel's hand-over between its phases (row owners publish q8_1 records as (word, tag) pairs into every rank's area over NVLink; b200_decode_chain_split), layer 0's wq / wk / wv fused into the GEMV epilogue" if peer is not None and chain else
                                                                              "fused into the GEMV epilogue (NVLink peer stores)" if peer is not None else "by NCCL"))})),
            "value_is": f"every MUL_MAT{'_ID' if cfg.get('n_expert') else ''} and FLASH_ATTN_EXT node of one decode token ({cfg['n_layer']} layers + output head) through the flat C-ABI, one CUDA graph, inputs resident in HBM" +
                        ("; the ADD / RMS_NORM / MUL / SwiGLU nodes between them run inside the chain kernel; ROPE, KV-cache writes and the embedding lookup are only in e2e" if chain else "; glue ops only in e2e"),
            "cuda_graph": bool(graph), "l2": f"weights {tg.weight_bytes * (world if replicas == 1 else 1) / 1e9:.1f} GB >> 126 MB L2, streamed once per step: no flush needed",
            "hbm_bytes_per_step": total_bytes, "hbm_frac_whole_step": total_bytes / (world if replicas == 1 else 1) / (ms_tg * 1e-3) / 1e9 / hbm_gbs,
            "roofline": {"kernel": kernel_name, "bound": "hbm", "achieved": achieved, "peak": hbm_gbs, "unit": "GB/s",
                         "frac": achieved / hbm_gbs, "traffic": traffic, "traffic_source": traffic_src, "algorithmic_bytes_per_launch": gemv_bytes_step / max(gemv_n, 1),
                         "peak_source": f"{peak_src} (MEASURED_PEAKS.json hbm_gbs: a COPY, half reads half writes; read-only streams reach 7.1-7.3 TB/s on this part, profiles/r2_membw_ceilings.txt)",
                         "launches_per_step": gemv_n, "other_launches_in_timed_graph": other_n, "ms_per_step": gemv_ms_step,
                         "algorithmic_bytes_per_step": gemv_bytes_step, "how": roof_how},
            "kernel_ms_per_step_event_bracketed": {k: v[0] / 3 for k, v in prof.items()},
            "e2e": {"value": replicas * 1e3 / e2e_ms, "unit": "tok/s", "h2d_bytes_per_step": e * 4, "d2h_bytes_per_step": cfg["n_vocab"] * 4, "ms_per_step": e2e_ms,
                    "how": "flat C-ABI (b200_*), pinned H2D of the embedding row + graph replay + D2H of the logits + sync per token"},
            "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step, "clocks": clocks,
        }
        if split_check is not None:
            result["split_check"] = split_check
    # ---- secondary: pp512 (prefill) on one GPU
    tg_weight_bytes, tg_kv_bytes = tg.weight_bytes, tg.kv_bytes_per_layer * cfg["n_layer"]
    if world == 1 and not args.no_pp:
        del tg
        pp = HotPathModel(B, ctx, cfg, 512, args.depth, quant=args.quant)
        ms_pp, _, _ = timed(pp, max(2, min(args.steps, 5)), 1, False)
        ctx.profile(True); pp.step(); prof_pp = ctx.profile_get(); ctx.profile(False)
        flops = pp.matmul_flops()
        mm_ms = prof_pp.get("gemm", (0.0, 0))[0] + prof_pp.get("gemv", (0.0, 0))[0] + prof_pp.get("gemv_id", (0.0, 0))[0]
        result["pp512"] = {"value": 512e3 / ms_pp, "unit": "tok/s", "ms_per_step": ms_pp, "matmul_tflop_per_step": flops / 1e12,
                           "roofline": {"kernel": "MUL_MAT prefill kernels", "bound": "tensor", "achieved": flops / (mm_ms * 1e-3) / 1e12 if mm_ms else 0.0,
                                        "peak": tf_sust, "unit": "TFLOP/s", "frac": (flops / (mm_ms * 1e-3) / 1e12 / tf_sust) if mm_ms else 0.0,
                                        "peak_source": f"{peak_src} bf16_tflops_sustained"},
                           "kernel_ms_per_step": {k: v[0] for k, v in prof_pp.items()}}
    if rank == 0:
        if not args.no_cpu and world == 1:
            try:
                v, kind, cores, sample = cpu_baseline(cfg, args.depth, quant=args.quant)
                result["cpu_baseline"] = {"value": v, "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample}
            except Exception as ex:                      # the baseline is a reported number, never a reason to lose the GPU line
                result["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": os.cpu_count(), "kind": "unavailable", "sample": repr(ex)}
        if not args.no_llama_bench:
            # the tool BASELINE.json's metric is named after, through the plugin, pinned to this rank's GPU
            lb = llama_bench_through_plugin(args, device=local)
            result["llama_bench"] = lb
            if "tg128" in lb:
                result["llama_bench_tg128"] = {"value": lb["tg128"]["value"], "stddev": lb["tg128"]["stddev"], "unit": "tok/s", "n_gpus": 1,
                                               "roofline_frac": lb["tg128"]["value"] * (tg_weight_bytes + tg_kv_bytes) / 1e9 / hbm_gbs,
                                               "note": "whole llama graph (hot ops + ROPE, KV writes, embedding, sampling sync) by the reference's llama-bench through GGML_BACKEND_PATH; frac = (weight + KV-cache bytes one token reads) x tok/s / measured HBM copy peak"}
                if world == 1:
                    result["e2e_flat_api"] = result["e2e"]
                    result["e2e"] = {"value": lb["tg128"]["value"], "unit": "tok/s", "h2d_bytes_per_step": 4, "d2h_bytes_per_step": cfg["n_vocab"] * 4, "ms_per_step": 1e3 / lb["tg128"]["value"],
                                     "how": "the reference's own llama-bench (-n 128 -fa 1 -ngl 99) through the plugin: host token id in, graph_compute, logits to host, sync per token"}
            if "pp512" in lb:
                fl = 2.0 * 512 * sum(K * M for (K, M) in weight_shapes(cfg).values()) * cfg["n_layer"]
                result["llama_bench_pp512"] = {"value": lb["pp512"]["value"], "stddev": lb["pp512"]["stddev"], "unit": "tok/s", "n_gpus": 1,
                                               "roofline_frac": fl / (512.0 / lb["pp512"]["value"]) / 1e12 / tf_sust, "note": "matmul flops of a 512-token ubatch / its whole time / sustained bf16 peak"}
        print(json.dumps(result), flush=True)
    if dist:
        dist.barrier(); dist.destroy_process_group()


def ref_build_dir():
    """oracle/_ref/<isa>: the AVX-512 build of the reference when this host runs it (the boxes have AVX-512 VNNI / BF16), else the AVX2 one"""
    try:
        flags = open("/proc/cpuinfo").read()
    except OSError:
        flags = ""
    for isa, need in (("avx512", ("avx512f", "avx512_vnni", "avx512_bf16", "avx512vbmi", "avx_vnni")), ("avx2", ())):
        d = os.path.join(ROOT, "oracle", "_ref", isa)
        if all(n in flags for n in need) and os.path.exists(os.path.join(d, "llama-bench")):
            return d, isa
    return os.path.join(ROOT, "oracle", "_ref", "avx2"), "avx2"


def llama_bench_through_plugin(args, budget_s=420, device=0, ngl=99):
    """Supporting evidence beside the hot-path numbers: the reference's OWN llama-bench binary (oracle/_ref/avx2/llama-bench, compiled
    in place from the unmodified reference by `make -C oracle llama`; nothing of /root/reference is read at run time) on a random-block
    GGUF of the same model (scripts/make_gguf.py), with libggml-b200.so loaded through GGML_BACKEND_PATH -- every node of the llama graph
    (hot ops + glue ops) runs on the device.  Absent binary or any failure -> {"unavailable": why}; never fatal."""
    import subprocess, tempfile, time
    refdir, isa = ref_build_dir() if ngl == 0 else (os.path.join(ROOT, "oracle", "_ref", "avx2"), "avx2")
    lb = os.path.join(refdir, "llama-bench")
    plugin = os.path.join(ROOT, "super-llama.cpp_b200", "libggml-b200.so")
    if not os.path.exists(lb):
        return {"unavailable": "oracle/_ref/avx2/llama-bench not built (make -C oracle llama, needs the reference sources at build time)"}
    if args.model not in ("8b", "tinyllama", "70b", "mixtral"):
        return {"unavailable": f"scripts/make_gguf.py does not write {args.model}"}
    t0 = time.time()
    gguf = os.path.join(tempfile.gettempdir(), f"b200-bench-{args.model}-{args.quant}-{os.getpid()}.gguf")
    try:
        subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "make_gguf.py"), "--model", args.model, "--quant", args.quant, "-o", gguf],
                       check=True, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=budget_s / 2)
        cmd = [lb, "-m", gguf, "-p", "512", "-n", "128", "-r", str(getattr(args, "lb_reps", 3)), "-fa", "1", "-ngl", str(ngl), "-t", str(os.cpu_count() or 1), "-o", "json"] + (["-d", str(args.depth)] if args.depth else [])
        env = dict(os.environ)
        if ngl:
            # one device only: llama-bench would otherwise spread the layers over every visible GPU (-sm layer)
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            env.update(GGML_BACKEND_PATH=plugin, CUDA_VISIBLE_DEVICES=(vis.split(",")[device] if vis else str(device)))
        else:
            env.pop("GGML_BACKEND_PATH", None)
        p = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=max(30, budget_s - (time.time() - t0)))
        rows = json.loads(p.stdout[p.stdout.index("["):])
        if ngl and "B200" not in str(rows[0].get("backends")):
            return {"unavailable": f"llama-bench did not pick the plugin up (backends: {rows[0].get('backends')})"}
        out = {"command": " ".join(cmd[:1] + ["-m", "<random-block GGUF>"] + cmd[3:]), "backend": rows[0].get("backends"), "model": rows[0].get("model_type"), "data": "random GGUF blocks, no_vocab tokenizer, random token ids",
               "build": f"oracle/_ref/{isa} (the reference compiled in place, {'AVX-512 VNNI/BF16' if isa == 'avx512' else 'x86-64-v3'})", "threads": os.cpu_count() or 1}
        for r in rows:
            name = f"pp{r['n_prompt']}" if r.get("n_prompt") else f"tg{r['n_gen']}"
            out[name] = {"value": r["avg_ts"], "stddev": r["stddev_ts"], "unit": "tok/s"}
        return out
    except Exception as ex:
        return {"unavailable": repr(ex)[:300]}
    finally:
        try: os.remove(gguf)
        except OSError: pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--depth", type=int, default=0, help="tokens already in the KV cache (llama-bench -d)")
    ap.add_argument("--model", default="8b", choices=sorted(MODELS), help="8b = the headline config; 70b = configs[3] scale point")
    ap.add_argument("--quant", default="q4_k_m", choices=["q4_k_m", "q8_0", "q4_0"], help="GGUF file type of the synthetic weights")
    ap.add_argument("--no-pp", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-llama-bench", action="store_true", help="skip the supporting run of the reference's llama-bench through the plugin")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
