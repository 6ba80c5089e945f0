#!/usr/bin/env python
"""Random-init GGUF of a llama-architecture model with the tensor types llama-quantize would pick, written directly as GGUF blocks
(random nibbles / bytes, small fp16 scales) with the `gguf` package -- no f16 stage and no quantize tool, so a 4.6 GB Q4_K_M file takes
well under a minute on the GPU box.  Tokenizer: "no_vocab" (src/llama-vocab.cpp:1706-1725 then adds llama.vocab_size dummy tokens;
llama-bench feeds random token ids).  Used by scripts/gpu_llama_bench.sh to run the reference's own llama-bench through the plugin.

  python scripts/make_gguf.py --model tinyllama --quant q4_0 -o /tmp/tiny-q4_0.gguf
"""
import argparse, os, sys
import numpy as np
import gguf
from gguf import GGMLQuantizationType as QT

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                       # model tables + the Q4_K_M tensor-type rule (src/llama-quant.cpp restated there)

QTYPE = {bench.Q4_0: QT.Q4_0, bench.Q8_0: QT.Q8_0, bench.Q4_K: QT.Q4_K, bench.Q6_K: QT.Q6_K}


def blocks(rng, t, K, M):
    """[M rows][K / blck * tsize bytes] of random blocks of ggml type t (same generator as bench.random_blocks, but one pool reused)"""
    nb = K // bench.BLCK[t] * M
    pool = bench.random_blocks(rng, t, min(nb, 1 << 18)).reshape(-1, bench.TSIZE[t])
    # block scales such that the weights are zero-mean with std ~ 1 / sqrt(K): activations then stay O(1..100) through the layers, as in a
    # trained model.  (With bench.random_blocks' scales -- fine for timing -- a q6_K matmul multiplies magnitudes by ~50 and two layers
    # push activations past 65504 * 127, where the fp16 block scale of q8_1 overflows: tests/test_gpu_llama_logits.py found that.)
    n = pool.shape[0]; target = 1.0 / np.sqrt(K)
    f16 = lambda v: v.astype(np.float16).view(np.uint8).reshape(n, 2)
    if t == bench.Q4_K:                                    # w = d * sc * q - dmin * m, sc, m in [0, 63], q in [0, 15]: dmin = 7.5 d centres it, std ~ 215 d
        d = rng.uniform(0.5, 1.0, n) * (target / 160.0)
        pool[:, 0:2] = f16(d); pool[:, 2:4] = f16(7.5 * d)
    elif t == bench.Q6_K:                                  # w = d * sc * (q - 32), sc int8, q in [0, 63]: std ~ 1370 |d|
        pool[:, 208:210] = f16(rng.uniform(-1, 1, n) * (target * np.sqrt(3) / 1370.0))
    elif t == bench.Q4_0:                                  # w = d * (q - 8): std ~ 4.6 |d|
        pool[:, 0:2] = f16(rng.uniform(-1, 1, n) * (target * np.sqrt(3) / 4.6))
    elif t == bench.Q8_0:                                  # w = d * q, q int8: std ~ 74 |d|
        pool[:, 0:2] = f16(rng.uniform(-1, 1, n) * (target * np.sqrt(3) / 74.0))
    reps = (nb + pool.shape[0] - 1) // pool.shape[0]
    raw = np.tile(pool, (reps, 1))[:nb]
    return raw.reshape(M, K // bench.BLCK[t] * bench.TSIZE[t])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="tinyllama", choices=["tinyllama", "8b", "70b", "mixtral"])
    ap.add_argument("--quant", default="q4_0", choices=["q4_0", "q8_0", "q4_k_m"])
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (smoke runs)")
    ap.add_argument("-o", "--out", required=True)
    a = ap.parse_args()
    cfg = dict(bench.MODELS[a.model])
    if a.layers:
        cfg["n_layer"] = a.layers
    e, f, nh, nkv, hd, nl, nv = cfg["n_embd"], cfg["n_ff"], cfg["n_head"], cfg["n_head_kv"], cfg["head_dim"], cfg["n_layer"], cfg["n_vocab"]
    types, out_type = bench.q4_k_m_layer_types(cfg, a.quant)
    emb_type = {"q4_0": bench.Q4_0, "q8_0": bench.Q8_0, "q4_k_m": bench.Q4_K}[a.quant]           # token_embd: the file's base type
    rng = np.random.default_rng(1234)
    w = gguf.GGUFWriter(a.out, "llama")
    w.add_name(f"{cfg['name']}-random-{a.quant}")
    w.add_context_length(2048 if a.model == "tinyllama" else 8192)
    w.add_embedding_length(e); w.add_block_count(nl); w.add_feed_forward_length(f)
    w.add_head_count(nh); w.add_head_count_kv(nkv); w.add_rope_dimension_count(hd)
    w.add_layer_norm_rms_eps(1e-5); w.add_rope_freq_base(10000.0 if a.model == "tinyllama" else 1e6 if a.model == "mixtral" else 500000.0)
    ne = cfg.get("n_expert", 0)
    if ne:                                                 # LLM_ARCH_LLAMA with experts = Mixtral (src/models/llama.cpp:100-137: build_moe_ffn)
        w.add_expert_count(ne); w.add_expert_used_count(cfg["n_expert_used"])
    w.add_vocab_size(nv)
    w.add_file_type({"q4_0": 2, "q8_0": 7, "q4_k_m": 15}[a.quant])                                 # LLAMA_FTYPE_MOSTLY_*
    w.add_tokenizer_model("no_vocab")
    kvd = nkv * hd

    def q(name, t, K, M):
        w.add_tensor(name, blocks(rng, t, K, M), raw_dtype=QTYPE[t])

    def q3(name, t, K, M, n):                              # stacked expert matrices [K, M, n_expert]
        w.add_tensor(name, np.stack([blocks(rng, t, K, M) for _ in range(n)]), raw_dtype=QTYPE[t])

    def norm(name, n):
        w.add_tensor(name, np.ones(n, dtype=np.float32))

    q("token_embd.weight", emb_type, e, nv)
    for i, lt in enumerate(types):
        p = f"blk.{i}."
        norm(p + "attn_norm.weight", e)
        q(p + "attn_q.weight", lt["wq"], e, e); q(p + "attn_k.weight", lt["wk"], e, kvd); q(p + "attn_v.weight", lt["wv"], e, kvd)
        q(p + "attn_output.weight", lt["wo"], e, e)
        norm(p + "ffn_norm.weight", e)
        if ne:
            w.add_tensor(p + "ffn_gate_inp.weight", (rng.standard_normal((ne, e)) / np.sqrt(e)).astype(np.float32))      # the router, f32 as llama-quantize leaves it
            q3(p + "ffn_gate_exps.weight", lt["gate"], e, f, ne); q3(p + "ffn_up_exps.weight", lt["up"], e, f, ne); q3(p + "ffn_down_exps.weight", lt["down"], f, e, ne)
        else:
            q(p + "ffn_gate.weight", lt["gate"], e, f); q(p + "ffn_up.weight", lt["up"], e, f); q(p + "ffn_down.weight", lt["down"], f, e)
    norm("output_norm.weight", e)
    q("output.weight", out_type, e, nv)
    w.write_header_to_file(); w.write_kv_data_to_file(); w.write_tensors_to_file(); w.close()
    print(f"wrote {a.out}: {os.path.getsize(a.out) / 1e9:.2f} GB, {nl} layers, {a.quant}")


if __name__ == "__main__":
    main()
