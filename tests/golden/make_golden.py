"""Generates tests/golden/hotpath_golden.npz from the UNMODIFIED reference (oracle/_ref/avx2, compiled in place
from /root/reference by oracle/Makefile).  Run in the build container:  python tests/golden/make_golden.py
The vectors pin the oracle (tests/test_oracle.py) and the CUDA path (tests/test_gpu_ops.py) on machines where the
reference library is absent.  ISA of the generating library: x86-64-v3 (AVX2 + F16C + FMA), gcc 13.3 -O3."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import oracle as O  # noqa: E402


def main():
    ref = O.Reference()
    rng = np.random.default_rng(20261104)
    out = {}

    # ---- q8_1 activation quantization (A1): random rows + the edge cases SURVEY 7.3 lists
    x = rng.uniform(-1, 1, (16, 256)).astype(np.float32)
    x[0, :64] = 0.0
    x[1, :32] = np.arange(32, dtype=np.float32) * 0.5          # .5 ties once scaled
    x[1, 31] = 127.0
    x[2, :64] = (rng.integers(-254, 255, 64) * 0.5).astype(np.float32); x[2, 0] = 127.0
    x[3, :32] = 1e-40                                            # denormal amax -> id = inf
    x[4, :32] = -1e-39
    x[5, :32] = 3.0e38
    x[6, :32] = np.float32(-3.4028235e38)
    x[7] *= 1e-6
    x[8] *= 1e4
    out["q8_1_x"] = x
    out["q8_1_x86"] = ref.quantize_q8_1(x, O.Q8_1_X86)
    out["q8_1_ref"] = ref.quantize_q8_1(x, O.Q8_1_REF)
    # rows with NaN / inf inputs: the AVX2 amax depends on WHERE a NaN sits (maxps returns its second operand), the scalar code on
    # whether it comes last; own generator so that the vectors above stay as they were
    rn = np.random.default_rng(20261105)
    xn = rn.standard_normal((72, 32)).astype(np.float32)
    xn[0, :] = np.nan
    for i in range(1, 33):
        xn[i, i - 1] = np.nan
    xn[33, :8] = np.nan; xn[34, 24:] = np.nan; xn[35, 8:16] = np.nan
    for i in range(36, 72):
        xn[i, rn.random(32) < rn.uniform(0.03, 0.6)] = np.nan
        if i % 5 == 0:
            xn[i, rn.integers(0, 32)] = np.inf
    out["q8_1_nan_x"] = xn
    out["q8_1_nan_x86"] = ref.quantize_q8_1(xn, O.Q8_1_X86)
    out["q8_1_nan_ref"] = ref.quantize_q8_1(xn, O.Q8_1_REF)
    out["q8_0_cpu"] = ref.from_float(O.Q8_0, x)
    out["q8_K_cpu"] = ref.from_float(O.Q8_K, x)

    # ---- weight decode + MUL_MAT per type
    K, M, N = 512, 24, 5
    xs = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    out["mm_x"] = xs
    for t in (O.Q4_0, O.Q8_0, O.Q4_K, O.Q5_K, O.Q6_K):
        name = O.TYPE_NAMES[t]
        wf = rng.uniform(-1, 1, (M, K)).astype(np.float32)
        w = ref.quantize_weights(t, wf, K)
        out[f"w_{name}"] = w
        out[f"deq_{name}"] = ref.dequantize(t, w, M * K)
        out[f"mm_{name}"] = ref.mul_mat(t, w, K, M, xs, N, threads=1)
    # batched / broadcast MUL_MAT (ne02=2, ne12=4 -> r2=2; ne03=1, ne13=2)
    wf = rng.uniform(-1, 1, (1, 2, M, 256)).astype(np.float32)
    w = ref.quantize_weights(O.Q4_K, wf, 256)
    xb = rng.uniform(-1, 1, (2, 4, 3, 256)).astype(np.float32)
    out["mmb_w"], out["mmb_x"] = w, xb
    out["mmb_out"] = ref.mul_mat(O.Q4_K, w, 256, M, xb, 3, ne02=2, ne03=1, ne12=4, ne13=2, threads=1)

    # ---- MUL_MAT_ID (8 experts, 2 used, 6 tokens; b broadcast (ne11=1) and per-slot (ne11=2))
    E, U, T, K2, M2 = 8, 2, 6, 256, 32
    for t in (O.Q4_0, O.Q4_K, O.Q8_0):
        name = O.TYPE_NAMES[t]
        as_ = ref.quantize_weights(t, rng.uniform(-1, 1, (E, M2, K2)).astype(np.float32), K2)
        ids = np.stack([rng.permutation(E)[:U] for _ in range(T)]).astype(np.int32)
        out[f"id_as_{name}"], out[f"id_ids_{name}"] = as_, ids
        for ne11 in (1, U):
            b = rng.uniform(-1, 1, (T, ne11, K2)).astype(np.float32)
            out[f"id_b{ne11}_{name}"] = b
            out[f"id_out{ne11}_{name}"] = ref.mul_mat_id(t, as_, K2, M2, E, b, ne11, T, ids, U, threads=1)

    # ---- FLASH_ATTN_EXT: GQA 4:2, ragged kv=45, mask with -inf holes, sinks, ALiBi, softcap
    S, H, Hk, Nq, KV, DK, DV = 1, 4, 2, 3, 45, 64, 64
    q = rng.uniform(-1, 1, (S, H, Nq, DK)).astype(np.float32)
    k = rng.uniform(-1, 1, (S, Hk, KV, DK)).astype(np.float16)
    v = rng.uniform(-1, 1, (S, Hk, KV, DV)).astype(np.float16)
    mask = rng.uniform(-1, 1, (1, 1, 4, KV)).astype(np.float16)
    mask[0, 0, :, rng.integers(0, KV, 12)] = -np.inf
    mask[0, 0, 1, :20] = -np.inf
    sinks = rng.uniform(-10, 10, H).astype(np.float32)
    out.update(fa_q=q, fa_k=k, fa_v=v, fa_mask=mask, fa_sinks=sinks)
    sc = 1.0 / np.sqrt(DK)
    out["fa_out_plain"] = ref.flash_attn_ext(q, k, v, None, None, sc, threads=1)
    out["fa_out_mask"] = ref.flash_attn_ext(q, k, v, mask, None, sc, threads=1)
    out["fa_out_mask_sinks"] = ref.flash_attn_ext(q, k, v, mask, sinks, sc, threads=1)
    out["fa_out_alibi"] = ref.flash_attn_ext(q, k, v, mask, None, sc, max_bias=8.0, threads=1)
    out["fa_out_softcap"] = ref.flash_attn_ext(q, k, v, mask, None, sc, softcap=10.0, threads=1)
    out["fa_out_f32prec"] = ref.flash_attn_ext(q, k, v, mask, sinks, sc, prec=10, threads=1)

    path = os.path.join(HERE, "hotpath_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
