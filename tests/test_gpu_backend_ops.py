"""GPU test (-m gpu): the reference's OWN parity harness, tests/test-backend-ops.cpp, compiled unmodified from
/root/reference into oracle/_ref/avx2/test-backend-ops (oracle/Makefile), loading our backend as a plugin through
GGML_BACKEND_PATH (ggml/src/ggml-backend-reg.cpp:611-615) and comparing every supported case with ggml-cpu at
NMSE < 5e-4.  This is the drop-in proof: no line of the harness or of ggml knows about ggml-b200.
"""
import os
import re
import subprocess
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARNESS = os.path.join(ROOT, "oracle", "_ref", "avx2", "test-backend-ops")
PLUGIN = os.path.join(ROOT, "super-llama.cpp_b200", "libggml-b200.so")
OUT = os.path.join(ROOT, "gpurun_out")


def run_harness(op, pattern=None, mode="test", timeout=1500, env_extra=None, tag=""):
    if not os.path.exists(HARNESS):
        pytest.skip("oracle/_ref/avx2/test-backend-ops not built (needs /root/reference at build time)")
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN, **(env_extra or {}))
    cmd = [HARNESS, mode, "-b", "B200_0", "-o", op]
    if pattern:
        cmd += ["-p", pattern]
    p = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    txt = re.sub(r"\x1b\[[0-9;]*m", "", p.stdout)
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, f"test-backend-ops_{mode}_{op}{tag}.log"), "w") as f:
        f.write(txt)
    return p.returncode, txt


def summarize(txt, must_support):
    ok = len(re.findall(r": OK$", txt, re.M))
    fail = [l for l in txt.splitlines() if re.search(r"FAIL|ERR = |NaN|inf mismatch", l)]
    unsupported = [l for l in txt.splitlines() if "not supported" in l]
    m = re.search(r"(\d+)/(\d+) tests passed", txt)
    assert m, txt[-2000:]
    passed, total = int(m.group(1)), int(m.group(2))
    # a north-star case reported "not supported" is a coverage hole, not a pass
    holes = [l for l in unsupported if must_support(l)]
    return ok, passed, total, fail, holes


NORTH_STAR_W = ("type_a=q4_0", "type_a=q8_0", "type_a=q4_K")


def test_backend_ops_mul_mat():
    rc, txt = run_harness("MUL_MAT")
    must = lambda l: any(t in l for t in NORTH_STAR_W) and "type_b=f32" in l and "per=[0,1,2,3]" in l
    ok, passed, total, fail, holes = summarize(txt, must)
    assert rc == 0 and passed == total and not fail, "\n".join(fail[:20])
    assert not holes, "north-star MUL_MAT cases declined by supports_op:\n" + "\n".join(holes[:20])
    assert ok >= 100, f"only {ok} MUL_MAT cases actually ran on the device"


def test_backend_ops_mul_mat_id():
    rc, txt = run_harness("MUL_MAT_ID")
    must = lambda l: any(t in l for t in NORTH_STAR_W) and "type_b=f32" in l
    ok, passed, total, fail, holes = summarize(txt, must)
    assert rc == 0 and passed == total and not fail, "\n".join(fail[:20])
    assert not holes, "north-star MUL_MAT_ID cases declined:\n" + "\n".join(holes[:20])
    assert ok >= 100, f"only {ok} MUL_MAT_ID cases ran on the device"


def test_backend_ops_flash_attn_ext():
    rc, txt = run_harness("FLASH_ATTN_EXT", timeout=3000)
    must = lambda l: re.search(r"type_KV=(f16|q8_0|q4_0)", l) and re.search(r"hsk=(64|128),", l) and re.search(r"hsv=(64|128),", l)   # (quantized caches: b200_fattn_q.cu)
    ok, passed, total, fail, holes = summarize(txt, must)
    assert rc == 0 and passed == total and not fail, "\n".join(fail[:20])
    assert not holes, "north-star FLASH_ATTN_EXT cases declined:\n" + "\n".join(holes[:20])
    assert ok >= 300, f"only {ok} FLASH_ATTN_EXT cases ran on the device"


@pytest.mark.parametrize("op", ["MUL_MAT", "MUL_MAT_ID"])
def test_backend_ops_with_load_time_repack(op):
    """the same harness with GGML_B200_REPACK=all: every whole-tensor set_tensor of q6_K / q8_0 / q4_0 weights is permuted into
    the device layout by the plugin's buffer hook, so the repacked-layout kernels (streaming GEMV, column-blocked GEMV,
    MUL_MAT_ID, tcgen05 GEMM) are the ones compared with ggml-cpu"""
    rc, txt = run_harness(op, env_extra={"GGML_B200_REPACK": "all"}, tag="_repack")
    ok, passed, total, fail, holes = summarize(txt, lambda l: False)
    assert rc == 0 and passed == total and not fail, "\n".join(fail[:20])
    assert ok >= 100


GLUE_OPS = ["RMS_NORM", "RMS_NORM_MUL_ADD", "ADD_RMS_NORM", "SOFT_MAX", "SILU", "ADD", "MUL", "SCALE", "ROPE", "SET_ROWS", "GET_ROWS", "CPY", "CONT", "SWIGLU",
            "ARGSORT", "SUM_ROWS", "CLAMP"]


@pytest.mark.parametrize("op", GLUE_OPS)
def test_backend_ops_glue(op):
    """the ops between the hot ops of a llama-architecture graph (b200_glue.cu), so that the whole decode graph
    stays on the device.  Cases the plugin declines (other types, non-contiguous rows) run nowhere and are not counted; every case
    it claims must pass against ggml-cpu"""
    rc, txt = run_harness(op, tag="_glue")
    ok, passed, total, fail, holes = summarize(txt, lambda l: False)
    assert rc == 0 and passed == total and not fail, "\n".join(fail[:20])
    assert ok >= 2, f"only {ok} {op} cases ran on the device"       # (the count includes the harness's final "Backend ...: OK" line)


@pytest.mark.parametrize("wtype,K,M,chunk", [(14, 4096, 1024, 1 << 20), (14, 4096, 64, 100000), (8, 4096, 512, 65536 + 34),
                                             (2, 2048, 256, 18 * 1000), (12, 4096, 256, 1 << 18), (14, 14336, 128, 1 << 22)])
def test_chunked_async_upload_repacks(wtype, K, M, chunk):
    """llama_model_loader without mmap uploads every weight in ascending chunks through set_tensor_async
    (src/llama-model-loader.cpp:1136-1186): the chunked upload must end in the device layout, read back as GGUF bytes and give
    the same MUL_MAT bits as a whole-tensor write (tests/upload_probe.cpp, built from include/ggml-b200-abi.h only)"""
    exe = os.path.join(OUT, "upload_probe")
    os.makedirs(OUT, exist_ok=True)
    if not os.path.exists(exe):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "upload_probe.cpp"),
                               "-o", exe, "-ldl"])
    p = subprocess.run([exe, PLUGIN, str(wtype), str(K), str(M), str(chunk)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert p.returncode == 0 and "upload_probe: OK" in p.stdout, p.stdout[-2000:]
    if wtype in (14, 8, 2):                                   # types with a device layout (b200_repack_supported)
        assert "repacked chunked=1 whole=1" in p.stdout, p.stdout


@pytest.mark.parametrize("ndev,wtype,K,M,split", [(2, 12, 4096, 4096, []), (4, 14, 4096, 1024, []), (3, 12, 2048, 6144, [3, 1, 2]), (8, 14, 4096, 14336, []),
                                                   (2, 8, 1024, 640, []), (4, 2, 512, 96, [])])
def test_split_buffer_type_row_slices(ndev, wtype, K, M, split):
    """split-mode=row through the plugin ABI (tests/split_probe.cpp): the registry's "ggml_backend_split_buffer_type" (src/llama-model.cpp:393-411)
    with ndev devices listed on the GPUs present (GGML_B200_FAKE_DEVICES): whole-tensor set / get, per-slice load-time repack, MUL_MAT with
    every device in turn as the main one, decode column and prefill batch, against the same weight in a plain buffer"""
    exe = os.path.join(OUT, "split_probe")
    os.makedirs(OUT, exist_ok=True)
    if not os.path.exists(exe):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "split_probe.cpp"), "-o", exe, "-ldl"])
    p = subprocess.run([exe, PLUGIN, str(wtype), str(K), str(M)] + [str(v) for v in split], env=dict(os.environ, GGML_B200_FAKE_DEVICES=str(ndev)),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert p.returncode == 0 and "split_probe: OK" in p.stdout, p.stdout[-3000:]


@pytest.mark.parametrize("vtype", [12, 14])
def test_layer_fusions_bit_identical(vtype):
    """what be_graph_optimize / be_graph_compute do to the head of a llama layer -- weight matmuls of one activation pulled together and
    launched as one, ADD + RMS_NORM + MUL as one kernel, ROPE(Q) + ROPE(K) + SET_ROWS(K) + SET_ROWS(V) as one kernel, ffn_gate + ffn_up + SwiGLU as one GEMV launch, the graph recorded
    into a CUDA graph and replayed -- changes no bit of any node's output against the same nodes run one by one
    (tests/fusion_probe.cpp, built from include/ggml-b200-abi.h only)"""
    exe = os.path.join(OUT, "fusion_probe")
    os.makedirs(OUT, exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "fusion_probe.cpp"), "-o", exe, "-ldl"])
    outs = []
    # "fused": every fusion but the persistent decode chain (bit-identical to node by node); "chain": the chain too (its GEMV sums in another
    # order: compared within 1e-6 NMSE per dump); "plain": node by node
    for tag, extra in (("fused", {"GGML_B200_GRAPH_STATS": "1", "GGML_B200_NO_CHAIN": "1"}), ("plain", {"GGML_B200_NO_FUSE": "1", "GGML_B200_GRAPHS": "0"}),
                       ("chain", {"GGML_B200_GRAPH_STATS": "1"})):
        f = os.path.join(OUT, f"fusion_probe_{vtype}_{tag}.bin")
        p = subprocess.run([exe, PLUGIN, f, str(vtype)], env=dict(os.environ, **extra), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert p.returncode == 0 and "fusion_probe: OK" in p.stdout, p.stdout[-2000:]
        if tag == "fused":
            order = [l for l in p.stdout.splitlines() if "node order" in l][0].split(":")[-1].split()
            iq, ik, iv = order.index("Qcur"), order.index("Kcur"), order.index("Vcur")
            assert max(iq, ik, iv) - min(iq, ik, iv) == 2, order            # the three matmuls adjacent
            assert abs(iq - ik) == 1 or vtype == 12, order                  # equal weight types side by side
            assert "gate+up+glu 2" in p.stdout and "rope/set_rows 4" in p.stdout and "add+norm+mul 4" in p.stdout, p.stdout   # every fusion fired (2 node walks)
            assert "1 recorded, 4 replayed" in p.stdout, p.stdout           # first run launch by launch, second recorded, four graph launches
        if tag == "chain":
            assert "decode chains 2" in p.stdout, p.stdout                  # ffn_gate / ffn_up / SwiGLU / ffn_down went out as one kernel (2 node walks)
        outs.append(open(f, "rb").read())
        os.remove(f)
    assert len(outs[0]) == len(outs[1]) == len(outs[2]) and len(outs[0]) > 0
    assert outs[0] == outs[1]
    # the dump is f32 node outputs and two f16 caches; the chain only touches f32 columns, the rest stays bit-identical
    import numpy as np
    a32, c32 = np.frombuffer(outs[1], dtype=np.float32), np.frombuffer(outs[2], dtype=np.float32)
    same = a32.view(np.uint32) == c32.view(np.uint32)
    diff = ~same
    assert diff.sum() > 0                                                   # (it really ran another kernel)
    ref = a32[diff].astype(np.float64); got = c32[diff].astype(np.float64)
    assert np.isfinite(got).all() and ((ref - got) ** 2).sum() / (ref ** 2).sum() < 1e-6

