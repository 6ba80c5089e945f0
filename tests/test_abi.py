"""CPU tests (-m "not gpu") for the drop-in boundary:
 * include/ggml-b200-abi.h restates the reference's plugin ABI (ggml-backend-impl.h, ggml.h) byte for byte:
   tests/abi_probe.cpp is compiled against both and the layouts are diffed; golden/abi_layout_reference.txt is the
   reference's output committed for machines without /root/reference;
 * libggml-b200.so loads and exports every symbol include/ggml-b200.h declares (no compute without a GPU);
 * the product has no CPU fallback: creating a context without a device fails loudly;
 * the product never touches oracle/.
"""
import ctypes
import os
import re
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")
PKG = os.path.join(ROOT, "super-llama.cpp_b200")
LIB = os.path.join(PKG, "libggml-b200.so")
REF = "/root/reference"


def _probe(tmp_path, name, flags):
    exe = str(tmp_path / name)
    subprocess.check_call(["g++", "-std=c++17", "-O1", os.path.join(ROOT, "tests", "abi_probe.cpp"), "-o", exe] + flags)
    return subprocess.check_output([exe], text=True)


def test_abi_header_matches_committed_reference_layout(tmp_path):
    ours = _probe(tmp_path, "probe_ours", ["-I", INC])
    want = open(os.path.join(ROOT, "tests", "golden", "abi_layout_reference.txt")).read()
    assert ours == want


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present on this machine")
def test_abi_header_matches_reference_headers_live(tmp_path):
    refdir = os.path.join(ROOT, "oracle", "_ref", "avx2")
    if not os.path.exists(os.path.join(refdir, "libggml-ref.so")):
        pytest.skip("oracle/_ref not built")
    ref = _probe(tmp_path, "probe_ref", ["-DPROBE_REFERENCE", "-I", f"{REF}/ggml/include", "-I", f"{REF}/ggml/src",
                                         "-L", refdir, "-lggml-ref", f"-Wl,-rpath,{refdir}"])
    ours = _probe(tmp_path, "probe_ours", ["-I", INC])
    assert ours == ref


def _declared_symbols():
    src = open(os.path.join(INC, "ggml-b200.h")).read()
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"#\s*define[^\n]*", "", src)
    return sorted(set(re.findall(r"B200_API\s+[^;(]*?\b(\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def built_lib():
    if not os.path.exists(LIB):
        subprocess.check_call(["make", "-C", PKG, "-j8"], stdout=subprocess.DEVNULL)
    return LIB


def test_library_exports_every_declared_symbol(built_lib):
    syms = _declared_symbols()
    assert "ggml_backend_init" in syms and "b200_mul_mat" in syms and len(syms) >= 20
    lib = ctypes.CDLL(built_lib)
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, f"declared in include/ggml-b200.h but not exported: {missing}"


def test_no_unintended_exports(built_lib):
    out = subprocess.check_output(["nm", "-D", "--defined-only", built_lib], text=True)
    names = {l.split()[-1] for l in out.splitlines() if l.split()[1] in "TD"}
    extra = {n for n in names if not (n.startswith("b200_") or n.startswith("ggml_backend_"))}
    assert not extra, extra


def test_registry_object_is_well_formed_without_gpu(built_lib):
    """ggml_backend_init() must return a registry with the right api_version even when no device is usable,
    and score 0 so ggml_backend_load() declines it (ggml-backend-reg.cpp:255-290)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = ctypes.CDLL(built_lib)
    lib.ggml_backend_init.restype = ctypes.c_void_p
    reg = lib.ggml_backend_init()
    assert reg
    assert ctypes.cast(reg, ctypes.POINTER(ctypes.c_int))[0] == 2        # GGML_BACKEND_API_VERSION
    assert lib.ggml_backend_score() == 0
    assert lib.b200_device_count() == 0


def test_no_cpu_fallback_without_device(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import ggml_b200 as B
    with pytest.raises(B.B200Error):
        B.Context(0)


def test_product_never_references_oracle():
    bad = []
    for dirpath, _, files in os.walk(PKG):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                if re.search(r"liboracle|ggml_oracle|ref_driver|libggml-ref|import oracle|from oracle", txt):
                    bad.append(f)
    assert not bad, bad


def test_graph_optimize_reorders_weight_matmuls_only(tmp_path):
    """host logic of the plugin's graph_optimize (no GPU): the weight matmuls of one activation become adjacent, equal weight types side by
    side; a matmul over a KV-cache view stays put; every node still follows its sources (tests/reorder_probe.cpp through get_proc_address)"""
    import subprocess
    plugin = os.path.join(ROOT, "super-llama.cpp_b200", "libggml-b200.so")
    if not os.path.exists(plugin):
        pytest.skip("libggml-b200.so not built")
    exe = str(tmp_path / "reorder_probe")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "reorder_probe.cpp"), "-o", exe, "-ldl"])
    for vtype in (12, 14):
        out = subprocess.run([exe, plugin, str(vtype)], stdout=subprocess.PIPE, text=True, timeout=60)
        assert out.returncode == 0 and "dependencies: OK" in out.stdout, out.stdout
        # the CUDA-graph cache key (graph_identity): deterministic, sensitive to everything a recording bakes in, prefill graphs ineligible
        assert "identity: eligible=1 stable=1" in out.stdout and "identity: 9 of 9 mutations changed the key, restored=1" in out.stdout, out.stdout
        assert "identity: prefill eligible=0" in out.stdout
        order = out.stdout.splitlines()[0].split()[1:]
        pos = {n: i for i, n in enumerate(order)}
        for il in (0, 1):
            q, k, v = pos[f"Q{il}"], pos[f"K{il}"], pos[f"V{il}"]
            assert max(q, k, v) - min(q, k, v) == 2, order                         # adjacent
            # ROPE(K) first, ROPE(Q) right behind it, then the two cache writes: one launch for the four (b200_glue_rope_kv)
            assert pos[f"Krope{il}"] > max(q, k, v) and pos[f"Qrope{il}"] == pos[f"Krope{il}"] + 1
            assert pos[f"setK{il}"] == pos[f"Qrope{il}"] + 1 and pos[f"setV{il}"] == pos[f"setK{il}"] + 1
            assert pos[f"KQ{il}"] > pos[f"setV{il}"]                               # not a weight: not hoisted over the cache write
            assert abs(pos[f"up{il}"] - pos[f"gate{il}"]) == 1
        if vtype == 14:
            assert abs(pos["Q1"] - pos["K1"]) == 1                                 # q4_K, q4_K, then the q6_K one
        with_env = subprocess.run([exe, plugin, str(vtype)], stdout=subprocess.PIPE, text=True, timeout=60, env=dict(os.environ, GGML_B200_NO_REORDER="1"))
        assert with_env.stdout.splitlines()[0].split()[1:6] == ["norm0", "Q0", "Q3_0", "Qrope0", "V0"]
