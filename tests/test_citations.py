"""Every reference citation `path/file.ext:line[-line]` in the headers, the oracle, the design documents and the csrc comments points at
a file that exists in the reference and is long enough to have that line (runs only where /root/reference is present: here, not on the GPU box)."""
import glob
import os
import re
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
PAT = re.compile(r"(?<![\w/.-])((?:ggml|src|tools|tests|include|common|gguf-py|examples|benches)/[\w./+-]+\.(?:cpp|cuh|cu|c|h|hpp|py|md)):(\d+)(?:-(\d+))?")


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree is not present")
def test_cited_reference_lines_exist():
    files = [os.path.join(ROOT, f) for f in ("DESIGN.md", "INTEGRATION.md", "README.md")]
    for pat in ("include/*.h", "oracle/*.c", "oracle/*.cpp", "oracle/*.py", "oracle/Makefile", "super-llama.cpp_b200/csrc/*", "super-llama.cpp_b200/*.py", "tests/*.py", "tests/*.cpp", "bench.py"):
        files += glob.glob(os.path.join(ROOT, pat))
    lengths, bad, n = {}, [], 0
    for f in files:
        txt = open(f, errors="replace").read()
        for m in PAT.finditer(txt):
            path, lo, hi = m.group(1), int(m.group(2)), int(m.group(3) or m.group(2))
            full = os.path.join(REF, path)
            n += 1
            if not os.path.isfile(full):
                bad.append(f"{os.path.relpath(f, ROOT)}: {m.group(0)} (no such file)")
                continue
            if full not in lengths:
                lengths[full] = sum(1 for _ in open(full, errors="replace"))
            if hi < lo or hi > lengths[full]:
                bad.append(f"{os.path.relpath(f, ROOT)}: {m.group(0)} (file has {lengths[full]} lines)")
    # short forms (`ggml-cuda.cu:777-803`, `ops.cpp:8047`): some file of that name in the reference must be long enough
    SHORT = re.compile(r"(?<![\w/.-])([\w+-]+\.(?:cpp|cuh|cu|c|h|hpp)):(\d+)(?:-(\d+))?")
    by_name = {}
    for dirpath, _, names in os.walk(REF):
        if "/.git" in dirpath:
            continue
        for nm in names:
            by_name.setdefault(nm, []).append(os.path.join(dirpath, nm))
    own = {os.path.basename(f) for f in files} | {os.path.basename(f) for f in glob.glob(os.path.join(ROOT, "**", "*.*"), recursive=True)}
    ns = 0
    for f in files:
        txt = open(f, errors="replace").read()
        for m in SHORT.finditer(txt):
            name, lo, hi = m.group(1), int(m.group(2)), int(m.group(3) or m.group(2))
            if name in own or name not in by_name:                 # our own files, or not a reference file name at all
                continue
            ns += 1
            best = 0
            for full in by_name[name]:
                if full not in lengths:
                    lengths[full] = sum(1 for _ in open(full, errors="replace"))
                best = max(best, lengths[full])
            if hi < lo or hi > best:
                bad.append(f"{os.path.relpath(f, ROOT)}: {m.group(0)} (longest file of that name has {best} lines)")
    assert n > 50 and ns > 50, (n, ns)                                         # the pattern really finds the citations
    assert not bad, "\n".join(bad[:40])
