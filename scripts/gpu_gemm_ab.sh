#!/bin/bash
# one gpurun call: prefill GEMM shapes of a Llama-3-8B layer at N = 512 under environment settings (scripts/gemv_bench.py)
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
S="--n 512 --iters 100 --shapes 4096x4096,4096x5120,4096x6144,4096x28672,14336x4096"
run() { echo "== $*"; env "$@" timeout 300 python scripts/gemv_bench.py --types ${TYPES:-q4_K} $S 2>&1 | grep -E "TFLOP" | tr '\n' ' '; echo; }
run A=0
for e in "$@"; do run $e; done
