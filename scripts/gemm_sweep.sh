#!/bin/bash
# timing experiments for the prefill GEMM: env-forced tile shapes / pipeline depths / debug modes
S="--types q4_K --n 512 --shapes 4096x14336,4096x4096,14336x4096"
run() { echo "== $*"; env "$@" timeout 200 python scripts/gemv_bench.py $S 2>&1 | grep TFLOP | tr '\n' ' '; echo; }
run A=0
run GGML_B200_GEMM_MT=1 GGML_B200_GEMM_BN=256
run GGML_B200_GEMM_MT=1 GGML_B200_GEMM_BN=256 GGML_B200_GEMM_RS=2 GGML_B200_GEMM_SB=6
run GGML_B200_GEMM_MT=1 GGML_B200_GEMM_BN=128
run GGML_B200_GEMM_MT=2 GGML_B200_GEMM_BN=128
run GGML_B200_GEMM_MT=1 GGML_B200_GEMM_BN=128 GGML_B200_GEMM_DBG=3
run GGML_B200_GEMM_MT=1 GGML_B200_GEMM_BN=128 GGML_B200_GEMM_DBG=7
run GGML_B200_GEMM_MT=1 GGML_B200_GEMM_BN=64
