#!/bin/bash
# where a tg128 token's wall time goes: device time of the replayed graph vs host time inside the plugin (GGML_B200_GRAPH_STATS)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
G=/tmp/8b-q4_k_m.gguf
python scripts/make_gguf.py --model 8b --quant q4_k_m -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
GGML_B200_GRAPH_STATS=1 GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 $LB -m $G -p 0 -n 128 -r 3 -t $(nproc) -fa 1 -ngl 99 > gpurun_out/lb_host.log 2>&1
grep -a -E "tg128|ggml-b200" gpurun_out/lb_host.log | tail -12
