#!/bin/bash
# why does -sm row arrive as ~50 scheduler splits per token?  GGML_SCHED_DEBUG of one short run
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
G=/tmp/8b-q4_k_m.gguf
python scripts/make_gguf.py --model 8b --quant q4_k_m --layers 4 -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
NG=$(nvidia-smi --query-gpu=index --format=csv,noheader | wc -l)
FAKE=""; [ "$NG" -lt 2 ] && FAKE="GGML_B200_FAKE_DEVICES=2"
env $FAKE GGML_SCHED_DEBUG=2 GGML_B200_GRAPH_STATS=1 GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 300 $LB -v -m $G -p 0 -n 2 -r 1 -t 8 -fa 1 -ngl 99 -sm row > gpurun_out/split_dbg.log 2>&1
grep -c "SPLIT #" gpurun_out/split_dbg.log
grep -a "SPLIT #" gpurun_out/split_dbg.log | tail -30
tail -c 20000 gpurun_out/split_dbg.log > gpurun_out/split_dbg_tail.log
