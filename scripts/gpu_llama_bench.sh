#!/bin/bash
# The reference's own llama-bench (oracle/_ref, compiled in place, unmodified) on a random-init GGUF: ggml-cpu vs the B200 plugin loaded
# through GGML_BACKEND_PATH (ops the plugin does not claim run on the CPU backend via ggml_backend_sched).
# usage: gpu_llama_bench.sh [model=tinyllama] [quant=q4_0] [extra llama-bench args]
set -u
MODEL=${1:-tinyllama}; QUANT=${2:-q4_0}; shift 2 2>/dev/null
mkdir -p gpurun_out
G=/tmp/$MODEL-$QUANT.gguf
python scripts/make_gguf.py --model $MODEL --quant $QUANT ${LAYERS:+--layers $LAYERS} -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
if [ -z "${SKIP_CPU:-}" ]; then
  timeout 900 $LB -m $G -p 512 -n 128 -r ${REPS:-2} -t $(nproc) -fa 1 -ngl 0 "$@" > gpurun_out/llama_bench_cpu_$MODEL-$QUANT.txt 2>&1; echo "cpu rc=$?"; tail -6 gpurun_out/llama_bench_cpu_$MODEL-$QUANT.txt
fi
GGML_B200_GRAPH_STATS=1 GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 900 $LB -m $G -p 512 -n 128 -r ${REPS:-2} -t $(nproc) -fa 1 -ngl 99 "$@" > gpurun_out/llama_bench_b200_$MODEL-$QUANT.txt 2>&1; echo "b200 rc=$?"; tail -12 gpurun_out/llama_bench_b200_$MODEL-$QUANT.txt
