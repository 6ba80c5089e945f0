#!/bin/bash
# one gpurun call: the reference's llama-bench tg128 through the plugin under several environment settings, same box, same GGUF
# usage: gpu_lb_ab.sh "ENV=V ..." ["ENV=V ..."]   (the first run is always the default; PLUGIN=path in a setting swaps the library)
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
G=/tmp/8b-q4_k_m.gguf
python scripts/make_gguf.py --model 8b --quant q4_k_m -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
run() { env GGML_B200_GRAPH_STATS=1 GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so "$@" timeout 600 $LB -m $G -p ${PP:-0} -n ${NTOK:-128} -r 3 -t $(nproc) -fa 1 -ngl 99 2>&1 | grep -E "tg128|tg8|pp512|replays timed|fusions|declined|decode chain" | tail -${TAILN:-5}; }
echo "== default"; run
for e in "$@"; do echo "== $e"; run $e; done
