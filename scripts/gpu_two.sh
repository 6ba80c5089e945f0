#!/bin/bash
# one gpurun --gpus 2 call: what needs a second GPU -- the peer-exchange test, logits with -sm row / -sm layer on two real devices, the
# reference's llama-bench with -sm row and -sm layer through the plugin, bench.py --gpus 2 under torchrun
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 600 python -m pytest tests/test_gpu_rowsplit_peer.py -x -q > gpurun_out/two_peer.log 2>&1; echo "peer rc=$?"; tail -3 gpurun_out/two_peer.log
timeout 900 python -m pytest tests/test_gpu_llama_logits.py -x -q -k "split_mode_row or (8b and 250)" > gpurun_out/two_logits.log 2>&1; echo "logits rc=$?"; tail -3 gpurun_out/two_logits.log
G=/tmp/8b-q4_k_m.gguf
python scripts/make_gguf.py --model 8b --quant q4_k_m -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
for sm in none layer row; do
  echo "== llama-bench -sm $sm (2 devices listed)"
  GGML_B200_GRAPH_STATS=1 GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 $LB -m $G -p 512 -n 128 -r 2 -t $(nproc) -fa 1 -ngl 99 -sm $sm 2>&1 | grep -E "pp512|tg128|error|failed" | tail -3
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 64 --warmup 5 --no-llama-bench > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench 2 rc=$?"; head -c 1800 gpurun_out/bench_2gpu.json; tail -c 400 gpurun_out/bench_2gpu.err
