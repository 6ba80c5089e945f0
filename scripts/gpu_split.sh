#!/bin/bash
# -sm row through the plugin: logits parity (split buffer type on 2 / 4 listed devices), the split test of test-backend-ops' harness, and
# the reference's llama-bench with the row-split MUL_MATs recorded into CUDA graphs (default) against launch by launch
# (GGML_B200_SPLIT_GRAPHS=0).  On a one-GPU box the two devices are listed on the same GPU (GGML_B200_FAKE_DEVICES=2).
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
NG=$(nvidia-smi --query-gpu=index --format=csv,noheader | wc -l)
echo "gpus: $NG"
timeout 900 python -m pytest tests/test_gpu_llama_logits.py -x -q -k "split_mode_row" > gpurun_out/split_logits.log 2>&1; echo "logits rc=$?"; tail -5 gpurun_out/split_logits.log
timeout 600 python -m pytest tests/test_gpu_backend_ops.py -x -q -k "split" > gpurun_out/split_ops.log 2>&1; echo "ops rc=$?"; tail -3 gpurun_out/split_ops.log
G=/tmp/8b-q4_k_m.gguf
python scripts/make_gguf.py --model 8b --quant q4_k_m -o $G || exit 1
LB=oracle/_ref/avx2/llama-bench
FAKE=""; [ "$NG" -lt 2 ] && FAKE="GGML_B200_FAKE_DEVICES=2"
for sg in 1 0; do
  echo "== llama-bench -sm row, GGML_B200_SPLIT_GRAPHS=$sg $FAKE"
  env $FAKE GGML_B200_SPLIT_GRAPHS=$sg GGML_B200_GRAPH_STATS=1 GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 $LB -m $G -p 512 -n 128 -r 2 -t $(nproc) -fa 1 -ngl 99 -sm row > gpurun_out/split_lb_$sg.log 2>&1
  grep -E "pp512|tg128|error|failed|recorded" gpurun_out/split_lb_$sg.log | tail -6
done
echo "== llama-bench -sm none"
GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 $LB -m $G -p 512 -n 128 -r 2 -t $(nproc) -fa 1 -ngl 99 -sm none 2>&1 | grep -E "pp512|tg128|error|failed" | tail -3
if [ "$NG" -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 64 --warmup 5 --no-llama-bench > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench 2 rc=$?"; head -c 1500 gpurun_out/bench_2gpu.json; grep -a "error" gpurun_out/bench_2gpu.err | head -3
fi
