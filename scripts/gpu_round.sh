#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the bench line (both arms), ncu launch list + one full capture, the reference's own harness.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
lscpu | head -20 > gpurun_out/cpu.txt; nproc >> gpurun_out/cpu.txt; free -g >> gpurun_out/cpu.txt
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 1800 ${PYTEST_ARGS:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py ${BENCH_ARGS:-} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference rc=$?"; cat gpurun_out/bench_reference.json
# ncu: launch list of the bench command (shares, not absolutes), then ONE full capture of the dominant kernels
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/ncu_launches.csv python bench.py --steps 2 --warmup 3 --no-pp --no-cpu --no-graph --no-llama-bench > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_gemv_q4k_mma|k_gemv_stream' -s 20 -c 6 -o gpurun_out/ncu_full_gemv -f python bench.py --steps 2 --warmup 3 --no-pp --no-cpu --no-graph --no-llama-bench > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
if [ -z "${SKIP_TBO:-}" ]; then
for op in MUL_MAT MUL_MAT_ID FLASH_ATTN_EXT; do
  GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 900 oracle/_ref/avx2/test-backend-ops test -b B200_0 -o $op > gpurun_out/tbo_$op.log 2>&1; echo "test-backend-ops $op rc=$?"; tail -3 gpurun_out/tbo_$op.log
done
fi
# the reference's own llama-bench (compiled in place by `make -C oracle llama`), ggml-cpu vs the plugin, on random-init GGUFs
if [ -x oracle/_ref/avx2/llama-bench ] && [ -z "${SKIP_LB:-}" ]; then
  REPS=3 bash scripts/gpu_llama_bench.sh 8b q4_k_m
  REPS=3 bash scripts/gpu_llama_bench.sh tinyllama q4_0
fi
