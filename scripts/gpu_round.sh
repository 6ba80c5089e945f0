#!/bin/bash
# One gpurun call: GPU parity tests, smoke, the bench line, the reference's own perf harness.  Logs -> gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
lscpu | head -20 > gpurun_out/cpu.txt; nproc >> gpurun_out/cpu.txt; free -g >> gpurun_out/cpu.txt
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 1800 ${PYTEST_ARGS:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py ${BENCH_ARGS:-} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 oracle/_ref/avx2/test-backend-ops perf -b B200_0 -o MUL_MAT -p "type_a=(q4_0|q8_0|q4_K|q6_K),type_b=f32,m=4096" > gpurun_out/perf_mul_mat.log 2>&1; echo "perf rc=$?"; grep -E "q4_K|q8_0" gpurun_out/perf_mul_mat.log | head -20
