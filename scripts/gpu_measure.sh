#!/bin/bash
# One gpurun call (1 GPU): the measurement pass -- bench line (both arms), tcgen05 issue-rate ceilings, ncu launch list, ncu --set full of the
# dominant kernels.  Results land in gpurun_out/; scripts/collect_profiles.py <tag> turns them into profiles/<tag>_*.
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,driver_version,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
lscpu | head -20 > gpurun_out/cpu.txt; nproc >> gpurun_out/cpu.txt; free -g >> gpurun_out/cpu.txt
if [ -n "${PYTEST_K:-}" ]; then timeout 1200 python -m pytest tests -m gpu -x -q -k "$PYTEST_K" > gpurun_out/pytest_k.log 2>&1; echo "pytest -k rc=$?"; tail -5 gpurun_out/pytest_k.log; fi
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Isuper-llama.cpp_b200/csrc -Iinclude scripts/umma_peak.cu -o gpurun_out/umma_peak && timeout 120 gpurun_out/umma_peak > gpurun_out/umma_peak.txt 2>&1; cat gpurun_out/umma_peak.txt
timeout 1500 python bench.py ${BENCH_ARGS:-} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench.err; head -c 7000 gpurun_out/bench.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"; head -c 600 gpurun_out/bench_reference.json
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --no-llama-bench"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/ncu_launches.csv $B --no-pp > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_decode_chain -s 4 -c 2 -o gpurun_out/ncu_full_chain -f $B --no-pp > gpurun_out/ncu_full_chain.log 2>&1; echo "ncu chain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemm_q -s 7 -c 8 -o gpurun_out/ncu_full_gemm -f $B > gpurun_out/ncu_full_gemm.log 2>&1; echo "ncu gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fattn_tc -s 2 -c 2 -o gpurun_out/ncu_full_fattn_tc -f $B > gpurun_out/ncu_full_fattn_tc.log 2>&1; echo "ncu fattn_tc rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fattn_gqa -s 40 -c 2 -o gpurun_out/ncu_full_fattn_gqa -f $B --no-pp --depth 8192 > gpurun_out/ncu_full_fattn_gqa.log 2>&1; echo "ncu fattn_gqa rc=$?"
B200_BENCH_NO_CHAIN=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_gemv_q4k_mma|k_gemv_stream' -s 20 -c 6 -o gpurun_out/ncu_full_gemv -f $B --no-pp > gpurun_out/ncu_full_gemv.log 2>&1; echo "ncu gemv rc=$?"
ls -la gpurun_out/*.ncu-rep
