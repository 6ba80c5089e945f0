#!/bin/bash
# One gpurun call (1 GPU): the round's closing pass -- whole GPU suite, smoke, bench (both arms), then the ncu launch list of the bench
# command and one --set full capture of the decode chain (each only after its own command ran clean without ncu).
# scripts/collect_profiles.py <tag> turns gpurun_out/ into profiles/<tag>_*.
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
bash scripts/gpu_check.sh
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-graph --no-llama-bench"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/ncu_launches.csv $B --no-pp > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_decode_chain -s 4 -c 2 -o gpurun_out/ncu_full_chain -f $B --no-pp > gpurun_out/ncu_full_chain.log 2>&1; echo "ncu chain rc=$?"
ls -la gpurun_out/*.ncu-rep
