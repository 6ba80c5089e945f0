#!/bin/bash
# one GPU: ncu --set full of the row-split form of the decode chain at world = 1 (template flag SP; the kernel the ranks run, minus the peers)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
FORMS=split timeout 200 python scripts/chain_split_check.py > gpurun_out/split_form.log 2>&1; echo "plain rc=$?"; tail -1 gpurun_out/split_form.log
FORMS=split timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_decode_chain -s 4 -c 2 -o gpurun_out/ncu_full_chain_split -f python scripts/chain_split_check.py > gpurun_out/ncu_full_chain_split.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/ncu_full_chain_split.ncu-rep --page raw --csv > gpurun_out/ncu_full_chain_split.csv 2>/dev/null; wc -l gpurun_out/ncu_full_chain_split.csv
