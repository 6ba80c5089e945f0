#!/bin/bash
# one gpurun call for the decode chain: parity tests, then the micro-benchmark with its ablations
# usage: gpu_chain.sh ["ENV=V ENV2=V" ...]   each argument = one extra chain_bench run under that environment
set -u
mkdir -p gpurun_out; rm -f gpurun_out/chain_bench.jsonl
timeout 600 python -m pytest tests/test_gpu_chain.py -m gpu -x -q --timeout 300 > gpurun_out/chain_pytest.log 2>&1; echo "chain pytest rc=$?"; tail -${TAILN:-4} gpurun_out/chain_pytest.log
timeout 300 python scripts/chain_bench.py --model 8b --mix q4_k_m ${TRACE:---trace} 2>&1 | tee gpurun_out/chain_bench.log
for extra in "$@"; do
  env $extra timeout 300 python scripts/chain_bench.py --model 8b --mix q4_k_m --no-separate ${XTRACE:-} --tag "$extra" 2>&1 | tee -a gpurun_out/chain_bench.log
done
timeout 300 python scripts/chain_bench.py --model 70b --mix q4_k_m --layers 4 --no-separate 2>&1 | tee -a gpurun_out/chain_bench.log
