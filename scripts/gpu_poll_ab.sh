#!/bin/bash
# row-split chain: pause between failed polls of a hand-over record (GGML_B200_CHAIN_POLL_NS), bench.py under torchrun on all GPUs of the box
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
NG=$(nvidia-smi --query-gpu=index --format=csv,noheader | wc -l)
for ns in ${POLLS:-0 50 200 500}; do
  GGML_B200_CHAIN_POLL_NS=$ns timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NG --steps 64 --warmup 5 --no-llama-bench --no-pp > gpurun_out/poll_$ns.json 2> gpurun_out/poll_$ns.err
  python -c "
import json; d=json.loads(open('gpurun_out/poll_$ns.json').read().strip().splitlines()[-1]); print('poll_ns $ns:', round(d['value'],1), 'tok/s', round(d['ms_per_step'],4), 'ms', d['split_check']['nmse_vs_unsplit_max'])"
done
