#!/bin/bash
# bench.py under torchrun on all the GPUs of the box, as the driver launches it for the scaling run
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
NG=$(nvidia-smi --query-gpu=index --format=csv,noheader | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NG --steps 64 --warmup 5 --no-llama-bench > gpurun_out/bench_${NG}gpu.json 2> gpurun_out/bench_${NG}gpu.err; echo "bench $NG rc=$?"
head -c 700 gpurun_out/bench_${NG}gpu.json; echo; grep -a "error" gpurun_out/bench_${NG}gpu.err | head -3
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${NG}gpu.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "n_gpus", "ms_per_step", "scaling", "split_check", "e2e")})
except Exception as ex:
    print("no json line:", ex)
PY
