#!/bin/bash
# BASELINE.json configs[3]: Llama-3-70B Q4_K_M row-split over all the GPUs of the box, bench.py under torchrun (MODE=nochain: separate row-split GEMV launches)
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
NG=$(nvidia-smi --query-gpu=index --format=csv,noheader | wc -l)
EXTRA=""; [ "${MODE:-chain}" = nochain ] && EXTRA="B200_BENCH_NO_SPLIT_CHAIN=1"
env $EXTRA timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29513 bench.py --model 70b --gpus $NG --steps 32 --warmup 3 --no-llama-bench --no-pp > gpurun_out/bench70b_${NG}gpu_${MODE:-chain}.json 2> gpurun_out/bench70b_${NG}gpu_${MODE:-chain}.err; echo "bench 70b $NG ${MODE:-chain} rc=$?"
grep -a "error\|declined\|Traceback\|Error" gpurun_out/bench70b_${NG}gpu_${MODE:-chain}.err | head -5
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench70b_${NG}gpu_${MODE:-chain}.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "n_gpus", "ms_per_step", "split_check", "gpu_launches")}, d.get("e2e", {}).get("value"), d.get("roofline", {}).get("frac"))
except Exception as ex:
    print("no json line:", ex)
PY
