#!/bin/bash
# bench.py under torchrun on all the GPUs of the box: the row-split decode chain (default) against the separate row-split GEMV launches
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out
NG=$(nvidia-smi --query-gpu=index --format=csv,noheader | wc -l)
for mode in ${MODES:-chain nochain}; do
  EXTRA=""; [ $mode = nochain ] && EXTRA="B200_BENCH_NO_SPLIT_CHAIN=1"; [ $mode = wholeqkv ] && EXTRA="B200_BENCH_WHOLE_QKV=1"
  env $EXTRA timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NG --steps 64 --warmup 5 --no-llama-bench --no-pp > gpurun_out/bench_${NG}gpu_$mode.json 2> gpurun_out/bench_${NG}gpu_$mode.err; echo "bench $NG $mode rc=$?"
  grep -a "error\|declined\|Traceback\|Error" gpurun_out/bench_${NG}gpu_$mode.err | head -5
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_${NG}gpu_$mode.json").read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ("value", "n_gpus", "ms_per_step", "split_check", "gpu_launches")}, d.get("e2e", {}).get("value"), d.get("roofline", {}).get("frac"))
except Exception as ex:
    print("no json line:", ex)
PY
done
