#!/bin/bash
# one gpurun call: quantized-KV attention parity (flat API + the reference's test-backend-ops through the plugin)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_fattn_kvq.py -x -q > gpurun_out/t_kvq.log 2>&1; echo "kvq rc=$?"; tail -15 gpurun_out/t_kvq.log
timeout 1500 python -m pytest tests/test_gpu_backend_ops.py -x -q -k "flash_attn_ext or SET_ROWS" > gpurun_out/t_kvq_ops.log 2>&1; echo "ops rc=$?"; tail -25 gpurun_out/t_kvq_ops.log
