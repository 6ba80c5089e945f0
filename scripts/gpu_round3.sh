#!/bin/bash
# one gpurun call: new parity tests, then the bench (both arms)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_gpu_rowsplit_one_gpu.py -x -q > gpurun_out/t_split1.log 2>&1; echo "split1 rc=$?"; tail -5 gpurun_out/t_split1.log
timeout 600 python -m pytest tests/test_gpu_fattn.py -x -q -k 8k_depth > gpurun_out/t_fa8k.log 2>&1; echo "fa8k rc=$?"; tail -5 gpurun_out/t_fa8k.log
timeout 900 python -m pytest tests/test_gpu_backend_ops.py -x -q -k "ARGSORT or SUM_ROWS or CLAMP" > gpurun_out/t_glue.log 2>&1; echo "glue rc=$?"; tail -5 gpurun_out/t_glue.log
timeout 1500 python -m pytest tests/test_gpu_llama_logits.py -x -q -k "70b or mixtral" > gpurun_out/t_logits.log 2>&1; echo "logits rc=$?"; tail -15 gpurun_out/t_logits.log
timeout 1500 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench.err; head -c 6000 gpurun_out/bench.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"; tail -c 1000 gpurun_out/bench_reference.err; cat gpurun_out/bench_reference.json
