#!/bin/bash
# quick iteration: selected GPU tests + bench (no CPU baseline / pp unless asked)
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 900 ${PYTEST_ARGS:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py ${BENCH_ARGS:---steps 32 --warmup 3 --no-pp --no-cpu} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench.json'))
    print('tg tok/s', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'hbm_frac_step', round(d['hbm_frac_whole_step'],3))
    print('roofline', {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['roofline'].items() if k in('achieved','frac','ms_per_step','launches_per_step')})
    print('kernel ms', {k:round(v,3) for k,v in d['kernel_ms_per_step_event_bracketed'].items()}, 'launches/step', d['gpu_launches_per_step'])
    if 'pp512' in d: print('pp512', round(d['pp512']['value'],1), d['pp512']['roofline']['frac'], d['pp512']['kernel_ms_per_step'])
except Exception as e: print('bench parse failed', e)
PY
tail -5 gpurun_out/bench.err
