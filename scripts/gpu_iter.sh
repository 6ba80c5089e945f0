#!/bin/bash
# quick iteration call: GEMV parity tests + timelines.  usage: gpu_iter.sh [skip | pytest -k expr]
set -u
mkdir -p gpurun_out
if [ "${1:-}" != "skip" ]; then timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -x -q --timeout 600 ${1:+-k "$1"} > gpurun_out/iter_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/iter_pytest.log; fi
run() { tag=$1; shift; env "$@" timeout 300 python scripts/gemv_trace.py --layers 32 --tag $tag > gpurun_out/$tag.txt 2> gpurun_out/$tag.err || echo "$tag failed rc=$?"; grep "^##" gpurun_out/$tag.txt; }
run i_default X=1

python scripts/trace_agg.py i_default
