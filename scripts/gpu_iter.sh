#!/bin/bash
# quick iteration call: GEMV parity tests + timelines.  usage: gpu_iter.sh [pytest -k expr]
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -x -q --timeout 600 ${1:+-k "$1"} > gpurun_out/iter_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/iter_pytest.log
run() { tag=$1; shift; env "$@" timeout 300 python scripts/gemv_trace.py --layers 32 --tag $tag > gpurun_out/$tag.txt 2> gpurun_out/$tag.err || echo "$tag failed rc=$?"; grep "^##" gpurun_out/$tag.txt; }
run i_default X=1
run i_fq16k   GGML_B200_FUSED_QUANT_MAX_K=16384
run i_allmma  GGML_B200_GEMV_MMA=1 GGML_B200_FUSED_QUANT_MAX_K=16384
sed -n 5,26p gpurun_out/i_fq16k.txt | cut -c1-230
