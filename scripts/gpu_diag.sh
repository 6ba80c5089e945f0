#!/bin/bash
# one gpurun call: device-side timelines + graph timings of the decode MUL_MAT chain under the tuning switches
set -u
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python scripts/gemv_trace.py --layers ${LAYERS:-32} --tag $tag > gpurun_out/$tag.txt 2> gpurun_out/$tag.err || echo "$tag failed rc=$?"; grep "^##" gpurun_out/$tag.txt; }
run t_default   X=1
run t_fq16k     GGML_B200_FUSED_QUANT_MAX_K=16384
run t_npf32k    GGML_B200_GS_NPF=32768 GGML_B200_GQ_NPF=8
run t_npf32k_fq GGML_B200_GS_NPF=32768 GGML_B200_GQ_NPF=8 GGML_B200_FUSED_QUANT_MAX_K=16384
run t_nomma     GGML_B200_GEMV_MMA=0
run t_allmma    GGML_B200_GEMV_MMA=1
run t_nopdl     GGML_B200_NO_PDL=1
run t_noq       GGML_B200_GS_DBG=1
run t_nodot     GGML_B200_GS_DBG=2
head -24 gpurun_out/t_default.txt
