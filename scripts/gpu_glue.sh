#!/bin/bash
# the reference's test-backend-ops for the glue ops the plugin claims, then llama-bench through the plugin
set -u
mkdir -p gpurun_out
for op in RMS_NORM RMS_NORM_MUL_ADD ADD MUL SUB DIV SCALE ROPE SET_ROWS GET_ROWS CPY CONT DUP GLU; do
  GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so timeout 600 oracle/_ref/avx2/test-backend-ops test -b B200_0 -o $op > gpurun_out/tbo_$op.log 2>&1; rc=$?
  echo "$op rc=$rc $(grep -c 'OK' gpurun_out/tbo_$op.log) ok-lines, $(grep -c 'not supported' gpurun_out/tbo_$op.log) unsupported, $(grep -c 'FAIL' gpurun_out/tbo_$op.log) FAIL; $(grep 'tests passed' gpurun_out/tbo_$op.log | head -1)"
  grep -m3 "FAIL" gpurun_out/tbo_$op.log | sed 's/\x1b\[[0-9;]*m//g' | cut -c1-220
done
