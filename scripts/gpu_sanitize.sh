#!/bin/bash
# One gpurun call (1 GPU): compute-sanitizer over a reduced set of the GPU parity tests (the reference runs its kernels under
# compute-sanitizer in CI).  memcheck: out-of-bounds / misaligned global, shared and local accesses, leaks of device allocations;
# racecheck: shared-memory hazards (the mbarrier-guarded rings of the TMA kernels are the interesting part).  The test selection touches
# every kernel family once: quantize, GEMV (dp4a, streaming, mma), the decode chain, tcgen05 GEMM, MUL_MAT_ID, the attention kernels
# (vec, gather, GQA, tcgen05 prefill, paged), the row-split exchange.  Summaries -> gpurun_out/sanitize_{memcheck,racecheck}.txt
set -u
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
SEL=${SANITIZE_K:-"golden or small_n or test_mul_mat_prefill_gemm or decode_ragged_rows or tensor_core_gemv_all or gate_up_swiglu or mixtral_decode or decode_chain_vs_oracle or flash_attn_kv_cache_views or prefill_causal or gqa_decode_kernel or paged_rejects or test_rowsplit_n_shards_on_one_gpu"}
for tool in memcheck racecheck; do
  extra=""; [ "$tool" = memcheck ] && extra="--leak-check no"
  timeout ${SANITIZE_TIMEOUT:-1500} compute-sanitizer --tool $tool $extra --target-processes application-only --print-limit 30 --error-exitcode 86 \
      python -m pytest tests -m gpu -q -x -k "$SEL" -p no:cacheprovider > gpurun_out/sanitize_$tool.log 2>&1
  rc=$?
  { echo "# compute-sanitizer --tool $tool $extra -- python -m pytest tests -m gpu -q -x -k \"$SEL\""; echo "# exit code $rc (86 = the tool reported errors; 124 = timeout)";
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" gpurun_out/sanitize_$tool.log | tail -8;
    grep -E "=========.*(Invalid|Race|hazard|Misaligned|out of bounds)" gpurun_out/sanitize_$tool.log | sort | uniq -c | sort -rn | head -20; } > gpurun_out/sanitize_$tool.txt
  echo "$tool rc=$rc"; cat gpurun_out/sanitize_$tool.txt
done
