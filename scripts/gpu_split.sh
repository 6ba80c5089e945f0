#!/bin/bash
# One gpurun call: the split-mode=row boundary (split buffer type + logits with -sm row), the paged attention tests, then the sanitizers.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q -k "split_buffer_type or split_mode_row or paged" > gpurun_out/t_split.log 2>&1; echo "split rc=$?"; tail -30 gpurun_out/t_split.log
if [ -n "${SANITIZE:-}" ]; then bash scripts/gpu_sanitize.sh; fi
