#!/bin/bash
# 2-GPU experiments: peer-exchange parity test, then bench variants; everything goes to gpurun_out/
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_rowsplit_peer.py -m gpu -x -q > gpurun_out/peer_test.log 2>&1; grep -E "rank|passed|failed" gpurun_out/peer_test.log | head
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus ${NG:-2} --steps 32 --warmup 3 --no-cpu --no-pp"
show() { grep '^{' "$1" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$2', round(d['value'],1), 'tok/s', round(d['ms_per_step'],3), 'ms', {k:round(v,2) for k,v in d['kernel_ms_per_step_event_bracketed'].items()})"; }
for x in ${XS:-0 1}; do GGML_B200_XDBG=$x timeout 300 $R > gpurun_out/n2_x$x.json 2> gpurun_out/n2_x$x.err; show gpurun_out/n2_x$x.json xdbg$x; done
B200_BENCH_NCCL=1 timeout 300 $R > gpurun_out/n2_nccl.json 2> gpurun_out/n2_nccl.err; show gpurun_out/n2_nccl.json nccl
