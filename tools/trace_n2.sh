#!/bin/bash
# Phase timelines (tools/trace_demo.py) + heartbeat check at N ranks; one process group per config.
N=${1:-2}
mkdir -p gpurun_out/trace
for c in hogwild_wide_mlp adag_wide_mlp ssgd_lr_gpt2_medium; do
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29541 tools/trace_demo.py --configs $c > gpurun_out/trace/$c.log 2>&1
  echo "$c rc=$?"; grep -E "^TRACE" gpurun_out/trace/$c.log | tail -1; grep -E "Error|error" gpurun_out/trace/$c.log | tail -3
done
timeout 90 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
  --master-port 29542 tools/heartbeat_check.py > gpurun_out/trace/heartbeat.log 2>&1
echo "heartbeat rc=$?"; grep -E "^HEARTBEAT|Error" gpurun_out/trace/heartbeat.log | tail -5
