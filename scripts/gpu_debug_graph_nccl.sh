#!/bin/bash
# usage (gpurun --gpus 2): each NCCL-in-graph variant under its own short timeout; logs -> gpurun_out/dbg_<mode>.log
mkdir -p gpurun_out
for mode in ${@:-default}; do
  NCCL_DEBUG=WARN timeout -k 5 80 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
      --master-addr 127.0.0.1 --master-port 29533 scripts/debug_graph_nccl.py > gpurun_out/dbg_$mode.log 2>&1
  echo "== $mode exit $?"; grep -E "rank 0|Error|error|Traceback|File" gpurun_out/dbg_$mode.log | tail -12
done
