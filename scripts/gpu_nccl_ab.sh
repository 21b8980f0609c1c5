#!/bin/bash
# usage (gpurun --gpus N): gpu_nccl_ab.sh N "0 2 4 8"   -- N-GPU bench under different NCCL CTA limits
N=${1:-2}
mkdir -p gpurun_out
for c in ${2:-0 2 4 8}; do
  UNETB200_NCCL_MAX_CTAS=$c timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 29541 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/nccl_ab_n${N}_c$c.json 2> gpurun_out/nccl_ab.err || tail -3 gpurun_out/nccl_ab.err
  python - gpurun_out/nccl_ab_n${N}_c$c.json $c <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(f"max_ctas {sys.argv[2]:>2s}: ups {d['ms_per_step']:.3f} ms  tconv {d.get('tconv', {}).get('ms_per_step', 0):.3f} ms  "
      f"allreduce alone {d['roofline']['collective']['ms_back_to_back']:.3f} ms")
PY
done
