#!/bin/bash
# usage (gpurun --gpus N): gpu_reserve_ab.sh N "reserve:max_ctas ..."  -- N-GPU bench, SMs reserved for NCCL x NCCL CTA cap
N=${1:-2}
mkdir -p gpurun_out
for rc in ${2:-0:0 4:4 8:8 8:0}; do
  r=${rc%%:*}; c=${rc##*:}
  UNETB200_RESERVE_SMS=$r UNETB200_NCCL_MAX_CTAS=$c timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
      --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 30 --warmup 5 --no-tconv-block \
      > gpurun_out/reserve_ab_n${N}_r${r}_c$c.json 2> gpurun_out/reserve_ab.err || tail -3 gpurun_out/reserve_ab.err
  python - gpurun_out/reserve_ab_n${N}_r${r}_c$c.json $r $c <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(f"reserve {sys.argv[2]:>2s} max_ctas {sys.argv[3]:>2s}: {d['ms_per_step']:.3f} ms/step  {d['value']:.0f} img/s  "
      f"allreduce alone {d['roofline']['collective']['ms_back_to_back']:.3f} ms")
PY
done
