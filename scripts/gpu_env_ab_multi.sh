#!/bin/bash
# usage (gpurun --gpus N): gpu_env_ab_multi.sh N "VAR=1" ...  -- N-GPU bench with and without environment settings
N=${1:-2}; shift
mkdir -p gpurun_out
for env in "" "$@"; do
  tag=$(echo "$env" | tr -c 'A-Za-z0-9=\n' '_')
  env $env timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 29561 bench.py --gpus $N --steps 30 --warmup 5 --no-tconv-block > gpurun_out/mab_n${N}_$tag.json 2> gpurun_out/mab.err || tail -3 gpurun_out/mab.err
  python - gpurun_out/mab_n${N}_$tag.json "$env" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(f"[{sys.argv[2] or 'default'}] {d['ms_per_step']:.3f} ms/step  {d['value']:.0f} img/s")
PY
done
