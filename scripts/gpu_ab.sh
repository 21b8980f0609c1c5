#!/bin/bash
# A/B a bench run under different environment settings: usage gpu_ab.sh "VAR=1" "VAR=2" ...
mkdir -p gpurun_out
for env in "" "$@"; do
  out=gpurun_out/ab_$(echo "$env" | tr -c 'A-Za-z0-9=\n' '_').json
  env $env timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline $AB_ARGS > $out 2> gpurun_out/ab.err || tail -3 gpurun_out/ab.err
  python - "$out" "$env" <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); r=d["roofline"]
print(f"[{sys.argv[2] or 'default'}] ms {d['ms_per_step']:.3f} img/s {d['value']:.0f} igemm {r['ms_per_step']:.3f} wgrad {r['wgrad']['ms_per_step']:.3f}")
PY
done
