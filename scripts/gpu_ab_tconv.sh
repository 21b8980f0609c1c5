#!/bin/bash
# A/B the transposed-convolution decoder (the `tconv` block of the bench line): usage gpu_ab_tconv.sh "VAR=1" ...
mkdir -p gpurun_out
for env in "" "$@"; do
  out=gpurun_out/abt_$(echo "$env" | tr -c 'A-Za-z0-9=\n' '_').json
  env $env timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $out 2> gpurun_out/abt.err || tail -3 gpurun_out/abt.err
  python - "$out" "$env" <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); t=d["tconv"]
print(f"[{sys.argv[2] or 'default'}] ups {d['ms_per_step']:.3f} ms | tconv {t['ms_per_step']:.3f} ms {t['value']:.0f} img/s frac {t['whole_step_frac']:.3f}")
PY
done
