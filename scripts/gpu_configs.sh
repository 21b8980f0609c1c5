#!/bin/bash
# the other BASELINE.json configurations on one GPU (parity cases, not the headline): configs[0], [3], [4]
mkdir -p gpurun_out
run() { name=$1; shift; timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-tconv-block "$@" > gpurun_out/cfg_$name.json 2> gpurun_out/cfg_$name.err; echo "$name exit $?"; python - gpurun_out/cfg_$name.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print("  img/s", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "loss", round(d["loss"],4), d["config"]["workload"][:60])
except Exception as e: print("  ERR", e)
PY
}
run tconv_bs128 --decoder tconv
run ups_240x240x4_bs64 --size 240 --channels 4 --batch 64
run tconv_240x240x4_bs64 --size 240 --channels 4 --batch 64 --decoder tconv
run ups_bs512_decay --batch 512 --const_learningrate 0 --decay_steps 150 --lr_fraction 0.25
