#!/bin/bash
# quick iteration: native numerics, op + model parity tests, then the bench with a per-launch CSV
mkdir -p gpurun_out; rm -f gpurun_out/prof_*.csv
(cd unet_distributed_b200/csrc && timeout 600 ./build/native_test > ../../gpurun_out/native.log 2>&1; echo "native exit $?" >> ../../gpurun_out/native.log)
grep -E "FAIL|SUMMARY|exit|error" gpurun_out/native.log | head -20
bash scripts/gpu_pytest.sh tests/test_gpu_ops.py tests/test_gpu_model.py tests/test_zz_no_alloc.py
UNETB200_PROFILE_CSV=$PWD/gpurun_out/prof_ups.csv timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; tail -3 gpurun_out/bench.err
UNETB200_PROFILE_CSV=$PWD/gpurun_out/prof_tconv.csv timeout 600 python bench.py --steps 20 --warmup 5 --decoder tconv --no-cpu-baseline > gpurun_out/bench_tconv.json 2> gpurun_out/bench_tconv.err; echo "bench tconv exit $?"; tail -3 gpurun_out/bench_tconv.err
python - <<'PY'
import json
for f in ("gpurun_out/bench.json","gpurun_out/bench_tconv.json"):
    try:
        d=json.load(open(f)); r=d["roofline"]
        print(f, "img/s", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "igemm ms", round(r["ms_per_step"],3), "TF", round(r["achieved"]), "wgrad ms", round(r["wgrad"]["ms_per_step"],3), "launches", d["gpu_launches"], "loss", round(d["loss"],4))
    except Exception as e: print(f, "ERR", e)
PY
