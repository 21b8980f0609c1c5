#!/bin/bash
# GPU parity tests (all of them, no -x: every failure is wanted), then a short bench of both decoders (no ncu)
mkdir -p gpurun_out
export UNETB200_TEST_REPORT=$PWD/gpurun_out/parity_report.txt
rm -f $UNETB200_TEST_REPORT
timeout 2400 python -m pytest tests -q -m gpu --timeout 1200 --durations=15 "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
UNETB200_PROFILE_CSV=$PWD/gpurun_out/prof_ups.csv timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
UNETB200_PROFILE_CSV=$PWD/gpurun_out/prof_tconv.csv timeout 600 python bench.py --steps 20 --warmup 5 --decoder tconv --no-cpu-baseline > gpurun_out/bench_tconv.json 2> gpurun_out/bench_tconv.err; echo "bench tconv exit $?"
cat gpurun_out/bench_tconv.json; tail -5 gpurun_out/bench_tconv.err
