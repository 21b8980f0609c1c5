#!/bin/bash
# What the driver runs at round end, in one call: all GPU tests, smoke(), the default bench line and the reference arm.
mkdir -p gpurun_out
export UNETB200_TEST_REPORT=$PWD/gpurun_out/parity_report.txt
rm -f $UNETB200_TEST_REPORT
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 900 --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -14 gpurun_out/pytest_gpu.log
unset UNETB200_TEST_REPORT
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('__SMOKE_OK__')" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; cut -c1-400 gpurun_out/bench_default.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference exit $?"; cut -c1-300 gpurun_out/bench_reference.json
