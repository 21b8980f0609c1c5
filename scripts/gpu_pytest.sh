#!/bin/bash
# usage: gpu_pytest.sh [pytest paths/args]   -- runs the GPU tests (all of tests/ by default), writes the parity report
mkdir -p gpurun_out
export UNETB200_TEST_REPORT=$PWD/gpurun_out/parity_report.txt
rm -f $UNETB200_TEST_REPORT
if [ $# -eq 0 ]; then set -- tests; fi
timeout 2400 python -m pytest -q -m gpu --timeout 1200 "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error|error" gpurun_out/pytest_gpu.log | tail -30
