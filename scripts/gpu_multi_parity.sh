#!/bin/bash
# usage (under `gpurun --gpus 2`): scripts/gpu_multi_parity.sh  -- the 2-rank NCCL parity test; log -> gpurun_out/
mkdir -p gpurun_out
export UNETB200_TEST_REPORT=$PWD/gpurun_out/multi_parity_report.txt
rm -f $UNETB200_TEST_REPORT
timeout 420 python -m pytest -q -m gpu --timeout 150 tests/test_gpu_multi.py -rs > gpurun_out/pytest_multi.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_multi.log
tail -15 gpurun_out/pytest_multi.log
cat $UNETB200_TEST_REPORT 2>/dev/null
