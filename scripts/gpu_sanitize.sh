#!/bin/bash
# usage: gpu_sanitize.sh memcheck|racecheck|synccheck|initcheck   (ONE tool per gpurun call, see B200_PROFILING.md)
# compute-sanitizer over the per-operator parity tests (every tile engine, both precisions) and two small whole-model
# cases. The summary lands in gpurun_out/sanitize_<tool>.txt; copy it to profiles/ when it is clean.
TOOL=${1:-memcheck}
mkdir -p gpurun_out
LOG=gpurun_out/sanitize_$TOOL.log
OPS='test_conv3x3_fwd_dgrad_wgrad or test_tconv2x2 or test_pool_and_upsample or test_first_conv_and_head or test_adam_and_sgd_kernels or test_conv3x3_tf32 or test_tconv2x2_tf32 or test_bandwidth_kernels_fp32'
MODEL='test_tconv_dropout or test_upsampling_dropout or test_model_tf32_tconv'
# plain run first: the sanitizer only runs on a command line that has just exited 0 without it
timeout 600 python -m pytest -q -m gpu -x tests/test_gpu_ops.py tests/test_gpu_tf32.py tests/test_gpu_model.py \
    -k "$OPS or $MODEL" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
timeout 2400 compute-sanitizer --tool $TOOL --error-exitcode 1 --print-limit 30 \
    python -m pytest -q -m gpu -x --timeout 2000 tests/test_gpu_ops.py tests/test_gpu_tf32.py tests/test_gpu_model.py \
    -k "$OPS or $MODEL" > $LOG 2>&1
RC=$?
{
  echo "compute-sanitizer --tool $TOOL over: $OPS or $MODEL"
  echo "exit code $RC"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" $LOG | tail -5
  grep -E "=========     (Invalid|Uninitialized|Race|Barrier|Error|Warning)" $LOG | sort | uniq -c | sort -rn | head -20
} > gpurun_out/sanitize_$TOOL.txt
cat gpurun_out/sanitize_$TOOL.txt
