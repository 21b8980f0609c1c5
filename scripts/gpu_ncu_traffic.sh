#!/bin/bash
# DRAM traffic and duration of every kernel of one bench run (cold-cache, serialised): feeds profiles/traffic.json
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-tconv-block --decoder ${1:-upsampling}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none --csv --log-file gpurun_out/traffic.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu.log; wc -l gpurun_out/traffic.csv
