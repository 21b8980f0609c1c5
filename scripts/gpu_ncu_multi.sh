#!/bin/bash
# full ncu capture of a few launches of several kernels in one bench run (keeps each report small):
# usage gpu_ncu_multi.sh "<regex> <skip> <count> <outname>" ...
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline --decoder ${DECODER:-upsampling}"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
for spec in "$@"; do
  set -- $spec
  ncu --set full --clock-control none --import-source on -k regex:$1 -s $2 -c $3 -o gpurun_out/$4 -f $CMD > gpurun_out/ncu_$4.log 2>&1
  echo "ncu $4 exit $?"; ls -la gpurun_out/$4.ncu-rep
done
