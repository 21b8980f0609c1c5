#!/bin/bash
# One gpurun call, ncu --set full over (a) the bandwidth / reduction kernels and (b) the tensor kernels that had no full
# capture in round 1 (conv_halo_kernel N=128, igemm_kernel at the 8x8 level, wgrad_row_kernel): one eager training
# step each, after the same command has exited 0 without ncu. The reports stay on the box (they exceed the 64 MiB
# return limit); their raw pages (all metrics per launch) and source pages (per-line stalls) come back as gzipped CSV.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline --no-tconv-block --decoder ${DECODER:-upsampling}"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on \
    -k regex:"conv_first32|maxpool_bwd|splitk_reduce|wgrad_halo_reduce|head32|rows_reduce|upsample_|adam_kernel|pack_all" \
    -s ${SMALL_SKIP:-150} -c ${SMALL_COUNT:-42} -o /tmp/r02_small -f $CMD > gpurun_out/ncu_small.log 2>&1
echo "ncu small exit $?"
ncu -i /tmp/r02_small.ncu-rep --page raw --csv | gzip > gpurun_out/r02_small_raw.csv.gz
ncu -i /tmp/r02_small.ncu-rep --page source --csv -k regex:"conv_first32" 2>/dev/null | gzip > gpurun_out/r02_conv_first_source.csv.gz
ncu --set full --clock-control none --import-source on -k regex:"conv_halo_kernel|igemm_kernel|wgrad_row_kernel" \
    -s ${TENSOR_SKIP:-87} -c ${TENSOR_COUNT:-29} -o /tmp/r02_tensor -f $CMD > gpurun_out/ncu_tensor.log 2>&1
echo "ncu tensor exit $?"
ncu -i /tmp/r02_tensor.ncu-rep --page raw --csv | gzip > gpurun_out/r02_tensor_raw.csv.gz
ls -la gpurun_out/*.gz; du -sh gpurun_out
