#!/bin/bash
# ncu --set full over the kernels the Conv2DTranspose decoder adds (igemm_kernel with the depth-to-space epilogue, the
# transposed-conv dgrad, wgrad_kernel): one eager training step, after the same command has exited 0 without ncu.
# Reports stay on the box; raw pages (all launches) and the igemm source page come back as gzipped CSV.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline --no-tconv-block --decoder tconv"
$CMD > gpurun_out/plain_tconv.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_tconv.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"igemm_kernel|wgrad_kernel" \
    -s ${SKIP:-69} -c ${COUNT:-23} -o /tmp/r02_tconv -f $CMD > gpurun_out/ncu_tconv.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/r02_tconv.ncu-rep --page raw --csv | gzip > gpurun_out/r02_tconv_raw.csv.gz
ncu -i /tmp/r02_tconv.ncu-rep --page source --csv -k regex:"igemm_kernel" 2>/dev/null | gzip > gpurun_out/r02_tconv_igemm_source.csv.gz
ncu -i /tmp/r02_tconv.ncu-rep --page source --csv -k regex:"wgrad_kernel" 2>/dev/null | gzip > gpurun_out/r02_tconv_wgrad_source.csv.gz
ls -la gpurun_out/*.gz; du -sh gpurun_out
