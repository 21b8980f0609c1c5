#!/bin/bash
mkdir -p gpurun_out
UNETB200_HALO_TIMING=1 timeout 600 python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline > gpurun_out/timing.json 2> gpurun_out/timing.err
grep "halo timing" gpurun_out/timing.err | tail -23
