#!/bin/bash
# Round-end check (scripts/gpu_final.sh) followed by the ncu launch lists of one training step of both decoders.
bash scripts/gpu_final.sh
for dec in upsampling tconv; do
  bash scripts/gpu_ncu_traffic.sh $dec
  mv gpurun_out/traffic.csv gpurun_out/traffic_$dec.csv
done
ls -la gpurun_out/*.csv
