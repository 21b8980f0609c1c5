#!/bin/bash
# first-light GPU run: hardware probe + native numerics check
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia-smi.txt 2>&1
(lscpu | head -25; nproc; free -g) > gpurun_out/host.txt 2>&1
cd unet_distributed_b200/csrc
timeout 300 ./build/umma_probe > ../../gpurun_out/probe.log 2>&1; echo "probe exit $?" >> ../../gpurun_out/probe.log
timeout 900 ./build/native_test > ../../gpurun_out/native.log 2>&1; echo "native exit $?" >> ../../gpurun_out/native.log
tail -5 ../../gpurun_out/probe.log; tail -30 ../../gpurun_out/native.log
