#!/bin/bash
# usage (gpurun --gpus 2): the training driver at world size 2, host-fed and device-resident input, then a resume
mkdir -p gpurun_out
D=/tmp/unet_cli_ckpt; rm -rf $D
common="--synthetic 640 --batch_size 64 --epochs 2 --learning_rate 0.0001 --checkpoint_dir $D --progress 0 --summary_every 5"
run() { timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 \
          -m unet_distributed_b200.test_dist $common "$@"; echo "exit $?"; }
run --max_steps 12 > gpurun_out/cli_n2_a.log 2>&1; tail -4 gpurun_out/cli_n2_a.log
run --resume 1 --resident_data 1 > gpurun_out/cli_n2_b.log 2>&1; tail -6 gpurun_out/cli_n2_b.log
ls -la $D $D/unet* | head -12
