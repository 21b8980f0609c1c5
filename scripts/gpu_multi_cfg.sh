#!/bin/bash
# BASELINE configs[3] and [4] on N GPUs through torchrun (bench.py with the config's shape)
N=${1:-8}
mkdir -p gpurun_out
run() { name=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/mcfg_${name}_n$N.json 2> gpurun_out/mcfg_${name}_n$N.err; echo "$name n=$N exit $?"; cut -c1-260 gpurun_out/mcfg_${name}_n$N.json; }
run 240x240x4_bs64 --size 240 --channels 4 --batch 64
run bs512_decay --batch 512 --const_learningrate 0 --decay_steps 150 --lr_fraction 0.25
