#!/bin/bash
# N-GPU bench through torchrun exactly as the driver launches it
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench n=$N exit $?"; cat gpurun_out/bench_n$N.json | cut -c1-700; tail -5 gpurun_out/bench_n$N.err
