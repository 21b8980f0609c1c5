"""2-rank check that a whole training step (NCCL all-reduces included) captures into ONE CUDA graph and replays.
torchrun --nproc-per-node 2 scripts/debug_graph_nccl.py"""
import faulthandler
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
faulthandler.dump_traceback_later(int(os.environ.get("DUMP_AFTER", "50")), exit=True)

import numpy as np
import torch

from unet_distributed_b200 import parallel
from unet_distributed_b200.synthetic import make_slices
from unet_distributed_b200.trainer import Trainer

world, rank, local = parallel.init_from_env("nccl")
dev = torch.device("cuda", local)
N, S = int(os.environ.get("DBG_N", "4")), int(os.environ.get("DBG_S", "64"))
t0 = time.time()


def log(msg):
    print(f"[rank {rank} +{time.time() - t0:5.1f}s] {msg}", flush=True)


tr = Trainer(N, S, S, dropout=0.2, use_cuda_graph=True)
imgs, msks = make_slices(N, S, S, 1, seed=5 + rank)
d_i, d_m = torch.tensor(imgs, device=dev), torch.tensor(msks, device=dev)
log("trainer built")
tr._capture(d_i, d_m)
torch.cuda.synchronize()
log(f"captured: {tr.graph_launches_per_step} launches")
for i in range(5):
    m = tr.train_step(d_i, d_m)
    torch.cuda.synchronize()
    log(f"replay {i}: loss {float(m[0]):.4f}")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.distributed.barrier()
torch.cuda.synchronize()
e0.record()
for i in range(20):
    tr.train_step(d_i, d_m)
e1.record()
torch.cuda.synchronize()
log(f"20 replays: {e0.elapsed_time(e1) / 20:.3f} ms/step")
p = tr.plan.params.clone()
torch.distributed.all_reduce(p, op=torch.distributed.ReduceOp.MAX)
log(f"replicas identical: {bool(torch.equal(p, tr.plan.params))}")
tr.close()
torch.distributed.barrier()
torch.distributed.destroy_process_group()
faulthandler.cancel_dump_traceback_later()
