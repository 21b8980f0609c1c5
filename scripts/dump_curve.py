"""Debug helper: 200 training steps on the GPU (no dropout) -> gpurun_out/curve_gpu.npy."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from util import O, synthetic_batch
from unet_distributed_b200.trainer import Trainer
N, H, W, steps = 4, 64, 64, 200
for ups in (False, True):
    weights = O.init_weights(1234, ups)
    batches = [synthetic_batch(N, H, W, seed=816 + i) for i in range(8)]
    tr = Trainer(N, H, W, use_upsampling=ups, dropout=0.0, weights=weights)
    dev = torch.device("cuda:0")
    db = [(torch.tensor(a, device=dev), torch.tensor(b, device=dev)) for a, b in batches]
    out = []
    for s in range(steps):
        out.append(tr.train_step(*db[s % 8]).clone())
    np.save(f"gpurun_out/curve_gpu_ups{int(ups)}.npy", torch.stack(out).cpu().numpy())
print("ok")
