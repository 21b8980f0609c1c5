"""Do the BASELINE configurations whose GPU runs end with dice ~0 (configs[3] with UpSampling2D, configs[4]) collapse in
the reference's own arithmetic too?  The same training loop on the GPU (bf16) and in the fp32 CPU oracle, from the
same weights on the same batches at a reduced batch size, constant dropout off so that the two can be compared step
by step for as long as their trajectories coincide.  Output: profiles/r02_collapse_check.txt (via gpurun_out/)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import unet_oracle as O  # noqa: E402
from unet_distributed_b200.synthetic import make_slices  # noqa: E402
from unet_distributed_b200.trainer import Trainer  # noqa: E402

torch.set_num_threads(os.cpu_count() or 1)
dev = torch.device("cuda:0")
CASES = [
    # name, N, H, C, ups, steps, optimizer settings
    ("configs[3] shape: 240x240x4, UpSampling2D, lr 5e-4 const", 4, 240, 4, True, 40,
     dict(lr0=5e-4, const_lr=True, lr_fraction=0.2, decay_steps=100)),
    ("configs[3] shape: 240x240x4, Conv2DTranspose, lr 5e-4 const", 4, 240, 4, False, 40,
     dict(lr0=5e-4, const_lr=True, lr_fraction=0.2, decay_steps=100)),
    ("configs[4] optimizer: 128x128x1, UpSampling2D, lr 5e-4 decaying (150 steps, x0.25)", 8, 128, 1, True, 60,
     dict(lr0=5e-4, const_lr=False, lr_fraction=0.25, decay_steps=150)),
]
out = []
for name, N, H, C, ups, steps, opt in CASES:
    t0 = time.time()
    weights = O.init_weights(1234, ups, C)
    batches = [make_slices(N, H, H, C, seed=816 + i) for i in range(8)]
    tr = Trainer(N, H, H, C, 1, ups, dropout=0.0, learning_rate=opt["lr0"], const_learningrate=opt["const_lr"],
                 lr_fraction=opt["lr_fraction"], decay_steps=opt["decay_steps"], weights=weights)
    got = []
    for s in range(steps):
        a, b = batches[s % 8]
        got.append(tr.train_step(torch.tensor(a, device=dev), torch.tensor(b, device=dev)).clone())
    got = torch.stack(got).cpu().numpy()
    _, hist = O.train(weights, batches, steps, ups, dropout=0.0, **opt)
    ref = np.array(hist)
    out.append(f"# {name}; batch {N}, {steps} steps, dropout off ({time.time() - t0:.0f} s)")
    out.append("step   oracle loss  GPU loss   oracle dice  GPU dice")
    for s in list(range(0, 8)) + list(range(8, steps, 4)):
        out.append(f"{s:4d}   {ref[s, 0]:10.4f} {got[s, 0]:9.4f}   {ref[s, 1]:10.4f} {got[s, 1]:9.4f}")
    out.append(f"last 8 steps, mean: oracle loss {ref[-8:, 0].mean():.4f} dice {ref[-8:, 1].mean():.4f} | "
               f"GPU loss {got[-8:, 0].mean():.4f} dice {got[-8:, 1].mean():.4f}")
    out.append("")
    tr.close()
text = "\n".join(out)
print(text)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
open(os.path.join(ROOT, "gpurun_out", "collapse_check.txt"), "w").write(text + "\n")
