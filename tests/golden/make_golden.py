"""Regenerates tests/golden/unet_small_{tconv,ups}.npz from the fp64 oracle.

The reference ships no fixtures for this path and cannot run here (TensorFlow 1.4 / Python 2.7), so these vectors
come from the CPU restatement (oracle/unet_oracle.py, float64).  They pin the restatement against silent edits and
give the GPU tests a fixed, file-based target.  Run from the repo root:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import unet_oracle as O  # noqa: E402
from unet_distributed_b200.synthetic import make_slices  # noqa: E402

N, H, W = 2, 32, 32
WEIGHT_SEED, DATA_SEED, MASK_SEED = 1234, 816, 2024
SMALL = ("conv1a/kernel", "conv1a/bias", "conv9b/bias", "Mask/kernel", "Mask/bias")


def build(ups):
    imgs, msks = make_slices(N, H, W, 1, DATA_SEED)
    weights = O.init_weights(WEIGHT_SEED, ups)
    rng = np.random.default_rng(MASK_SEED)
    masks = {"conv3a": (rng.uniform(size=(N, H // 4, W // 4, 128)) < 0.8).astype(np.float32),
             "conv4a": (rng.uniform(size=(N, H // 8, W // 8, 256)) < 0.8).astype(np.float32)}
    res = O.loss_and_grads(weights, imgs, msks, ups, masks, 0.2, torch.float64)
    out = {"imgs": imgs, "msks": msks, "mask_conv3a": masks["conv3a"], "mask_conv4a": masks["conv4a"],
           "loss": np.float64(res["loss"]), "dice": np.float64(res["dice"]),
           "sensitivity": np.float64(res["sensitivity"]), "specificity": np.float64(res["specificity"]),
           "preds": res["preds"].astype(np.float32)}
    names, norms, sums = [], [], []
    for k, g in res["grads"].items():
        names.append(k)
        norms.append(np.linalg.norm(g))
        sums.append(g.sum())
        if k in SMALL:
            out["grad:" + k] = g.astype(np.float64)
    out["grad_names"] = np.array(names)
    out["grad_norms"] = np.array(norms, np.float64)
    out["grad_sums"] = np.array(sums, np.float64)
    # one Adam update (TF semantics) of the smallest tensors, decaying LR, step 3
    w = {k: weights[k].astype(np.float64) for k in SMALL}
    m = {k: np.zeros_like(w[k]) for k in SMALL}
    v = {k: np.zeros_like(w[k]) for k in SMALL}
    O.adam_update(w, m, v, {k: res["grads"][k] for k in SMALL}, 3, lr0=0.0005, const_lr=False, lr_fraction=0.25,
                  decay_steps=150)
    for k in SMALL:
        out["adam:" + k] = w[k]
    return out


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    for ups, tag in ((False, "tconv"), (True, "ups")):
        np.savez_compressed(os.path.join(here, f"unet_small_{tag}.npz"), **build(ups))
        print("wrote", tag)
