"""BraTS-shaped synthetic slices (no dataset can be fetched here).  Statistics follow what preprocess.py produces:
z-scored modalities inside an elliptical "brain" on a constant background (preprocess.py:117-124), one-hot label
channels whose sum is a binary lesion mask (preprocess.py:305-307), empty masks on ~40 % of the slices (those are
why dice has smooth=1).  The default seed 816 is the one in the reference's commented initialiser (model.py:49)."""
import os

import numpy as np


def make_slices(n, H=128, W=128, in_channels=1, seed=816):
    """-> (imgs float32 [n,H,W,in_channels], msks float32 [n,H,W,1] in {0,1})."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    imgs = np.full((n, H, W, in_channels), -0.5, np.float32)
    msks = np.zeros((n, H, W, 1), np.float32)
    s = H / 128.0
    for i in range(n):
        cy, cx = H / 2 + rng.uniform(-6, 6) * s, W / 2 + rng.uniform(-6, 6) * s
        ay, ax = rng.uniform(40, 56) * s, rng.uniform(40, 56) * s
        brain = ((yy - cy) / ay) ** 2 + ((xx - cx) / ax) ** 2 <= 1.0
        for c in range(in_channels):
            tex = rng.standard_normal((H, W)).astype(np.float32)
            imgs[i, :, :, c] = np.where(brain, tex, -0.5)
        if rng.uniform() < 0.6:
            ly, lx = rng.uniform(4, 16) * s, rng.uniform(4, 16) * s
            oy, ox = cy + rng.uniform(-0.5, 0.5) * ay, cx + rng.uniform(-0.5, 0.5) * ax
            lesion = (((yy - oy) / ly) ** 2 + ((xx - ox) / lx) ** 2 <= 1.0) & brain
            msks[i, :, :, 0] = lesion
            for c in range(in_channels):      # lesions are brighter, so the task is learnable
                imgs[i, :, :, c] += 1.5 * lesion
    return imgs, msks


def write_npy_dataset(out_path, n_train=256, n_test=64, H=128, W=128, seed=816):
    """Write imgs_train/msks_train/imgs_test/msks_test .npy in the reference's on-disk format (README.md:40-45):
    4 modality channels and 4 one-hot label channels, NHWC."""
    os.makedirs(out_path, exist_ok=True)
    for prefix, n, sd in (("_train", n_train, seed), ("_test", n_test, seed + 1)):
        imgs, msks = make_slices(n, H, W, 4, sd)
        rng = np.random.default_rng(sd + 7)
        label = rng.integers(0, 4, size=msks.shape[:3])
        onehot = np.zeros(msks.shape[:3] + (4,), np.float32)
        for k in range(4):
            onehot[..., k] = (label == k) * msks[..., 0]
        np.save(os.path.join(out_path, "imgs" + prefix + ".npy"), imgs)
        np.save(os.path.join(out_path, "msks" + prefix + ".npy"), onehot)
