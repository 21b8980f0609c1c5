"""The two preprocess.py entry points the training path uses (preprocess.py:281-350): loading the four .npy files
and selecting modality / merging label channels.  The offline BraTS .mha conversion is out of scope (SimpleITK)."""
import os

import numpy as np


def load_data(data_path, prefix="_train"):
    """preprocess.py:281-285 -- memory-mapped imgs<prefix>.npy / msks<prefix>.npy, NHWC."""
    imgs = np.load(os.path.join(data_path, "imgs" + prefix + ".npy"), mmap_mode="r", allow_pickle=False)
    msks = np.load(os.path.join(data_path, "msks" + prefix + ".npy"), mmap_mode="r", allow_pickle=False)
    return imgs, msks


def update_channels(imgs, msks, input_no=3, output_no=3, mode=1, CHANNEL_LAST=True):
    """preprocess.py:287-350.  mode 1: image = channel 2 (FLAIR), mask = sum of the four label channels;
    mode 2: image = channel 0 (T1 post), mask = label 3; mode 3: image = channel 1 (T2), mask = labels 0+2+3;
    any other mode leaves the image zero and merges all labels.  Returns float64 arrays like the reference
    (np.zeros default dtype, preprocess.py:301-302)."""
    imgs = np.asarray(imgs).astype("float32")
    msks = np.asarray(msks).astype("float32")
    ax = 3 if CHANNEL_LAST else 1

    def ch(a, i):
        return np.take(a, i, axis=ax)

    shp = list(imgs.shape)
    ishape, mshape = list(shp), list(shp)
    ishape[ax], mshape[ax] = input_no, output_no
    new_imgs, new_msks = np.zeros(ishape), np.zeros(mshape)
    sel = [slice(None)] * 4
    sel[ax] = 0
    sel = tuple(sel)
    if mode == 1:
        new_imgs[sel] = ch(imgs, 2)
        new_msks[sel] = ch(msks, 0) + ch(msks, 1) + ch(msks, 2) + ch(msks, 3)
    elif mode == 2:
        new_imgs[sel] = ch(imgs, 0)
        new_msks[sel] = ch(msks, 3)
    elif mode == 3:
        new_imgs[sel] = ch(imgs, 1)
        new_msks[sel] = ch(msks, 0) + ch(msks, 2) + ch(msks, 3)
    else:
        new_msks[sel] = ch(msks, 0) + ch(msks, 1) + ch(msks, 2) + ch(msks, 3)
    return new_imgs, new_msks
