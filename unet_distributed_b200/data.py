"""data.py of the reference (data.py:5-54): load the four .npy files and cut a shuffled epoch into batches.
Python-3 correct (the reference shuffles a zip object and uses integer `/`, data.py:38,41-42) and seedable."""
import numpy as np

from . import settings_dist
from .preprocess import load_data, update_channels


def load_all_data(out_path=None):
    """data.py:5-26.  Returns imgs_train, msks_train, imgs_test, msks_test (NHWC)."""
    path = settings_dist.OUT_PATH if out_path is None else out_path
    print("-" * 42)
    print("Loading and preprocessing training data...")
    print("-" * 42)
    imgs_train, msks_train = load_data(path, "_train")
    imgs_train, msks_train = update_channels(imgs_train, msks_train, settings_dist.IN_CHANNEL_NO,
                                             settings_dist.OUT_CHANNEL_NO, settings_dist.MODE)
    print("-" * 38)
    print("Loading and preprocessing test data...")
    print("-" * 38)
    imgs_test, msks_test = load_data(path, "_test")
    imgs_test, msks_test = update_channels(imgs_test, msks_test, settings_dist.IN_CHANNEL_NO,
                                           settings_dist.OUT_CHANNEL_NO, settings_dist.MODE)
    print("Training images shape: {}".format(imgs_train.shape))
    print("Training masks shape:  {}".format(msks_train.shape))
    print("Testing images shape:  {}".format(imgs_test.shape))
    print("Testing masks shape:   {}".format(msks_test.shape))
    return imgs_train, msks_train, imgs_test, msks_test


class EpochOfBatches:
    """Result of get_epoch when images and masks differ in channel count (a ragged pair cannot be one ndarray).
    Indexes like the reference's array: epoch[batch_idx, 0] -> images, epoch[batch_idx, 1] -> masks."""

    def __init__(self, data, labels):
        self.data, self.labels = data, labels
        self.shape = (data.shape[0], 2) + tuple(data.shape[1:])

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, idx):
        if isinstance(idx, tuple):
            b, which = idx
            return (self.data, self.labels)[which][b]
        return self.data[idx], self.labels[idx]


def get_epoch(batch_size, imgs_train, msks_train, rng=None):
    """data.py:29-54: drop the remainder, shuffle image/mask pairs together, reshape to
    [num_batches, batch_size, H, W, C] and pair them -> array [num_batches, 2, batch_size, H, W, C].
    Each array keeps its own channel count (the reference reshapes the labels with the image's, data.py:49).
    rng: optional numpy Generator / seed; default = unseeded like the reference."""
    train_size = imgs_train.shape[0]
    epoch_length = train_size - train_size % batch_size
    batch_count = epoch_length // batch_size
    if not isinstance(rng, np.random.Generator):
        rng = np.random.default_rng(rng)
    perm = rng.permutation(train_size)[:epoch_length]
    data = np.asarray(imgs_train)[perm]
    labels = np.asarray(msks_train)[perm]
    data = data.reshape((batch_count, batch_size) + tuple(imgs_train.shape[1:]))
    labels = labels.reshape((batch_count, batch_size) + tuple(msks_train.shape[1:]))
    if data.shape == labels.shape:
        return np.stack([data, labels], axis=1)
    return EpochOfBatches(data, labels)


class ResidentEpochs:
    """The training set kept on one device (8,192 BraTS slices of 128x128 are ~1 GB: a fraction of 180 GB of HBM) and
    cut into the same epochs as get_epoch (data.py:29-54): same permutation for the same seed, same dropped
    remainder, same batch order. `batch(i, start, end)` gathers this replica's shard of batch i
    (test_dist.py:390-394) with one unetb200_gather_batch launch per array straight into fp32 step inputs, so the
    step loop never touches the host.

    imgs / msks: NHWC arrays or tensors. dtype: torch dtype of the resident copy -- float32 (default: cast once at
    upload) or float64 (update_channels hands out float64 arrays, preprocess.py:301-302; the gather kernel then does
    the cast TensorFlow does at feed time). On a CPU device (how the epoch logic is tested without a GPU) the gather
    is torch.index_select; on a CUDA device it is the library's kernel and nothing else."""

    def __init__(self, batch_size, imgs_train, msks_train, device, dtype=None):
        import torch
        self._torch = torch
        self.batch_size = int(batch_size)
        dtype = torch.float32 if dtype is None else dtype
        if dtype not in (torch.float32, torch.float64):
            raise ValueError("ResidentEpochs: dtype must be torch.float32 or torch.float64")
        self.imgs = torch.as_tensor(np.ascontiguousarray(imgs_train)).to(device=device, dtype=dtype).contiguous()
        self.msks = torch.as_tensor(np.ascontiguousarray(msks_train)).to(device=device, dtype=dtype).contiguous()
        self.train_size = int(self.imgs.shape[0])
        self.epoch_length = self.train_size - self.train_size % self.batch_size
        self.num_batches = self.epoch_length // self.batch_size
        self.perm = None

    def __len__(self):
        return self.num_batches

    def shuffle(self, rng=None):
        """New epoch: the permutation get_epoch(batch_size, imgs, msks, rng) would draw."""
        if not isinstance(rng, np.random.Generator):
            rng = np.random.default_rng(rng)
        perm = rng.permutation(self.train_size)[:self.epoch_length]
        self.perm = self._torch.as_tensor(perm, dtype=self._torch.int64).to(self.imgs.device)
        return self

    def _gather(self, src, lo, rows, out):
        torch = self._torch
        shape = (rows,) + tuple(src.shape[1:])
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=src.device)
        elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError(f"ResidentEpochs.batch: out must be a contiguous float32 tensor of shape {shape}")
        if not src.is_cuda:
            out.copy_(src.index_select(0, self.perm[lo:lo + rows]))
            return out
        import ctypes
        from . import _lib
        row_elems = int(np.prod(src.shape[1:]))
        _lib.call("unetb200_gather_batch", ctypes.c_void_p(src.data_ptr()),
                  _lib.F64 if src.dtype == torch.float64 else _lib.F32, ctypes.c_longlong(self.train_size),
                  ctypes.c_void_p(self.perm.data_ptr()), ctypes.c_longlong(lo), int(rows), ctypes.c_longlong(row_elems),
                  ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        return out

    def batch(self, batch_idx, start=0, end=None, out=None):
        """(images, masks) of samples [start, end) of batch `batch_idx` of the current epoch, fp32 on the device.
        out: optional (imgs, msks) pair of preallocated fp32 tensors to fill (e.g. the trainer's input slots)."""
        if self.perm is None:
            raise RuntimeError("ResidentEpochs.shuffle() has not been called")
        end = self.batch_size if end is None else end
        lo = batch_idx * self.batch_size + start
        o_i, o_m = out if out is not None else (None, None)
        return self._gather(self.imgs, lo, end - start, o_i), self._gather(self.msks, lo, end - start, o_m)
