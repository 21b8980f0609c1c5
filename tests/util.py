"""Helpers shared by the parity tests: the oracle on one side, the C-ABI plan on the other, same seeded inputs."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import unet_oracle as O  # noqa: E402

DROPOUT_SEED = 816


def rel_err(got, ref):
    got = np.asarray(got, np.float64)
    ref = np.asarray(ref, np.float64)
    return float(np.linalg.norm(got - ref) / (np.linalg.norm(ref) + 1e-30))


def synthetic_batch(N, H, W, in_channels=1, seed=816):
    from unet_distributed_b200.synthetic import make_slices
    return make_slices(N, H, W, in_channels, seed)


def plan_dropout_masks(plan, seed):
    """The keep masks the conv epilogues will use for Philox seed `seed` (layers conv3a/conv4a = streams 2/3)."""
    import torch
    from unet_distributed_b200 import _lib
    masks = {}
    for name, stream_id, L in (("conv3a", 2, 2), ("conv4a", 3, 3)):
        shape = (plan.N, plan.H >> L, plan.W >> L, 32 << L)
        n = int(np.prod(shape))
        keep = torch.empty(n, dtype=torch.uint8, device=plan.device)
        _lib.call("unetb200_dropout_mask", ctypes.c_void_p(keep.data_ptr()), ctypes.c_longlong(n),
                  ctypes.c_float(1.0 - plan.dropout), ctypes.c_uint64(seed), ctypes.c_uint32(stream_id),
                  ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        masks[name] = keep.cpu().numpy().reshape(shape).astype(np.float32)
    return masks


# plan tensor name -> oracle activation name
def act_names(use_upsampling):
    m = {}
    for L in range(1, 6):
        m[f"a{L}"] = f"conv{L}a"
        m[f"c{L}"] = f"conv{L}b"
    for L in range(6, 10):
        m[f"a{L}"] = f"conv{L}a"
        m[f"c{L}"] = f"conv{L}b"
        if not use_upsampling:
            m[f"u{L}"] = f"transConv{L}"
    return m
