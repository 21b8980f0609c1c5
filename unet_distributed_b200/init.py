"""Host-side weight initialisation with the initialisers the reference names (model.py:47-49 he_uniform for the 3x3
convs; Keras defaults -- glorot_uniform kernels, zero biases -- for Conv2DTranspose and the 1x1 Mask conv).
TensorFlow's RNG stream cannot be reproduced and need not be: weights are generated once on the host from a seeded
NumPy generator and fed to every replica (and to the oracle in parity tests)."""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np


def initial_weights(variables, seed=1234):
    """variables: OrderedDict name -> (offset, shape) as reported by the plan. Returns name -> float32 array."""
    rng = np.random.default_rng(seed)
    out = OrderedDict()
    for name, (_, shape) in variables.items():
        if name.endswith("/bias"):
            out[name] = np.zeros(shape, np.float32)
            continue
        kh, kw = shape[0], shape[1]
        if name.startswith("transConv"):          # [kh, kw, out, in]
            fan_in, fan_out = kh * kw * shape[3], kh * kw * shape[2]
            limit = math.sqrt(6.0 / (fan_in + fan_out))
        elif name.startswith("Mask"):             # [1, 1, in, out], glorot_uniform
            limit = math.sqrt(6.0 / (shape[2] + shape[3]))
        else:                                     # [3, 3, in, out], he_uniform = VarianceScaling(2, fan_in, uniform)
            limit = math.sqrt(6.0 / (kh * kw * shape[2]))
        out[name] = rng.uniform(-limit, limit, size=shape).astype(np.float32)
    return out
