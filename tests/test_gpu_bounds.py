"""Out-of-bounds WRITE detection without compute-sanitizer (closed on this GPU pool): the plan's arena is created with a
guard gap after every sub-buffer (UNETB200_ARENA_GUARD), filled with a byte pattern, bracketed by guard tensors, and a
full training step (forward, backward with the side stream, optimizer, re-pack) must leave every byte outside the
plan's regions -- and the guard tensors around the parameter / gradient / Adam buffers -- exactly as they were.
Covers both decoders, both precisions, ragged 240x240 tiles and the shapes whose tiles straddle images."""
import ctypes
import os

import numpy as np
import pytest
import torch

from util import synthetic_batch

pytestmark = pytest.mark.gpu

GUARD = 8192
PATTERN = 0x5A


@pytest.mark.parametrize("N,H,C,ups,precision", [
    (2, 64, 1, False, "bf16"), (3, 32, 1, True, "bf16"), (1, 240, 4, False, "bf16"), (1, 240, 4, True, "bf16"),
    (5, 16, 1, False, "bf16"), (2, 64, 1, False, "tf32"), (1, 48, 4, True, "tf32"), (8, 128, 1, True, "bf16"),
])
def test_a_training_step_writes_only_inside_its_buffers(cuda_device, N, H, C, ups, precision):
    from unet_distributed_b200 import _lib
    from unet_distributed_b200.trainer import Trainer
    os.environ["UNETB200_ARENA_GUARD"] = str(GUARD)
    try:
        tr = Trainer(N, H, H, C, 1, ups, dropout=0.2, precision=precision)
    finally:
        del os.environ["UNETB200_ARENA_GUARD"]
    p = tr.plan
    L = _lib.lib()
    regions = []
    off, nb = ctypes.c_longlong(), ctypes.c_longlong()
    for i in range(L.unetb200_plan_num_regions(p._h)):
        _lib.call("unetb200_plan_region", p._h, i, ctypes.byref(off), ctypes.byref(nb))
        regions.append((off.value, nb.value))
    assert regions and all(b > 0 for _, b in regions)
    inside = torch.zeros(p.arena_bytes, dtype=torch.bool, device=cuda_device)
    end = 0
    for o, b in regions:
        assert o >= end, "regions overlap"
        inside[o:o + b] = True
        end = o + b
    assert int((~inside).sum()) >= GUARD * len(regions)
    # pattern everywhere outside the regions; the regions keep their contents (packed weights, tables)
    p.arena[~inside] = PATTERN
    # flat buffers: re-home them between guard tensors
    n = p.num_params_padded
    big = torch.full((4 * (n + 2 * 1024) + 1024,), float("nan"), dtype=torch.float32, device=cuda_device)
    views = []
    for k, src in enumerate((p.params, p.grads, tr.m, tr.v)):
        base = k * (n + 2 * 1024) + 1024
        v = big[base:base + n]
        v.copy_(src)
        views.append(v)
    p.bind(views[0], views[1])
    tr.m, tr.v = views[2], views[3]
    tr.reducer.flat = views[1]
    guard_mask = torch.ones_like(big, dtype=torch.bool)
    for k in range(4):
        base = k * (n + 2 * 1024) + 1024
        guard_mask[base:base + n] = False
    imgs, msks = synthetic_batch(N, H, H, C)
    d_i, d_m = torch.tensor(imgs, device=cuda_device), torch.tensor(msks, device=cuda_device)
    before = p.params.clone()
    for _ in range(2):
        met = tr.train_step(d_i, d_m)
    tr.predict(d_i, d_m)
    torch.cuda.synchronize()
    assert torch.isfinite(met).all() and not torch.equal(before, p.params)
    bad = (p.arena[~inside] != PATTERN)
    assert not bool(bad.any()), f"{int(bad.sum())} bytes written outside the plan's regions"
    assert bool(torch.isnan(big[guard_mask]).all()), "a flat-buffer guard was overwritten"
    assert bool(torch.isfinite(big[~guard_mask]).all())
