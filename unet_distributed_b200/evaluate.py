"""Test-set evaluation with the reference's semantics (test_dist.py:406-446, 461-490; metrics model.py:4-21, 133-143).

The reference evaluates on the chief only, in batches of the GLOBAL batch_size, walks
`range(0, n_test - batch_size, batch_size)` (so the last batch is dropped even when it is full), divides every
accumulated metric by `n_test // batch_size` (which counts that dropped batch), and leaves dropout ON (the Keras
learning phase is forced to 1, model.py:40).  All of that is reproduced here, quirks included, so that numbers
printed by the two programs mean the same thing.

The device side only has to supply the three dice sums of a sub-batch (sum(y*p), sum(y), sum(p)); a test batch larger
than the plan's per-GPU batch is evaluated in sub-batches whose sums are added before the metrics are formed, which is
exactly the metric of the whole batch tensor."""
from __future__ import annotations

import numpy as np


def metrics_from_sums(inter, sum_true, sum_pred, smooth=1.0):
    """(loss, dice, sensitivity, specificity) from I = sum(y*p), Sy = sum(y), Sp = sum(p).
    dice_coef / dice_coef_loss: model.py:4-21; sensitivity / specificity: model.py:133-143 (note that the reference's
    'specificity' divides by sum(p), not by the true negatives)."""
    inter, sum_true, sum_pred = float(inter), float(sum_true), float(sum_pred)
    loss = -np.log(2.0 * inter + smooth) + np.log(sum_true + sum_pred + smooth)
    dice = (2.0 * inter + smooth) / (sum_true + sum_pred + smooth)
    sens = (inter + smooth) / (sum_true + smooth)
    spec = (inter + smooth) / (sum_pred + smooth)
    return float(loss), float(dice), float(sens), float(spec)


def test_batch_starts(n_test, batch_size):
    """Start indices of the evaluated test batches -- test_dist.py:413 `range(0, imgs_test.shape[0] - batch_size,
    batch_size)`."""
    return list(range(0, n_test - batch_size, batch_size))


test_batch_starts.__test__ = False   # (not a pytest test despite the name)


def evaluate(sums_fn, imgs_test, msks_test, batch_size, sub_batch):
    """Average test metrics the way test_dist.py:406-424 forms them.

    sums_fn(imgs, msks) -> (I, Sy, Sp) for `sub_batch` samples (NHWC float32 arrays).  Returns a dict with loss,
    dice, sensitivity, specificity and the number of batches evaluated."""
    n_test = int(imgs_test.shape[0])
    if batch_size % sub_batch != 0:
        raise ValueError("batch_size %d is not a multiple of the per-device batch %d" % (batch_size, sub_batch))
    divisor = n_test // batch_size
    acc = np.zeros(4, np.float64)
    starts = test_batch_starts(n_test, batch_size)
    for idx in starts:
        sums = np.zeros(3, np.float64)
        for off in range(idx, idx + batch_size, sub_batch):
            x = np.ascontiguousarray(imgs_test[off:off + sub_batch], dtype=np.float32)
            y = np.ascontiguousarray(msks_test[off:off + sub_batch], dtype=np.float32)
            sums += np.asarray(sums_fn(x, y), np.float64)
        acc += np.asarray(metrics_from_sums(*sums), np.float64) / max(divisor, 1)
    return {"loss": float(acc[0]), "dice": float(acc[1]), "sensitivity": float(acc[2]), "specificity": float(acc[3]),
            "batches": len(starts), "divisor": divisor}
