"""The chief's TensorBoard summaries and progress bar (test_dist.py:275-298, 311-326, 401-404, 436-443, 456-459).

The reference's chief runs a SECOND forward pass per step (`sess.run(summary_op)`, test_dist.py:401-404) just to fill
the summaries.  Here everything they need is already on the device after the training step: the four metrics
(plan.metrics), the step's predictions (plan.preds) and its input batch, so logging costs one small device->host copy
on the steps that log and nothing on the others.  Tags are the reference's: scalars and histograms `loss`, `dice`,
`sensitivity`, `specificity`, scalar `percent_complete`, images `predictions`, `ground_truth`, `images`
(settings_dist.TENSORBOARD_IMAGES of each), and the test-pass scalars `loss_test`, `dice_test`, `sensitivity_test`,
`specificity_test`.  Only rank 0 (the chief) writes."""
from __future__ import annotations

import numpy as np

from . import settings_dist

METRIC_TAGS = ("loss", "dice", "sensitivity", "specificity")


def _to_image(x):
    """NHWC slice -> HWC uint8-scalable float in [0, 1]; channels beyond 1 or 3 keep the first one (tf.summary.image
    accepts 1, 3 or 4 channels; the 4-modality input shows its first modality)."""
    x = np.asarray(x, dtype=np.float32)
    if x.shape[-1] not in (1, 3):
        x = x[..., :1]
    lo, hi = float(x.min()), float(x.max())
    return (x - lo) / (hi - lo) if hi > lo else np.zeros_like(x)


class ChiefSummaries:
    """SummaryWriter wrapper; `enabled=False` (non-chief ranks, --tensorboard=0) turns every call into a no-op."""

    def __init__(self, logdir, total_steps, enabled=True, max_images=settings_dist.TENSORBOARD_IMAGES):
        self.total_steps = max(int(total_steps), 1)
        self.max_images = int(max_images)
        self.writer = None
        if enabled:
            from torch.utils.tensorboard import SummaryWriter
            self.writer = SummaryWriter(log_dir=logdir)

    def percent_complete(self, global_step):
        """test_dist.py:295-298: 100 * global_step / (num_batches * epochs)."""
        return 100.0 * float(global_step) / self.total_steps

    def train_step(self, global_step, metrics, imgs=None, msks=None, preds=None):
        """metrics: the step's [loss, dice, sensitivity, specificity]; imgs/msks/preds: NHWC arrays (any may be None)."""
        if self.writer is None:
            return
        w = self.writer
        for tag, v in zip(METRIC_TAGS, metrics):
            w.add_scalar(tag, float(v), global_step)
            w.add_histogram(tag, np.asarray([float(v)], dtype=np.float32), global_step)
        w.add_scalar("percent_complete", self.percent_complete(global_step), global_step)
        for tag, arr in (("predictions", preds), ("ground_truth", msks), ("images", imgs)):
            if arr is None:
                continue
            for i in range(min(self.max_images, len(arr))):
                w.add_image(f"{tag}/image/{i}", _to_image(arr[i]), global_step, dataformats="HWC")

    def test_pass(self, global_step, m):
        """test_dist.py:322-326, 436-443."""
        if self.writer is None:
            return
        for key in METRIC_TAGS:
            self.writer.add_scalar(key + "_test", float(m[key]), global_step)

    def close(self):
        if self.writer is not None:
            self.writer.flush()
            self.writer.close()
            self.writer = None
