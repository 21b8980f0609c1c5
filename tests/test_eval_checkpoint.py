"""Host logic of the evaluation pass and the checkpoint files, on the CPU:
the batching / averaging quirks of test_dist.py:406-424, the metric formulas against the oracle, the .npz round trip."""
import os

import numpy as np
import torch

from util import O
from unet_distributed_b200 import checkpoint
from unet_distributed_b200.evaluate import evaluate, metrics_from_sums, test_batch_starts


def test_metrics_from_sums_match_oracle_formulas():
    rng = np.random.default_rng(3)
    y = (rng.random((4, 16, 16, 1)) > 0.8).astype(np.float32)
    p = rng.random((4, 16, 16, 1)).astype(np.float32)
    yt, pt = torch.tensor(y, dtype=torch.float64), torch.tensor(p, dtype=torch.float64)
    got = metrics_from_sums((y.astype(np.float64) * p).sum(), y.sum(dtype=np.float64), p.sum(dtype=np.float64))
    want = (float(O.dice_coef_loss(yt, pt)), float(O.dice_coef(yt, pt)), float(O.sensitivity(yt, pt)),
            float(O.specificity(yt, pt)))
    np.testing.assert_allclose(got, want, rtol=1e-12)
    # empty mask, empty prediction: smooth = 1 keeps everything finite (model.py:4-21)
    assert metrics_from_sums(0.0, 0.0, 0.0) == (0.0, 1.0, 1.0, 1.0)


def test_batch_walk_is_the_references():
    # test_dist.py:413: range(0, n - B, B) -- the last batch is dropped even when it is full
    assert test_batch_starts(512, 128) == [0, 128, 256]
    assert test_batch_starts(600, 128) == [0, 128, 256, 384]
    assert test_batch_starts(128, 128) == []


def test_evaluate_accumulates_sub_batches_and_keeps_the_divisor_quirk():
    rng = np.random.default_rng(5)
    n, B, sub = 40, 8, 4
    imgs = rng.standard_normal((n, 4, 4, 1)).astype(np.float32)
    msks = (rng.random((n, 4, 4, 1)) > 0.7).astype(np.float32)
    calls = []

    def sums_fn(x, y):   # a stand-in "network": p = sigmoid(x)
        assert x.shape[0] == sub and y.shape[0] == sub and x.dtype == np.float32
        calls.append(x.shape[0])
        p = 1.0 / (1.0 + np.exp(-x.astype(np.float64)))
        return (y * p).sum(), y.sum(), p.sum()

    got = evaluate(sums_fn, imgs, msks, B, sub)
    starts = list(range(0, n - B, B))           # 0, 8, 16, 24 -- the batch at 32 is never evaluated
    assert got["batches"] == len(starts) == 4 and got["divisor"] == n // B == 5
    assert len(calls) == len(starts) * (B // sub)
    want = np.zeros(4)
    for i in starts:   # metrics of the WHOLE batch tensor, averaged with the reference's divisor
        x, y = imgs[i:i + B].astype(np.float64), msks[i:i + B].astype(np.float64)
        p = 1.0 / (1.0 + np.exp(-x))
        want += np.array(metrics_from_sums((y * p).sum(), y.sum(), p.sum())) / (n // B)
    np.testing.assert_allclose([got["loss"], got["dice"], got["sensitivity"], got["specificity"]], want, rtol=1e-10)


def test_checkpoint_round_trip_uses_tf_names(tmp_path):
    w = O.init_weights(11)
    state = {"global_step": 37}
    for k, v in w.items():
        state[k] = v
        state[k + "/Adam"] = np.full_like(v, 0.25)
        state[k + "/Adam_1"] = np.full_like(v, 0.5)
    path = os.path.join(tmp_path, "ckpt", checkpoint.LAST_GOOD)
    assert checkpoint.latest(os.path.dirname(path)) is None
    checkpoint.save_checkpoint(path, state)
    assert checkpoint.latest(os.path.dirname(path)) == path
    back = checkpoint.load_checkpoint(path)
    assert back["global_step"] == 37 and isinstance(back["global_step"], int)
    assert "conv1a/kernel" in back and "transConv6/kernel/Adam_1" in back and "Mask/bias/Adam" in back
    assert back["conv1a/kernel"].shape == (3, 3, 1, 32) and back["transConv6/kernel"].shape == (2, 2, 256, 512)
    for k in state:
        np.testing.assert_array_equal(np.asarray(state[k]), back[k])
    assert not [f for f in os.listdir(os.path.dirname(path)) if f.endswith(".tmp.npz")]


def test_latest_prefers_the_newer_of_last_good_and_autosave(tmp_path):
    d = str(tmp_path)
    assert checkpoint.latest(d) is None
    checkpoint.save_checkpoint(os.path.join(d, checkpoint.LAST_GOOD), {"global_step": 10, "conv1a/bias": np.zeros(32)})
    assert checkpoint.latest(d).endswith(checkpoint.LAST_GOOD)
    checkpoint.save_checkpoint(os.path.join(d, checkpoint.AUTOSAVE), {"global_step": 14, "conv1a/bias": np.ones(32)})
    assert checkpoint.latest(d).endswith(checkpoint.AUTOSAVE)
    checkpoint.save_checkpoint(os.path.join(d, checkpoint.LAST_GOOD), {"global_step": 20, "conv1a/bias": np.zeros(32)})
    assert checkpoint.latest(d).endswith(checkpoint.LAST_GOOD)


def test_epoch_seed_is_a_function_of_the_global_step():
    """A resumed run must re-draw the permutation of the epoch it stopped in: the seed depends only on
    which epoch the restored global_step lies in."""
    from unet_distributed_b200.test_dist import epoch_seed
    nb = 7
    assert [epoch_seed(s, nb) for s in (0, 3, 6)] == [1234] * 3
    assert [epoch_seed(s, nb) for s in (7, 13)] == [1241] * 2 and epoch_seed(14, nb) == 1248


def test_chief_summaries_write_the_references_tags(tmp_path):
    """test_dist.py:275-298, 322-326: scalar + histogram per metric, percent_complete, three images of each kind, and
    the *_test scalars; a disabled writer (non-chief ranks) writes nothing."""
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    from unet_distributed_b200.summaries import ChiefSummaries
    rng = np.random.default_rng(0)
    imgs, msks = rng.standard_normal((5, 16, 16, 4)).astype(np.float32), (rng.random((5, 16, 16, 1)) > 0.5) * 1.0
    preds = rng.random((5, 16, 16, 1)).astype(np.float32)
    s = ChiefSummaries(str(tmp_path / "run"), total_steps=40)
    assert s.percent_complete(10) == 25.0
    for step in (1, 2):
        s.train_step(step, [2.5 / step, 0.1 * step, 0.2, 0.3], imgs, msks, preds)
    s.test_pass(2, {"loss": 1.5, "dice": 0.4, "sensitivity": 0.5, "specificity": 0.6})
    s.close()
    acc = EventAccumulator(str(tmp_path / "run"), size_guidance={"images": 0, "scalars": 0, "histograms": 0})
    acc.Reload()
    tags = acc.Tags()
    for t in ("loss", "dice", "sensitivity", "specificity", "percent_complete", "loss_test", "dice_test",
              "sensitivity_test", "specificity_test"):
        assert t in tags["scalars"], t
    for t in ("loss", "dice", "sensitivity", "specificity"):
        assert t in tags["histograms"], t
    for kind in ("predictions", "ground_truth", "images"):
        assert sorted(t for t in tags["images"] if t.startswith(kind + "/")) == [f"{kind}/image/{i}" for i in range(3)]
    assert [e.value for e in acc.Scalars("loss")] == [2.5, 1.25]
    assert [e.value for e in acc.Scalars("percent_complete")] == [2.5, 5.0]
    off = ChiefSummaries(str(tmp_path / "none"), total_steps=40, enabled=False)
    off.train_step(1, [1, 1, 1, 1], imgs, msks, preds)
    off.close()
    assert not os.path.exists(str(tmp_path / "none"))
