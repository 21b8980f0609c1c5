"""Checkpoint files: one .npz holding every variable under its TensorFlow name and layout (`conv1a/kernel` HWIO,
`transConv6/kernel` [2,2,Cout,Cin], `Mask/bias`, ...), the Adam slots under TF's slot names (`<var>/Adam`,
`<var>/Adam_1`) and `global_step` -- the contents of the reference's Saver checkpoints (test_dist.py:271, 347-355,
446, 490), so a weight set read from a TF checkpoint with TensorFlow loads here unchanged and vice versa.
The file format itself (TF's tensor bundle) is out of scope."""
from __future__ import annotations

import os

import numpy as np

LAST_GOOD = "last_good_model.npz"      # test_dist.py:446,490 saves `last_good_model.cpkt` after every test pass
AUTOSAVE = "model.autosave.npz"        # tf.train.Supervisor(save_model_secs=60) (test_dist.py:347-355)


def save_checkpoint(path, state):
    """Atomically write a Trainer.state_dict() (name -> ndarray, plus 'global_step')."""
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    tmp = path + ".tmp.npz"
    arrays = {k: np.asarray(v) for k, v in state.items()}
    with open(tmp, "wb") as f:
        np.savez(f, **arrays)
    os.replace(tmp, path)
    return path


def load_checkpoint(path):
    """Inverse of save_checkpoint: dict name -> ndarray, 'global_step' as int."""
    with np.load(path) as z:
        state = {k: z[k] for k in z.files}
    state["global_step"] = int(state.get("global_step", 0))
    return state


def latest(checkpoint_dir):
    """Path of the checkpoint to resume from -- whichever of the last good model and the periodic autosave holds the
    larger global_step (the Supervisor restores the most recent checkpoint of its logdir) -- or None."""
    best, best_step = None, -1
    for name in (LAST_GOOD, AUTOSAVE):
        p = os.path.join(checkpoint_dir, name)
        if not os.path.exists(p):
            continue
        with np.load(p) as z:
            step = int(z["global_step"]) if "global_step" in z.files else 0
        if step > best_step:
            best, best_step = p, step
    return best
