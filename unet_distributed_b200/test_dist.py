"""Training driver with the reference's flag surface (test_dist.py:64-85; README.md:61-72 spellings accepted too).

    python -m unet_distributed_b200.test_dist --batch_size 128 --epochs 1 --use_upsampling=False
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m unet_distributed_b200.test_dist --batch_size 1024

`--batch_size` is the GLOBAL batch (test_dist.py:390): every rank trains on batch_size / world_size samples of each
batch.  The parameter server and the SavedModel export of the reference are out of scope; this is the loop of
test_dist.py:380-398, 448-459 with the chief's TensorBoard summaries and tqdm bar (summaries.py), the chief's test pass
after every epoch and at the end (test_dist.py:406-446, 461-490), `last_good_model` checkpoints under TF variable
names plus the Supervisor's periodic autosave (.npz, see checkpoint.py), and --resume, which continues at the
restored global_step."""
from __future__ import annotations

import argparse
import os
import time

import numpy as np
import torch

from . import checkpoint, parallel, settings_dist
from .data import ResidentEpochs, get_epoch, load_all_data
from .summaries import ChiefSummaries
from .synthetic import make_slices
from .trainer import Trainer


def _bool(v):
    if isinstance(v, bool):
        return v
    return str(v).lower() in ("1", "true", "t", "yes", "y")


def build_parser():
    p = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument("--const_learningrate", type=_bool, default=settings_dist.CONST_LEARNINGRATE)
    p.add_argument("--learning_rate", "--learningrate", type=float, default=settings_dist.LEARNINGRATE,
                   dest="learning_rate")
    p.add_argument("--lr_fraction", type=float, default=settings_dist.LR_FRACTION)
    p.add_argument("--decay_steps", type=int, default=settings_dist.DECAY_STEPS)
    p.add_argument("--is_sync", type=int, default=1, help="only synchronous updates are implemented")
    p.add_argument("--ip", type=str, default="127.0.0.1", help="accepted for compatibility; ranks come from torchrun")
    p.add_argument("--batch_size", type=int, default=settings_dist.BATCH_SIZE)
    p.add_argument("--epochs", type=int, default=settings_dist.EPOCHS)
    p.add_argument("--use_upsampling", type=_bool, default=settings_dist.USE_UPSAMPLING)
    p.add_argument("--num_threads", type=int, default=settings_dist.NUM_INTRA_THREADS, help="ignored (CPU knob)")
    p.add_argument("--synthetic", type=int, default=0,
                   help="train on this many synthetic BraTS-shaped slices instead of the .npy files")
    p.add_argument("--data_path", type=str, default=None, help="directory with imgs_/msks_{train,test}.npy")
    p.add_argument("--max_steps", type=int, default=0, help="stop early after this many steps (0 = all)")
    p.add_argument("--optimizer", type=str, default="adam", choices=["adam", "sgd"])
    p.add_argument("--precision", type=str, default="bf16", choices=["bf16", "tf32", "fp32"],
                   help="bf16 (default) or the fp32/tf32 mode: fp32 activations, tf32 tensor-core MMAs")
    p.add_argument("--resident_data", type=_bool, default=False,
                   help="keep the training set on the GPU and gather batches there (no host copies in the step loop)")
    p.add_argument("--checkpoint_dir", type=str, default=settings_dist.CHECKPOINT_DIRECTORY)
    p.add_argument("--resume", type=_bool, default=False, help="start from <checkpoint_dir>/last_good_model.npz")
    p.add_argument("--test", type=_bool, default=True, help="run the chief's test pass after each epoch and at the end")
    p.add_argument("--tensorboard", type=_bool, default=True,
                   help="chief writes TensorBoard scalars / histograms / images under <checkpoint_dir>/<run name>")
    p.add_argument("--summary_every", type=int, default=1,
                   help="log the chief's summaries every this many steps (the reference logs every step)")
    p.add_argument("--save_model_secs", type=float, default=60.0,
                   help="periodic checkpoint interval of the chief, as tf.train.Supervisor(save_model_secs=60); 0 = off")
    p.add_argument("--progress", type=_bool, default=True, help="tqdm progress bar with (loss, dice) on the chief")
    return p


def run_name(FLAGS):
    """Sub-directory of the checkpoint directory, named like the reference's (test_dist.py:342-345)."""
    return "unet,lr={},{},b200".format(FLAGS.learning_rate, FLAGS.optimizer)


def epoch_seed(step, num_batches):
    """Shuffle seed of the epoch that contains global step `step` (1234 for the first, then 1234 + its first step):
    a resumed run re-draws the permutation of the epoch it stopped in."""
    return 1234 + (step // num_batches) * num_batches


def main(argv=None):
    FLAGS = build_parser().parse_args(argv)
    if not FLAGS.is_sync:
        raise SystemExit("asynchronous parameter-server updates (--is_sync=0) are not implemented")
    world, rank, _ = parallel.init_from_env()
    chief = rank == 0
    if FLAGS.synthetic:
        imgs_train, msks_train = make_slices(FLAGS.synthetic, settings_dist.IMG_ROWS, settings_dist.IMG_COLS,
                                             settings_dist.IN_CHANNEL_NO)
        n_test = max(FLAGS.synthetic // 8, 2 * FLAGS.batch_size)
        imgs_test, msks_test = make_slices(n_test, settings_dist.IMG_ROWS, settings_dist.IMG_COLS,
                                           settings_dist.IN_CHANNEL_NO, seed=4242)
    else:
        imgs_train, msks_train, imgs_test, msks_test = load_all_data(FLAGS.data_path)
    batch_size = FLAGS.batch_size
    start, end = parallel.shard_range(batch_size, world, rank)           # test_dist.py:390-392
    _, H, W, C = imgs_train.shape
    trainer = Trainer(end - start, H, W, C, msks_train.shape[3], FLAGS.use_upsampling,
                      learning_rate=FLAGS.learning_rate, const_learningrate=FLAGS.const_learningrate,
                      lr_fraction=FLAGS.lr_fraction, decay_steps=FLAGS.decay_steps, optimizer=FLAGS.optimizer,
                      precision=FLAGS.precision)
    if FLAGS.resume:
        path = checkpoint.latest(FLAGS.checkpoint_dir)
        if path is None:
            raise SystemExit("--resume: no checkpoint under %s" % FLAGS.checkpoint_dir)
        trainer.load_state_dict(checkpoint.load_checkpoint(path))       # every rank reads the same file
        if chief:
            print("resumed from {} at global_step {}".format(path, trainer.global_step()), flush=True)

    train_size = imgs_train.shape[0]
    num_batches = (train_size - train_size % batch_size) // batch_size
    total = num_batches * FLAGS.epochs
    if FLAGS.max_steps:
        total = min(total, FLAGS.max_steps)
    # the loop is bounded on the RESTORED global_step (`while step < num_batches*epochs`, test_dist.py:380-398):
    # a resumed run only does the steps that are left, in the epoch permutation it stopped in
    step = trainer.global_step()
    summaries = ChiefSummaries(os.path.join(FLAGS.checkpoint_dir, run_name(FLAGS)), num_batches * FLAGS.epochs,
                               enabled=chief and FLAGS.tensorboard)

    def test_pass(epoch_no):
        """Chief only (test_dist.py:401-446): metrics over the test set, then the last-good-model checkpoint."""
        if not chief:
            return
        if FLAGS.test and imgs_test.shape[0] > batch_size:
            m = trainer.evaluate(imgs_test, msks_test, batch_size)
            print("\nEpoch {} of {}: TEST DATASET\nloss = {:.4f}\nDice = {:.4f}\nSensitivity = {:.4f}\n"
                  "Specificity = {:.4f}".format(epoch_no, FLAGS.epochs, m["loss"], m["dice"], m["sensitivity"],
                                                m["specificity"]), flush=True)
            summaries.test_pass(step, m)
        checkpoint.save_checkpoint(os.path.join(FLAGS.checkpoint_dir, checkpoint.LAST_GOOD), trainer.state_dict())

    resident = epoch = None
    if FLAGS.resident_data:   # same epochs as get_epoch, cut on the device by unetb200_gather_batch
        resident = ResidentEpochs(batch_size, imgs_train, msks_train, trainer.plan.device)
        resident.shuffle(epoch_seed(step, num_batches))
        assert len(resident) == num_batches
    else:
        epoch = get_epoch(batch_size, imgs_train, msks_train, rng=epoch_seed(step, num_batches))  # same on every rank
        assert len(epoch) == num_batches
    progressbar = None
    if chief and FLAGS.progress:
        from tqdm import tqdm
        progressbar = tqdm(total=total, initial=min(step, total), desc="(loss=?, dice=?)")   # test_dist.py:369,456-459
    t0, step0, last_save = time.time(), step, time.time()
    while step < total:
        batch_idx = step % num_batches
        summarize = summaries.writer is not None and (step + 1) % max(FLAGS.summary_every, 1) == 0
        report = summarize or (step + 1) % 10 == 0 or step + 1 == total
        if resident is not None:
            d_data, d_labels = resident.batch(batch_idx, start, end)
            met = trainer.train_step(d_data, d_labels)
            met = met.cpu().numpy() if report else None
        else:
            data = np.ascontiguousarray(epoch[batch_idx, 0][start:end], dtype=np.float32)
            labels = np.ascontiguousarray(epoch[batch_idx, 1][start:end], dtype=np.float32)
            # steps whose numbers are shown synchronise; the others are pipelined (the copy of the next batch
            # overlaps the kernels of this one, trainer.train_step_host)
            met = trainer.train_step_host(data, labels, wait=report)
        step += 1
        if summarize:   # chief only: this step's metrics, predictions and inputs -- no second forward pass
            k = settings_dist.TENSORBOARD_IMAGES
            if resident is not None:
                s_imgs, s_msks = d_data[:k].cpu().numpy(), d_labels[:k].cpu().numpy()
            else:
                s_imgs, s_msks = data[:k], labels[:k]
            summaries.train_step(step, met, s_imgs, s_msks, trainer.plan.preds[:k].cpu().numpy())
        if chief and report:
            loss_v, dice_v = float(met[0]), float(met[1])
            if progressbar is not None:
                progressbar.set_description("(loss={:.4f}, dice={:.4f})".format(loss_v, dice_v))
                progressbar.update(step - progressbar.n)
            else:
                dt = time.time() - t0
                print("step {}/{} (loss={:.4f}, dice={:.4f}) {:.1f} images/sec".format(
                    step, total, loss_v, dice_v, (step - step0) * batch_size / dt), flush=True)
        if chief and FLAGS.save_model_secs > 0 and time.time() - last_save >= FLAGS.save_model_secs:
            # Supervisor(save_model_secs=60) autosave (test_dist.py:347-355); --resume prefers the newer file
            checkpoint.save_checkpoint(os.path.join(FLAGS.checkpoint_dir, checkpoint.AUTOSAVE), trainer.state_dict())
            last_save = time.time()
        if batch_idx == num_batches - 1 and step < total:
            test_pass(step // num_batches)                                           # test_dist.py:406-446
            if resident is not None:
                resident.shuffle(epoch_seed(step, num_batches))
            else:
                epoch = get_epoch(batch_size, imgs_train, msks_train,
                                  rng=epoch_seed(step, num_batches))                 # test_dist.py:448-452
    if progressbar is not None:
        progressbar.close()
    if chief and step > step0:
        print("{} steps, {:.1f} images/sec".format(step - step0, (step - step0) * batch_size / (time.time() - t0)),
              flush=True)
    test_pass(max(step // max(num_batches, 1), 1))                                   # test_dist.py:461-490
    summaries.close()
    trainer.close()
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
