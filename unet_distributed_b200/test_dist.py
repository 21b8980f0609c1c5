"""Training driver with the reference's flag surface (test_dist.py:64-85; README.md:61-72 spellings accepted too).

    python -m unet_distributed_b200.test_dist --batch_size 128 --epochs 1 --use_upsampling=False
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m unet_distributed_b200.test_dist --batch_size 1024

`--batch_size` is the GLOBAL batch (test_dist.py:390): every rank trains on batch_size / world_size samples of each
batch.  The parameter server, Supervisor, TensorBoard and SavedModel export of the reference are out of scope; this
is the loop of test_dist.py:380-398, 448-459 plus the chief's test pass after every epoch and at the end
(test_dist.py:406-446, 461-490) and `last_good_model` checkpoints under TF variable names (.npz, see checkpoint.py)."""
from __future__ import annotations

import argparse
import os
import time

import numpy as np
import torch

from . import checkpoint, parallel, settings_dist
from .data import ResidentEpochs, get_epoch, load_all_data
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
    p.add_argument("--resident_data", type=_bool, default=False,
                   help="keep the training set on the GPU and gather batches there (no host copies in the step loop)")
    p.add_argument("--checkpoint_dir", type=str, default=settings_dist.CHECKPOINT_DIRECTORY)
    p.add_argument("--resume", type=_bool, default=False, help="start from <checkpoint_dir>/last_good_model.npz")
    p.add_argument("--test", type=_bool, default=True, help="run the chief's test pass after each epoch and at the end")
    return p


def main(argv=None):
    FLAGS = build_parser().parse_args(argv)
    if not FLAGS.is_sync:
        raise SystemExit("asynchronous parameter-server updates (--is_sync=0) are not implemented")
    world, rank, _ = parallel.init_from_env()
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
                      lr_fraction=FLAGS.lr_fraction, decay_steps=FLAGS.decay_steps, optimizer=FLAGS.optimizer)
    if FLAGS.resume:
        path = checkpoint.latest(FLAGS.checkpoint_dir)
        if path is None:
            raise SystemExit("--resume: no %s under %s" % (checkpoint.LAST_GOOD, FLAGS.checkpoint_dir))
        trainer.load_state_dict(checkpoint.load_checkpoint(path))       # every rank reads the same file
        if rank == 0:
            print("resumed from {} at global_step {}".format(path, trainer.global_step()), flush=True)

    def test_pass(epoch_no):
        """Chief only (test_dist.py:401-446): metrics over the test set, then the last-good-model checkpoint."""
        if rank != 0:
            return
        if FLAGS.test and imgs_test.shape[0] > batch_size:
            m = trainer.evaluate(imgs_test, msks_test, batch_size)
            print("\nEpoch {} of {}: TEST DATASET\nloss = {:.4f}\nDice = {:.4f}\nSensitivity = {:.4f}\n"
                  "Specificity = {:.4f}".format(epoch_no, FLAGS.epochs, m["loss"], m["dice"], m["sensitivity"],
                                                m["specificity"]), flush=True)
        checkpoint.save_checkpoint(os.path.join(FLAGS.checkpoint_dir, checkpoint.LAST_GOOD), trainer.state_dict())

    resident = None
    if FLAGS.resident_data:   # same epochs as get_epoch, cut on the device
        resident = ResidentEpochs(batch_size, imgs_train, msks_train, trainer.plan.device).shuffle(1234)
        epoch, num_batches = None, len(resident)
    else:
        epoch = get_epoch(batch_size, imgs_train, msks_train, rng=1234)   # same shuffle on every rank
        num_batches = len(epoch)
    total = num_batches * FLAGS.epochs
    if FLAGS.max_steps:
        total = min(total, FLAGS.max_steps)
    t0 = time.time()
    step = 0
    while step < total:
        batch_idx = step % num_batches
        report = (step + 1) % 10 == 0 or step + 1 == total
        if resident is not None:
            d_data, d_labels = resident.batch(batch_idx, start, end)
            met = trainer.train_step(d_data.contiguous(), d_labels.contiguous())
            met = met.cpu().numpy() if report else None
        else:
            data = np.ascontiguousarray(epoch[batch_idx, 0][start:end], dtype=np.float32)
            labels = np.ascontiguousarray(epoch[batch_idx, 1][start:end], dtype=np.float32)
            # steps whose numbers are printed synchronise; the others are pipelined (the copy of the next batch
            # overlaps the kernels of this one, trainer.train_step_host)
            met = trainer.train_step_host(data, labels, wait=report)
        step += 1
        if report:
            loss_v, dice_v = float(met[0]), float(met[1])
        if rank == 0 and report:
            dt = time.time() - t0
            print("step {}/{} (loss={:.4f}, dice={:.4f}) {:.1f} images/sec".format(
                step, total, loss_v, dice_v, step * batch_size / dt), flush=True)
        if batch_idx == num_batches - 1 and step < total:
            test_pass(step // num_batches)                                           # test_dist.py:406-446
            if resident is not None:
                resident.shuffle(1234 + step)
            else:
                epoch = get_epoch(batch_size, imgs_train, msks_train, rng=1234 + step)   # test_dist.py:448-452
    test_pass(max(step // max(num_batches, 1), 1))                                   # test_dist.py:461-490
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
