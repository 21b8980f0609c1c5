"""Synchronous data parallelism: the replacement for the reference's parameter-server SyncReplicasOptimizer
(test_dist.py:249-262) and its per-worker batch split (test_dist.py:388-394).

Semantics kept: every replica computes the dice loss on ITS OWN shard of the global batch and the applied gradient
is the MEAN of the replica gradients; global_step advances once per aggregated update.  Mechanism replaced: one
process per GPU, gradients live in one flat fp32 buffer that is all-reduced in a few contiguous buckets, each
launched as soon as the backward segment that completes it has been enqueued (overlap with the rest of backward).
The helpers are backend-agnostic (NCCL on GPUs, gloo in the CPU tests)."""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as dist


# NCCL may use 8 CTAs and the backward kernels that run beside an all-reduce leave 8 SMs out of their grids
# (profiles/r02_reserved_sms.txt, r02_nccl_max_ctas.txt); UNETB200_NCCL_MAX_CTAS / UNETB200_RESERVE_SMS override
# (0 = NCCL's default / no reservation)
NCCL_MAX_CTAS = 8
RESERVED_SMS = 8


def shard_range(batch_size, num_replicas, rank):
    """Slice of the global batch that replica `rank` trains on -- test_dist.py:390-392
    (data_range = int(batch_size / len(worker_hosts)); start = data_range * task_index)."""
    data_range = int(batch_size / num_replicas)
    start = data_range * rank
    return start, start + data_range


def replica_seed(seed, rank):
    """Dropout seed of replica `rank`: the reference's workers draw independent masks (every sess.run samples its own
    RandomUniform, model.py:60,66), so the rank is folded into the Philox seed (64-bit golden-ratio stride)."""
    return (int(seed) + int(rank) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


def init_from_env(backend=None):
    """Initialise torch.distributed from torchrun's environment (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*)."""
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws <= 1:
        return 1, 0, 0
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local)
    if not dist.is_initialized():
        kw = {"device_id": torch.device("cuda", local)} if backend == "nccl" else {}
        max_ctas = int(os.environ.get("UNETB200_NCCL_MAX_CTAS", str(NCCL_MAX_CTAS)))
        if backend == "nccl" and max_ctas > 0:
            # The gradient buckets are latency-bound (0.3-14 MB) and run beside persistent one-CTA-per-SM conv kernels:
            # every SM NCCL occupies delays one of those CTAs by a whole tile loop, so NCCL gets few CTAs.
            opts = dist.ProcessGroupNCCL.Options()
            opts.config.max_ctas = max_ctas
            opts.config.min_ctas = 1
            opts._timeout = datetime.timedelta(seconds=int(os.environ.get("UNETB200_DIST_TIMEOUT", "180")))
            kw["pg_options"] = opts
        # a mismatched collective should fail in minutes, not after the 10-minute default
        dist.init_process_group(backend=backend, rank=rank, world_size=ws,
                                timeout=datetime.timedelta(seconds=int(os.environ.get("UNETB200_DIST_TIMEOUT", "180"))),
                                **kw)
    return ws, rank, local


class BucketedAllReduce:
    """All-reduce (sum) of contiguous [begin, end) ranges of one flat buffer; the 1/N of the mean is folded into
    the optimizer's grad_scale, so nothing else touches the gradients."""

    def __init__(self, flat, buckets, group=None):
        self.flat = flat
        self.buckets = [(int(b), int(e)) for b, e in buckets]
        self.group = group
        self._pending = []

    def launch(self, bucket_index, track=True):
        """Enqueue the all-reduce of one bucket behind the work already on the current stream. Returns the work
        handle (None at world size 1); with track=True it is also kept for wait()."""
        if world()[0] == 1:
            return None
        b, e = self.buckets[bucket_index]
        if e <= b:
            return None
        work = dist.all_reduce(self.flat[b:e], op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        if track:
            self._pending.append(work)
        return work

    def reduce_now(self, bucket_index):
        """All-reduce of one bucket ordered on the CURRENT stream: work already enqueued on it comes first, work
        enqueued on it afterwards sees the reduced values (NCCL runs on its own stream in between)."""
        if world()[0] == 1:
            return
        b, e = self.buckets[bucket_index]
        if e > b:
            dist.all_reduce(self.flat[b:e], op=dist.ReduceOp.SUM, group=self.group)

    def wait(self):
        """Make the current stream (or, with gloo, the host) wait for every launched bucket."""
        for w in self._pending:
            w.wait()
        self._pending = []

    def bytes_per_step(self):
        return sum(e - b for b, e in self.buckets) * self.flat.element_size()


def broadcast_flat(flat, src=0, group=None):
    """Replicas start from identical weights: rank 0's initial values are broadcast once."""
    if world()[0] > 1:
        dist.broadcast(flat, src=src, group=group)
