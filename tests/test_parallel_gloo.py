"""The N>1 host path on CPU: two gloo ranks, each with the oracle's gradient of ITS shard of the global batch,
averaged through the same bucketed all-reduce the GPU trainer uses (flat buffer laid out by the native planner).
The result must be SyncReplicasOptimizer's: the mean of the per-replica gradients (test_dist.py:251-258, 388-394)."""
import os
import socket
import tempfile

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from util import O, synthetic_batch

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(WORLD),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(2)
    from unet_distributed_b200 import parallel
    from unet_distributed_b200.plan import describe_plan
    ws, rk, _ = parallel.init_from_env("gloo")
    assert (ws, rk) == (WORLD, rank) and parallel.world() == (WORLD, rank)
    imgs, msks = synthetic_batch(4, 16, 16)                  # the global batch, identical on every rank
    weights = O.init_weights(5)
    s, e = parallel.shard_range(4, ws, rk)
    res = O.loss_and_grads(weights, imgs[s:e], msks[s:e], False, None, 0.0, torch.float32)
    d = describe_plan(e - s, 16, 16)
    flat = torch.zeros(d["num_params_padded"], dtype=torch.float32)
    for k, (off, shape) in d["variables"].items():
        flat[off:off + int(np.prod(shape))] = torch.from_numpy(res["grads"][k].reshape(-1))
    # weights start identical everywhere: rank 0 broadcasts
    wflat = torch.full((8,), float(rank))
    parallel.broadcast_flat(wflat)
    assert torch.all(wflat == 0)
    red = parallel.BucketedAllReduce(flat, d["buckets"])
    for seg in range(len(d["buckets"])):                      # launched in the order backward completes them
        red.launch(seg)
    red.wait()
    flat *= 1.0 / ws                                          # folded into the optimizer's grad_scale on the GPU
    assert red.bytes_per_step() == d["num_params_padded"] * 4
    if rank == 0:
        np.save(os.path.join(outdir, "avg.npy"), flat.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_reproduces_sync_replicas():
    from unet_distributed_b200.plan import describe_plan
    with tempfile.TemporaryDirectory() as td:
        mp.spawn(_worker, args=(_free_port(), td), nprocs=WORLD, join=True)
        got = np.load(os.path.join(td, "avg.npy"))
    imgs, msks = synthetic_batch(4, 16, 16)
    weights = O.init_weights(5)
    mean, _ = O.sync_replica_grads(weights, imgs, msks, WORLD, use_upsampling=False, dropout=0.0, dtype=torch.float32)
    d = describe_plan(2, 16, 16)
    for k, (off, shape) in d["variables"].items():
        n = int(np.prod(shape))
        a, b = got[off:off + n].reshape(shape).astype(np.float64), mean[k]
        assert np.linalg.norm(a - b) <= 1e-6 * np.linalg.norm(b) + 1e-12, k


def test_buckets_tile_the_flat_buffer():
    """The gradient buckets (one per backward segment) are contiguous, ordered from the end of the flat buffer to its
    start (the order backward completes them) and tile [0, num_params) exactly, for both decoders and both inputs;
    every variable lies inside exactly one bucket."""
    from unet_distributed_b200.plan import describe_plan
    for ups in (False, True):
        for (H, C) in ((128, 1), (240, 4)):
            d = describe_plan(2, H, H, in_channels=C, use_upsampling=ups)
            b = d["buckets"]
            assert len(b) == 5 and b[0][1] == d["num_params_padded"] and b[-1][0] == 0
            for (lo, hi), (lo2, hi2) in zip(b[:-1], b[1:]):
                assert lo < hi and hi2 == lo and lo % 4 == 0
            for name, (off, shape) in d["variables"].items():
                n = int(np.prod(shape))
                assert sum(1 for lo, hi in b if lo <= off and off + n <= hi) == 1, name
            first = {next(i for i, (lo, hi) in enumerate(b) if lo <= off < hi) for name, (off, _) in
                     d["variables"].items() if name.startswith(("conv1", "conv2"))}
            assert first == {4}, "the last (exposed) bucket holds conv1a..conv2b only"
            assert (b[4][1] - b[4][0]) * 4 < 300_000
