"""SyncReplicasOptimizer semantics ON THE GPUS (test_dist.py:251-262, shard split :388-394): a 2-rank NCCL Trainer step
-- backward segments, bucket all-reduces on NCCL's stream, per-bucket Adam + operand re-pack on the update stream,
1/N folded into Adam's grad_scale, eager and replayed from ONE CUDA graph -- must equal, on one GPU, "run the two
shards one after the other, sum their gradients, apply one Adam update with grad_scale 1/2".  Reductions run in a
fixed order and a two-term sum is commutative, so the parameters are required to be bit-identical.

Skipped on a box with fewer than two GPUs (scripts/gpu_multi_parity.sh runs it under `gpurun --gpus 2`)."""
import ctypes
import os
import socket
import tempfile

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from util import O, DROPOUT_SEED, act_names, rel_err, synthetic_batch

pytestmark = pytest.mark.gpu

WORLD = 2
N, H, W = 4, 64, 64           # per-replica batch; the global batch is WORLD * N
STEPS = 3
OPT = dict(learning_rate=1e-3, const_learningrate=False, lr_fraction=0.25, decay_steps=5)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _batches(steps=STEPS):
    return [synthetic_batch(WORLD * N, H, W, seed=300 + i) for i in range(steps)]


def _worker(rank, port, outdir, ups, graph, reserve):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(WORLD),
                      LOCAL_RANK=str(rank), UNETB200_RESERVE_SMS=str(reserve))
    from unet_distributed_b200 import parallel
    from unet_distributed_b200.trainer import Trainer
    ws, rk, local = parallel.init_from_env("nccl")
    dev = torch.device("cuda", local)
    # rank 1 starts from DIFFERENT weights: the initial broadcast has to overwrite them
    weights = O.init_weights(1234 if rank == 0 else 99, ups)
    tr = Trainer(N, H, W, use_upsampling=ups, dropout=0.2, weights=weights, dropout_seed=DROPOUT_SEED,
                 use_cuda_graph=graph, **OPT)
    assert tr.dropout_seed == parallel.replica_seed(DROPOUT_SEED, rank)
    s, e = parallel.shard_range(WORLD * N, ws, rk)
    mets = []
    for imgs, msks in _batches():
        met = tr.train_step(torch.tensor(imgs[s:e], device=dev), torch.tensor(msks[s:e], device=dev))
        mets.append(met.clone())
    torch.cuda.synchronize()
    np.save(os.path.join(outdir, f"params{rank}.npy"), tr.plan.params.cpu().numpy())
    np.save(os.path.join(outdir, f"grads{rank}.npy"), tr.plan.grads.cpu().numpy())
    np.save(os.path.join(outdir, f"mets{rank}.npy"), torch.stack(mets).cpu().numpy())
    assert tr.global_step() == STEPS
    tr.close()                     # graphs that captured NCCL kernels must go before the process group does
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def _single_gpu_reference(ups, device, steps=STEPS):
    """One GPU, no process group: per step, each shard's forward/backward with that replica's dropout seed, gradients
    summed in fp32, ONE Adam update with grad_scale = 1/WORLD, weights re-packed."""
    from unet_distributed_b200 import _lib, parallel
    from unet_distributed_b200.plan import UNetPlan
    plan = UNetPlan(N, H, W, 1, 1, ups, 0.2)
    plan.set_weights(O.init_weights(1234, ups))
    m, v = torch.zeros_like(plan.params), torch.zeros_like(plan.params)
    step = torch.zeros(1, dtype=torch.int64, device=device)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    mets, last_sum, shard_state = [], None, None
    for imgs, msks in _batches(steps):
        total = torch.zeros_like(plan.grads)
        row, shard_state = [], []
        for r in range(WORLD):
            s, e = parallel.shard_range(WORLD * N, WORLD, r)
            d_i, d_m = torch.tensor(imgs[s:e], device=device), torch.tensor(msks[s:e], device=device)
            plan.grads.zero_()
            plan.forward(d_i, d_m, seed=parallel.replica_seed(DROPOUT_SEED, r), seed_add=step)
            plan.backward(d_i, d_m)
            total += plan.grads
            row.append(plan.metrics.clone())
            shard_state.append({o: plan.tensor(p).float().cpu().numpy() for p, o in act_names(ups).items()})
        mets.append(torch.stack(row))
        last_sum = total.clone()
        n = ctypes.c_longlong(plan.num_params_padded)
        _lib.call("unetb200_adam_step", ctypes.c_void_p(plan.params.data_ptr()), ctypes.c_void_p(m.data_ptr()),
                  ctypes.c_void_p(v.data_ptr()), ctypes.c_void_p(total.data_ptr()), n, ctypes.c_void_p(step.data_ptr()),
                  OPT["learning_rate"], 0, OPT["lr_fraction"], float(OPT["decay_steps"]), 0.9, 0.999, 1e-8,
                  1.0 / WORLD, stream)
        _lib.call("unetb200_step_increment", ctypes.c_void_p(step.data_ptr()), stream)
        plan.pack_weights()
    torch.cuda.synchronize()
    return plan, (None if last_sum is None else last_sum.cpu().numpy()), \
        (torch.stack(mets).cpu().numpy() if mets else None), shard_state


# reserve = SMs the backward kernels leave to NCCL (unetb200_set_reserved_sms). With a reservation the split-K factors of
# the weight-gradient kernels differ from the single-GPU reference's, i.e. partial sums are added in another (still
# fixed) order: bit-equality is required without it, 1e-4 (parameter vector, relative) with the default reservation.
@pytest.mark.parametrize("ups,graph,reserve", [(False, False, 0), (False, True, 0), (True, True, 0), (False, True, 8)])
def test_two_rank_step_equals_mean_of_shard_gradients(cuda_device, ups, graph, reserve):
    if torch.cuda.device_count() < WORLD:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    with tempfile.TemporaryDirectory() as td:
        mp.spawn(_worker, args=(_free_port(), td, ups, graph, reserve), nprocs=WORLD, join=True)
        params = [np.load(os.path.join(td, f"params{r}.npy")) for r in range(WORLD)]
        grads = [np.load(os.path.join(td, f"grads{r}.npy")) for r in range(WORLD)]
        mets = [np.load(os.path.join(td, f"mets{r}.npy")) for r in range(WORLD)]
    # replicas stay identical (same start after the broadcast, same reduced gradients, same update)
    np.testing.assert_array_equal(params[0], params[1])
    np.testing.assert_array_equal(grads[0], grads[1])
    plan, want_sum, want_mets, shard_state = _single_gpu_reference(ups, cuda_device)
    for r in range(WORLD):   # each replica reported the dice loss of ITS shard
        if reserve == 0:
            np.testing.assert_array_equal(mets[r], want_mets[:, r])
        else:
            np.testing.assert_allclose(mets[r], want_mets[:, r], rtol=1e-4, atol=1e-5)
    # all-reduced sum == sum of the shard gradients (variables only: the padding between them is never written)
    for k, (off, shape) in plan.variables.items():
        n = int(np.prod(shape))
        if reserve == 0:
            np.testing.assert_array_equal(grads[0][off:off + n], want_sum[off:off + n], err_msg=k)
        else:
            assert rel_err(grads[0][off:off + n], want_sum[off:off + n]) < 1e-4, k
    want_params = plan.params.cpu().numpy()
    bit_equal = np.array_equal(params[0], want_params)
    err = rel_err(params[0], want_params)
    out = os.environ.get("UNETB200_TEST_REPORT")
    if out:
        with open(out, "a") as f:
            f.write(f"# 2-rank NCCL step vs single-GPU mean-of-shard-gradients, ups={ups} graph={graph} "
                    f"reserved_sms={reserve}: parameters bit-equal={bit_equal} rel_err={err:.3e} after {STEPS} steps\n")
    assert err <= (1e-4 if reserve else 1e-6) and (bit_equal or reserve > 0)

    # and the averaged gradient is the oracle's SyncReplicasOptimizer gradient (mean over replicas of the gradient of
    # each replica's own shard loss), at the bf16 bar: oracle backward evaluated at each shard's GPU forward state
    imgs, msks = _batches()[-1]
    weights = _single_gpu_reference(ups, cuda_device, STEPS - 1)[0].get_weights()   # parameters before the last step
    mean = None
    for r in range(WORLD):
        s, e = r * N, (r + 1) * N
        emu = O.loss_and_grads(weights, imgs[s:e], msks[s:e], ups, None, 0.2, torch.float32, emulate_bf16=True,
                               forced_acts=shard_state[r])
        mean = emu["grads"] if mean is None else {k: mean[k] + g for k, g in emu["grads"].items()}
    bad = []
    for k, (off, shape) in plan.variables.items():
        n = int(np.prod(shape))
        got = grads[0][off:off + n].reshape(shape) / WORLD
        e = rel_err(got, mean[k] / WORLD)
        if not e < 2e-2:
            bad.append(f"{k}: {e:.3e}")
    assert not bad, "; ".join(bad)
