"""The device-resident input pipeline on the GPU: unetb200_gather_batch, called through the
C ABI, against data.get_epoch (data.py:29-54) and the per-replica shard slice of test_dist.py:390-394.
Bit-exact: a gather of fp32 rows copies, a gather of float64 rows rounds like numpy's astype(float32)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _sets(n=37, H=16, W=12, C=4, seed=0):
    rng = np.random.default_rng(seed)
    imgs = rng.standard_normal((n, H, W, C))                       # float64, like update_channels' output
    msks = (rng.random((n, H, W, 1)) > 0.5).astype(np.float64)
    return imgs, msks


@pytest.mark.parametrize("resident_dtype", [torch.float32, torch.float64])
def test_resident_epochs_match_get_epoch_on_gpu(cuda_device, resident_dtype):
    from unet_distributed_b200.data import ResidentEpochs, get_epoch
    from unet_distributed_b200.parallel import shard_range
    imgs, msks = _sets()
    B = 8
    ep = get_epoch(B, imgs, msks, rng=1234)                         # EpochOfBatches: image / mask channels differ
    res = ResidentEpochs(B, imgs, msks, cuda_device, dtype=resident_dtype).shuffle(1234)
    assert len(res) == len(ep) == 4
    for b in range(len(ep)):
        x, y = res.batch(b)
        assert x.dtype == torch.float32 and x.is_cuda
        np.testing.assert_array_equal(x.cpu().numpy(), ep[b, 0].astype(np.float32))
        np.testing.assert_array_equal(y.cpu().numpy(), ep[b, 1].astype(np.float32))
        for world in (2, 4):
            for rank in range(world):
                s, e = shard_range(B, world, rank)
                xs, ys = res.batch(b, s, e)
                np.testing.assert_array_equal(xs.cpu().numpy(), ep[b, 0][s:e].astype(np.float32))
                np.testing.assert_array_equal(ys.cpu().numpy(), ep[b, 1][s:e].astype(np.float32))
    # preallocated step-input slots are filled in place
    slot = (torch.zeros(B, 16, 12, 4, device=cuda_device), torch.zeros(B, 16, 12, 1, device=cuda_device))
    got = res.batch(1, out=slot)
    assert got[0].data_ptr() == slot[0].data_ptr()
    np.testing.assert_array_equal(slot[0].cpu().numpy(), ep[1, 0].astype(np.float32))
    res.shuffle(1235)
    assert not np.array_equal(res.batch(0)[0].cpu().numpy(), ep[0, 0].astype(np.float32))


def test_gather_batch_abi_edges(cuda_device):
    """Identity gather (perm = NULL), an empty batch, a full-size BraTS row, and the error paths."""
    from unet_distributed_b200 import _lib
    L = _lib.lib()
    s = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    src = torch.randn(9, 128 * 128, device=cuda_device, dtype=torch.float64)
    dst = torch.full((4, 128 * 128), -1.0, device=cuda_device)
    _lib.call("unetb200_gather_batch", ctypes.c_void_p(src.data_ptr()), _lib.F64, ctypes.c_longlong(9), None,
              ctypes.c_longlong(5), 4, ctypes.c_longlong(128 * 128), ctypes.c_void_p(dst.data_ptr()), s)
    assert torch.equal(dst, src[5:9].float())
    before = dst.clone()
    _lib.call("unetb200_gather_batch", ctypes.c_void_p(src.data_ptr()), _lib.F64, ctypes.c_longlong(9), None,
              ctypes.c_longlong(0), 0, ctypes.c_longlong(128 * 128), ctypes.c_void_p(dst.data_ptr()), s)
    assert torch.equal(dst, before)
    perm = torch.tensor([8, 0, 3, 3], device=cuda_device, dtype=torch.int64)
    _lib.call("unetb200_gather_batch", ctypes.c_void_p(src.data_ptr()), _lib.F64, ctypes.c_longlong(9),
              ctypes.c_void_p(perm.data_ptr()), ctypes.c_longlong(0), 4, ctypes.c_longlong(128 * 128),
              ctypes.c_void_p(dst.data_ptr()), s)
    assert torch.equal(dst, src[perm].float())
    # rows outside the set without a permutation, an unknown dtype and a ragged row are refused
    for args in ((_lib.F64, 7, 4, 128 * 128), (_lib.BF16, 0, 4, 128 * 128), (_lib.F32, 0, 4, 6)):
        rc = L.unetb200_gather_batch(ctypes.c_void_p(src.data_ptr()), args[0], ctypes.c_longlong(9), None,
                                     ctypes.c_longlong(args[1]), args[2], ctypes.c_longlong(args[3]),
                                     ctypes.c_void_p(dst.data_ptr()), s)
        assert rc < 0 and "gather_batch" in _lib.last_error()
    torch.cuda.synchronize()


def test_resident_training_matches_host_fed(cuda_device):
    """Three steps fed by ResidentEpochs (gather kernel into the step's input tensors) give the same metrics and
    parameters as the same steps fed from get_epoch's host batches."""
    from unet_distributed_b200.data import ResidentEpochs, get_epoch
    from unet_distributed_b200.synthetic import make_slices
    from unet_distributed_b200.trainer import Trainer
    imgs, msks = make_slices(10, 32, 32, 1, seed=5)
    B = 4
    runs = []
    for resident in (False, True):
        tr = Trainer(B, 32, 32, dropout=0.2, learning_rate=1e-3)
        mets = []
        if resident:
            res = ResidentEpochs(B, imgs.astype(np.float64), msks.astype(np.float64), cuda_device,
                                 dtype=torch.float64).shuffle(7)
            for b in range(len(res)):
                x, y = res.batch(b)
                mets.append(tr.train_step(x, y).clone())
        else:
            ep = get_epoch(B, imgs, msks, rng=7)
            for b in range(len(ep)):
                x = torch.tensor(np.ascontiguousarray(ep[b, 0], dtype=np.float32), device=cuda_device)
                y = torch.tensor(np.ascontiguousarray(ep[b, 1], dtype=np.float32), device=cuda_device)
                mets.append(tr.train_step(x, y).clone())
        torch.cuda.synchronize()
        runs.append((torch.stack(mets).cpu().numpy(), tr.plan.params.cpu().numpy()))
    np.testing.assert_array_equal(runs[0][0], runs[1][0])
    np.testing.assert_array_equal(runs[0][1], runs[1][1])
