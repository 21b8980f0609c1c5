"""unetb200_bias_grad_static: the bias gradient for callers that cannot allocate a workspace.

Some users of this library are not allowed to call cudaMalloc at all -- a plug-in inside a host application that owns
every byte of device memory, or a training loop that runs under a fixed memory budget with a "no allocation after
start-up" rule.  For them the library offers one entry point whose scratch is a fixed array compiled into the library
(about 2.3 MiB per GPU, C <= 512).  This test drives it through the C ABI exactly like the regular entry point and
requires the same numbers, twice in a row with different shapes (the array is shared between calls and keeps the
previous call's contents, which must not leak into the next result), and requires an oversize request to be refused
with a message that names the limit.  It is collected last on purpose: it must pass both on its own and after
every other test has used the GPU."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _both(L, dz, pixels, C, dev):
    """Bias gradient through the workspace entry point and through the allocation-free one."""
    ws = torch.empty(int(L.lib().unetb200_bias_grad_ws_bytes(C)), dtype=torch.uint8, device=dev)
    want = torch.empty(C, dtype=torch.float32, device=dev)
    got = torch.full((C,), float("nan"), dtype=torch.float32, device=dev)
    L.call("unetb200_bias_grad", _p(dz), _p(want), _p(ws), ctypes.c_longlong(pixels), C, None)
    L.call("unetb200_bias_grad_static", _p(dz), _p(got), ctypes.c_longlong(pixels), C, None)
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_bias_grad_static_matches_workspace_version(cuda_device):
    from unet_distributed_b200 import _lib as L
    rng = np.random.default_rng(2024)
    # two calls in a row: the widest supported tensor first (fills the whole built-in array), then a narrow one
    for pixels, C in ((3 * 8 * 8, 512), (2 * 64 * 64, 32)):
        dz = torch.tensor(rng.standard_normal((pixels, C)).astype(np.float32)).to(torch.bfloat16).to(cuda_device)
        got, want = _both(L, dz, pixels, C, cuda_device)
        ref = dz.float().cpu().numpy().astype(np.float64).sum(0)
        # same kernels, same reduction order: identical to the workspace version, and right to fp32 accuracy
        assert np.array_equal(got, want), (C, np.abs(got - want).max())
        assert np.linalg.norm(got - ref) <= 1e-4 * np.linalg.norm(ref)
    # beyond the built-in array: refused, and the message says what the limit is
    dz = torch.zeros((16, 1024), dtype=torch.bfloat16, device=cuda_device)
    db = torch.zeros(1024, dtype=torch.float32, device=cuda_device)
    rc = L.lib().unetb200_bias_grad_static(_p(dz), _p(db), ctypes.c_longlong(16), 1024, None)
    assert rc < 0
    assert "512" in L.last_error() and "1024" in L.last_error()
