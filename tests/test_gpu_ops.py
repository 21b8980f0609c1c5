"""Per-operator parity through the C ABI (ctypes -> libunetb200.so) against the oracle's operators on the same
bf16-rounded inputs.  Shapes cover every tile path: 32-channel (SWIZZLE_64B) and 64+-channel (SWIZZLE_128B) K blocks,
N tiles of 32..256, two-source (concat-free) K loops, two-destination dgrad, multi-image tiles on 8x8 and 4x4 grids,
ragged grids (15x15, odd batch), more tiles than SMs (persistent loop, TMEM double buffering), split-K wgrad."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from util import O, rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-2   # bf16 output rounding (2^-9) plus fp32 accumulation-order differences


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _bf(x, dev):
    return torch.tensor(x, dtype=torch.float32).to(torch.bfloat16).to(dev).contiguous()


def _lib():
    from unet_distributed_b200 import _lib as L
    return L


def _rand(shape, rng, relu_like=False, scale=1.0):
    a = rng.standard_normal(shape).astype(np.float32) * scale
    if relu_like:
        a = np.maximum(a, 0)
    return torch.tensor(a).to(torch.bfloat16).float().numpy()


def _pack(name, w, *dims):
    L = _lib()
    out = torch.empty(w.numel(), dtype=torch.bfloat16, device=w.device)
    L.call(name, _p(w), _p(out), *dims, None)
    return out


CONV_CASES = [
    # N, H, W, C0, C1, Cout
    (1, 16, 16, 32, 0, 32),
    (2, 32, 32, 64, 0, 64),
    (2, 16, 16, 32, 32, 32),       # concat of two 32-channel sources
    (3, 8, 8, 256, 0, 512),        # two images per tile, odd batch, two N tiles of 256
    (2, 16, 16, 256, 256, 256),    # the conv6a shape
    (5, 4, 4, 128, 0, 128),        # four images per tile (64x64 inputs reach 4x4)
    (1, 15, 15, 64, 0, 128),       # ragged grid (240x240 inputs reach 15x15)
    (5, 64, 64, 32, 0, 64),        # 160 tiles > 148 SMs
    (2, 64, 32, 32, 32, 32),       # row-pair kernel: two 32-channel sources, N = 32, four strips
    (1, 32, 64, 64, 0, 32),        # row-pair kernel: 64-channel blocks, N = 32
    (1, 96, 40, 32, 0, 32),        # row-pair kernel: three row tiles, width 40 -> one strip per tile
]


@pytest.mark.parametrize("N,H,W,C0,C1,Co", CONV_CASES)
def test_conv3x3_fwd_dgrad_wgrad(cuda_device, N, H, W, C0, C1, Co):
    L = _lib()
    rng = np.random.default_rng(N * 1000 + H + C0 + Co)
    C = C0 + C1
    x0 = _rand((N, H, W, C0), rng, True)
    x1 = _rand((N, H, W, C1), rng, True) if C1 else None
    w = _rand((3, 3, C, Co), rng, scale=1.0 / np.sqrt(9 * C))
    b = rng.standard_normal(Co).astype(np.float32) * 0.1
    dz = _rand((N, H, W, Co), rng)
    dev = cuda_device
    d_x0, d_x1 = _bf(x0, dev), (_bf(x1, dev) if C1 else None)
    d_w = torch.tensor(w, device=dev)
    d_b = torch.tensor(b, device=dev)
    pk_f = _pack("unetb200_pack_conv3x3_fwd", d_w, C, Co)
    pk_d = _pack("unetb200_pack_conv3x3_dgrad", d_w, C, Co)
    y = torch.empty((N, H, W, Co), dtype=torch.bfloat16, device=dev)
    L.call("unetb200_conv3x3_fwd", _p(d_x0), C0, _p(d_x1), C1, _p(pk_f), _p(d_b), _p(y), N, H, W, Co, 1,
           ctypes.c_float(1.0), ctypes.c_uint64(0), ctypes.c_uint32(0), None)
    # oracle operators on the same values
    xc = np.concatenate([x0, x1], axis=3) if C1 else x0
    xt = torch.tensor(xc, requires_grad=True)
    wt = torch.tensor(w, requires_grad=True)
    z = O._conv3x3(xt.permute(0, 3, 1, 2), wt, torch.tensor(b))
    yr = F.relu(z)
    assert rel_err(y.float().cpu().numpy(), yr.permute(0, 2, 3, 1).detach().numpy()) < TOL
    z.backward(torch.tensor(dz).permute(0, 3, 1, 2))
    # dgrad: source 0 unmasked when there are two sources (a transposed-conv output), the skip source ReLU-masked
    d_dz = _bf(dz, dev)
    dx0 = torch.empty_like(d_x0)
    dx1 = torch.empty_like(d_x1) if C1 else None
    scale = 1.25
    if C1:
        L.call("unetb200_conv3x3_dgrad", _p(d_dz), Co, _p(pk_d), _p(dx0), C0, None, _p(dx1), C1, _p(d_x1),
               ctypes.c_float(scale), N, H, W, None)
        r0 = xt.grad.numpy()[..., :C0]
        r1 = xt.grad.numpy()[..., C0:] * (x1 > 0) * scale
        assert rel_err(dx1.float().cpu().numpy(), r1) < TOL
    else:
        L.call("unetb200_conv3x3_dgrad", _p(d_dz), Co, _p(pk_d), _p(dx0), C0, _p(d_x0), None, 0, None,
               ctypes.c_float(scale), N, H, W, None)
        r0 = xt.grad.numpy() * (x0 > 0) * scale
    assert rel_err(dx0.float().cpu().numpy(), r0) < TOL
    # wgrad (+ workspace protocol) and bias grad
    nws = L.lib().unetb200_conv3x3_wgrad_ws_bytes(N, H, W, C, Co)
    ws = torch.empty(max(int(nws), 16), dtype=torch.uint8, device=dev)
    dw = torch.full((3, 3, C, Co), float("nan"), dtype=torch.float32, device=dev)
    L.call("unetb200_conv3x3_wgrad", _p(d_dz), Co, _p(d_x0), C0, _p(d_x1), C1, _p(dw), _p(ws),
           ctypes.c_size_t(int(nws)), N, H, W, None)
    assert rel_err(dw.cpu().numpy(), wt.grad.numpy()) < 2e-3
    if nws > 0:
        rc = L.lib().unetb200_conv3x3_wgrad(_p(d_dz), Co, _p(d_x0), C0, _p(d_x1), C1, _p(dw), _p(ws),
                                            ctypes.c_size_t(64), N, H, W, None)
        assert rc < 0 and "workspace" in L.last_error()
    bws = torch.empty(int(L.lib().unetb200_bias_grad_ws_bytes(Co)), dtype=torch.uint8, device=dev)
    db = torch.empty(Co, dtype=torch.float32, device=dev)
    L.call("unetb200_bias_grad", _p(d_dz), _p(db), _p(bws), ctypes.c_longlong(N * H * W), Co, None)
    assert rel_err(db.cpu().numpy(), dz.reshape(-1, Co).sum(0)) < 1e-4
    torch.cuda.synchronize()


def test_conv3x3_dropout_matches_mask_generator(cuda_device):
    L = _lib()
    rng = np.random.default_rng(0)
    N, H, W, C, Co = 2, 32, 32, 64, 128
    x = _rand((N, H, W, C), rng, True)
    w = _rand((3, 3, C, Co), rng, scale=1.0 / np.sqrt(9 * C))
    dev = cuda_device
    d_x, d_w = _bf(x, dev), torch.tensor(w, device=dev)
    pk = _pack("unetb200_pack_conv3x3_fwd", d_w, C, Co)
    y0 = torch.empty((N, H, W, Co), dtype=torch.bfloat16, device=dev)
    y1 = torch.empty_like(y0)
    args = (_p(d_x), C, None, 0, _p(pk), None)
    L.call("unetb200_conv3x3_fwd", *args, _p(y0), N, H, W, Co, 1, ctypes.c_float(1.0), ctypes.c_uint64(0),
           ctypes.c_uint32(0), None)
    L.call("unetb200_conv3x3_fwd", *args, _p(y1), N, H, W, Co, 1, ctypes.c_float(0.8), ctypes.c_uint64(77),
           ctypes.c_uint32(3), None)
    keep = torch.empty(y0.numel(), dtype=torch.uint8, device=dev)
    L.call("unetb200_dropout_mask", _p(keep), ctypes.c_longlong(keep.numel()), ctypes.c_float(0.8),
           ctypes.c_uint64(77), ctypes.c_uint32(3), None)
    k = keep.view(N, H, W, Co).float()
    assert abs(float(k.mean()) - 0.8) < 5e-3
    # floor(0.8 + U) / 0.8 semantics [TF-1.4 dropout]: kept values are scaled by 1/keep, the rest are zero
    assert rel_err(y1.float().cpu().numpy(), (y0.float() * k / 0.8).cpu().numpy()) < 6e-3
    other = torch.empty_like(keep)
    L.call("unetb200_dropout_mask", _p(other), ctypes.c_longlong(keep.numel()), ctypes.c_float(0.8),
           ctypes.c_uint64(78), ctypes.c_uint32(3), None)
    assert not torch.equal(other, keep)


@pytest.mark.parametrize("N,H,W,Ci,Co", [(2, 8, 8, 512, 256), (2, 16, 16, 64, 32), (1, 32, 32, 128, 64),
                                         (3, 4, 4, 512, 256), (1, 15, 15, 512, 256)])
def test_tconv2x2(cuda_device, N, H, W, Ci, Co):
    L = _lib()
    rng = np.random.default_rng(H + Ci)
    x = _rand((N, H, W, Ci), rng, True)
    w = _rand((2, 2, Co, Ci), rng, scale=1.0 / np.sqrt(Ci))
    b = rng.standard_normal(Co).astype(np.float32) * 0.1
    dy = _rand((N, 2 * H, 2 * W, Co), rng)
    dev = cuda_device
    d_x, d_w, d_b, d_dy = _bf(x, dev), torch.tensor(w, device=dev), torch.tensor(b, device=dev), _bf(dy, dev)
    pk_f = _pack("unetb200_pack_tconv2x2_fwd", d_w, Ci, Co)
    pk_d = _pack("unetb200_pack_tconv2x2_dgrad", d_w, Ci, Co)
    y = torch.empty((N, 2 * H, 2 * W, Co), dtype=torch.bfloat16, device=dev)
    L.call("unetb200_tconv2x2_fwd", _p(d_x), Ci, _p(pk_f), _p(d_b), _p(y), Co, N, H, W, None)
    xt = torch.tensor(x, requires_grad=True)
    wt = torch.tensor(w, requires_grad=True)
    yr = O._tconv2x2(xt.permute(0, 3, 1, 2), wt, torch.tensor(b))
    assert rel_err(y.float().cpu().numpy(), yr.permute(0, 2, 3, 1).detach().numpy()) < TOL
    assert rel_err(yr.permute(0, 2, 3, 1).detach().numpy(), O.np_tconv2x2(x, w, b)) < 1e-5 or H > 8
    yr.backward(torch.tensor(dy).permute(0, 3, 1, 2))
    dx = torch.empty_like(d_x)
    L.call("unetb200_tconv2x2_dgrad", _p(d_dy), Co, _p(pk_d), _p(dx), Ci, _p(d_x), ctypes.c_float(1.0), N, H, W, None)
    assert rel_err(dx.float().cpu().numpy(), xt.grad.numpy() * (x > 0)) < TOL
    nws = int(L.lib().unetb200_tconv2x2_wgrad_ws_bytes(N, H, W, Ci, Co))
    ws = torch.empty(max(nws, 16), dtype=torch.uint8, device=dev)
    dw = torch.empty((2, 2, Co, Ci), dtype=torch.float32, device=dev)
    L.call("unetb200_tconv2x2_wgrad", _p(d_dy), Co, _p(d_x), Ci, _p(dw), _p(ws), ctypes.c_size_t(nws), N, H, W, None)
    assert rel_err(dw.cpu().numpy(), wt.grad.numpy()) < 2e-3
    # weight and bias gradient from one pass over dy: same dW bit for bit, db = column sums of the bf16 dy
    dw2 = torch.empty_like(dw)
    db = torch.full((Co,), float("nan"), dtype=torch.float32, device=dev)
    L.call("unetb200_tconv2x2_wgrad_bias_dt", _p(d_dy), Co, _p(d_x), Ci, _p(dw2), _p(db), _p(ws), ctypes.c_size_t(nws),
           N, H, W, L.BF16, None)
    assert torch.equal(dw2, dw)
    db_ref = d_dy.float().cpu().numpy().astype(np.float64).reshape(-1, Co).sum(axis=0)
    assert rel_err(db.cpu().numpy(), db_ref) < 1e-5
    db_again = torch.empty_like(db)
    L.call("unetb200_tconv2x2_wgrad_bias_dt", _p(d_dy), Co, _p(d_x), Ci, _p(dw2), _p(db_again), _p(ws),
           ctypes.c_size_t(nws), N, H, W, L.BF16, None)
    assert torch.equal(db_again, db)                                    # fixed summation order


@pytest.mark.parametrize("N,H,W,C", [(2, 16, 16, 32), (1, 8, 8, 512), (3, 30, 30, 128)])
def test_pool_and_upsample(cuda_device, N, H, W, C):
    L = _lib()
    rng = np.random.default_rng(C)
    x = _rand((N, H, W, C), rng, True)
    x.reshape(-1)[::7] = 0.5                                            # force ties inside windows
    dp = _rand((N, H // 2, W // 2, C), rng)
    skip = _rand((N, H, W, C), rng)
    dev = cuda_device
    d_x, d_dp, d_skip = _bf(x, dev), _bf(dp, dev), _bf(skip, dev)
    y = torch.empty((N, H // 2, W // 2, C), dtype=torch.bfloat16, device=dev)
    L.call("unetb200_maxpool2x2_fwd", _p(d_x), _p(y), N, H, W, C, None)
    assert np.array_equal(y.float().cpu().numpy(), O.np_maxpool2x2(x))
    L.call("unetb200_maxpool2x2_bwd", _p(d_x), _p(d_dp), _p(d_skip), _p(d_skip), N, H, W, C, None)   # in place
    xt = torch.tensor(x, requires_grad=True)
    F.max_pool2d(xt.permute(0, 3, 1, 2), 2).backward(torch.tensor(dp).permute(0, 3, 1, 2))
    ref = skip + xt.grad.numpy() * (x > 0)                              # first-max routing + ReLU gate + skip grad
    assert rel_err(d_skip.float().cpu().numpy(), ref) < 6e-3
    up = torch.empty((N, H, W, C), dtype=torch.bfloat16, device=dev)
    L.call("unetb200_upsample2x2_fwd", _p(y), _p(up), N, H // 2, W // 2, C, None)
    assert np.array_equal(up.float().cpu().numpy(), O.np_upsample2x2(O.np_maxpool2x2(x)))
    dx = torch.empty_like(y)
    d_g = _bf(skip, dev)
    L.call("unetb200_upsample2x2_bwd", _p(d_g), _p(y), _p(dx), N, H // 2, W // 2, C, 1, None)
    ref = skip.reshape(N, H // 2, 2, W // 2, 2, C).sum(axis=(2, 4)) * (O.np_maxpool2x2(x) > 0)
    assert rel_err(dx.float().cpu().numpy(), ref) < 6e-3


@pytest.mark.parametrize("Cin,ncl", [(1, 1), (4, 2)])
def test_first_conv_and_head(cuda_device, Cin, ncl):
    L = _lib()
    rng = np.random.default_rng(Cin)
    N, H, W, Co = 2, 32, 32, 32
    dev = cuda_device
    x = rng.standard_normal((N, H, W, Cin)).astype(np.float32)
    w = (rng.standard_normal((3, 3, Cin, Co)) * 0.5).astype(np.float32)
    b = (rng.standard_normal(Co) * 0.1).astype(np.float32)
    dz = _rand((N, H, W, Co), rng)
    d_x, d_w, d_b = (torch.tensor(a, device=dev) for a in (x, w, b))
    d_dz = _bf(dz, dev)
    y = torch.empty((N, H, W, Co), dtype=torch.bfloat16, device=dev)
    L.call("unetb200_conv_first_fwd", _p(d_x), _p(d_w), _p(d_b), _p(y), N, H, W, Cin, Co, None)
    xt, wt = torch.tensor(x), torch.tensor(w, requires_grad=True)
    z = O._conv3x3(xt.permute(0, 3, 1, 2), wt, torch.tensor(b))
    assert rel_err(y.float().cpu().numpy(), F.relu(z).permute(0, 2, 3, 1).detach().numpy()) < 6e-3
    z.backward(torch.tensor(dz).permute(0, 3, 1, 2))
    ws = torch.empty(int(L.lib().unetb200_conv_first_wgrad_ws_bytes(Cin, Co)), dtype=torch.uint8, device=dev)
    dw = torch.empty((3, 3, Cin, Co), dtype=torch.float32, device=dev)
    db = torch.empty(Co, dtype=torch.float32, device=dev)
    L.call("unetb200_conv_first_wgrad", _p(d_x), _p(d_dz), _p(dw), _p(db), _p(ws), N, H, W, Cin, Co, None)
    assert rel_err(dw.cpu().numpy(), wt.grad.numpy()) < 1e-4
    assert rel_err(db.cpu().numpy(), dz.reshape(-1, Co).sum(0)) < 1e-4
    # head: 1x1 conv + sigmoid + dice sums, and its backward
    P = N * H * W
    hx = _rand((N, H, W, Co), rng, True)
    hw = (rng.standard_normal((Co, ncl)) * 0.3).astype(np.float32)
    hb = (rng.standard_normal(ncl) * 0.1).astype(np.float32)
    yt = (rng.uniform(size=(N, H, W, ncl)) < 0.1).astype(np.float32)
    d_hx, d_hw, d_hb, d_yt = _bf(hx, dev), torch.tensor(hw, device=dev), torch.tensor(hb, device=dev), \
        torch.tensor(yt, device=dev)
    p = torch.empty((N, H, W, ncl), dtype=torch.float32, device=dev)
    sums = torch.empty(3, dtype=torch.float32, device=dev)
    out4 = torch.empty(4, dtype=torch.float32, device=dev)
    hws = torch.empty(int(L.lib().unetb200_head_ws_bytes(ctypes.c_longlong(P), Co, ncl)), dtype=torch.uint8, device=dev)
    L.call("unetb200_head_fwd", _p(d_hx), _p(d_hw), _p(d_hb), _p(d_yt), _p(p), _p(sums), _p(hws),
           ctypes.c_longlong(P), Co, ncl, None)
    L.call("unetb200_dice_finalize", _p(sums), ctypes.c_float(1.0), _p(out4), None)
    xt = torch.tensor(hx, dtype=torch.float64, requires_grad=True)
    wt = torch.tensor(hw, dtype=torch.float64, requires_grad=True)
    bt = torch.tensor(hb, dtype=torch.float64, requires_grad=True)
    pr = torch.sigmoid(xt @ wt + bt)
    ytt = torch.tensor(yt, dtype=torch.float64)
    loss = O.dice_coef_loss(ytt, pr)
    assert rel_err(p.cpu().numpy(), pr.detach().numpy()) < 1e-5
    ref4 = [float(f(ytt, pr).detach()) for f in (O.dice_coef_loss, O.dice_coef, O.sensitivity, O.specificity)]
    assert np.allclose(out4.cpu().numpy(), ref4, rtol=1e-5, atol=1e-6)
    loss.backward()
    dx = torch.empty_like(d_hx)
    dw2 = torch.empty((Co, ncl), dtype=torch.float32, device=dev)
    db2 = torch.empty(ncl, dtype=torch.float32, device=dev)
    L.call("unetb200_head_bwd", _p(d_hx), _p(d_hw), _p(d_yt), _p(p), _p(sums), ctypes.c_float(1.0),
           ctypes.c_float(1.0), _p(dx), _p(dw2), _p(db2), _p(hws), ctypes.c_longlong(P), Co, ncl, None)
    assert rel_err(dx.float().cpu().numpy(), xt.grad.numpy() * (hx > 0)) < 6e-3
    assert rel_err(dw2.cpu().numpy(), wt.grad.numpy()) < 1e-4
    assert rel_err(db2.cpu().numpy(), bt.grad.numpy()) < 1e-4


def test_adam_and_sgd_kernels(cuda_device):
    L = _lib()
    rng = np.random.default_rng(9)
    n = 100003                                                        # odd tail after the float4 body
    th = rng.standard_normal(n).astype(np.float32)
    g = (rng.standard_normal(n) * 0.01).astype(np.float32)
    dev = cuda_device
    d_th, d_g = torch.tensor(th, device=dev), torch.tensor(g, device=dev)
    d_m, d_v = torch.zeros_like(d_th), torch.zeros_like(d_th)
    step = torch.zeros(1, dtype=torch.int64, device=dev)
    w, m, v = {"a": th.astype(np.float64)}, {"a": np.zeros(n)}, {"a": np.zeros(n)}
    for s in range(4):
        L.call("unetb200_adam_step", _p(d_th), _p(d_m), _p(d_v), _p(d_g), ctypes.c_longlong(n), _p(step),
               ctypes.c_float(5e-4), 0, ctypes.c_float(0.25), ctypes.c_float(150.0), ctypes.c_float(0.9),
               ctypes.c_float(0.999), ctypes.c_float(1e-8), ctypes.c_float(0.5), None)
        L.call("unetb200_step_increment", _p(step), None)
        O.adam_update(w, m, v, {"a": g * 0.5}, s, lr0=5e-4, const_lr=False, lr_fraction=0.25, decay_steps=150)
    assert int(step.item()) == 4
    assert rel_err(d_th.cpu().numpy() - th, w["a"] - th) < 1e-3   # fp32 theta: the update is ~1e-3 of |theta|
    ws = {"a": th.astype(np.float64)}
    d_th2 = torch.tensor(th, device=dev)
    L.call("unetb200_sgd_step", _p(d_th2), _p(d_g), ctypes.c_longlong(n), _p(step), ctypes.c_float(0.1), 0,
           ctypes.c_float(0.5), ctypes.c_float(2.0), ctypes.c_float(1.0), None)
    O.sgd_update(ws, {"a": g}, 4, lr0=0.1, const_lr=False, lr_fraction=0.5, decay_steps=2)
    assert rel_err(d_th2.cpu().numpy() - th, ws["a"] - th) < 1e-3


def test_errors_are_reported_not_swallowed(cuda_device):
    L = _lib()
    x = torch.zeros((1, 16, 16, 24), dtype=torch.bfloat16, device=cuda_device)
    y = torch.zeros((1, 16, 16, 32), dtype=torch.bfloat16, device=cuda_device)
    rc = L.lib().unetb200_conv3x3_fwd(_p(x), 24, None, 0, _p(x), None, _p(y), 1, 16, 16, 32, 1, ctypes.c_float(1.0),
                                      ctypes.c_uint64(0), ctypes.c_uint32(0), None)
    assert rc < 0 and "multiples of 32" in L.last_error()
    with pytest.raises(L.UnetB200Error):
        L.call("unetb200_maxpool2x2_fwd", _p(x), _p(y), 1, 15, 16, 24, None)


def test_tensor_maps_are_cached(cuda_device):
    """CUtensorMaps are cached by (pointer, dims, box) -- the second identical call encodes none."""
    L = _lib()
    rng = np.random.default_rng(1)
    N, H, W, C, Co = 1, 16, 16, 64, 64
    d_x = _bf(_rand((N, H, W, C), rng, True), cuda_device)
    d_w = torch.tensor(_rand((3, 3, C, Co), rng, scale=0.05), device=cuda_device)
    pk = _pack("unetb200_pack_conv3x3_fwd", d_w, C, Co)
    y = torch.empty((N, H, W, Co), dtype=torch.bfloat16, device=cuda_device)
    h, m = ctypes.c_longlong(), ctypes.c_longlong()

    def run():
        L.call("unetb200_conv3x3_fwd", _p(d_x), C, None, 0, _p(pk), None, _p(y), N, H, W, Co, 1, ctypes.c_float(1.0),
               ctypes.c_uint64(0), ctypes.c_uint32(0), None)
        L.call("unetb200_tmap_cache_stats", ctypes.byref(h), ctypes.byref(m))
        return h.value, m.value

    h0, m0 = run()
    y0 = y.clone()
    h1, m1 = run()
    assert m1 == m0 and h1 >= h0 + 2, (h0, m0, h1, m1)
    assert torch.equal(y, y0)
    torch.cuda.synchronize()
