"""The fp32/tf32 precision mode (north_star: "forward activations and per-layer gradients within 1e-3 relative in the
fp32/tf32 mode"; the reference computes in fp32, test_dist.py:196-202).  Activations, gradients and packed weights are
fp32 words rounded to tf32 by their producers, the convolutions are tcgen05 kind::tf32 MMAs with fp32 accumulation.

Three levels, all through the C ABI:
  * per operator (`_dt` entry points): every conv / transposed-conv pass against the oracle's operator on the same
    tf32-representable inputs -- outputs that are activations are themselves rounded to tf32 (half an ulp = 2.4e-4),
    weight gradients are plain fp32 sums (1e-5);
  * per layer inside the network: the oracle's fp32 backward evaluated at the GPU's forward state, every activation
    gradient and weight gradient within TOL_TF32 = 1e-3;
  * free-running against the fp32 oracle: every forward activation within TOL_TF32 x depth allowance, and never worse
    than TWIN_FACTOR x the oracle's own tf32-storage twin (the fp32 oracle with activations / weights rounded to tf32):
    rounding 23 layers deep compounds for ANY tf32 implementation, the twin shows by how much."""
import ctypes
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from util import O, DROPOUT_SEED, act_names, plan_dropout_masks, rel_err, synthetic_batch

pytestmark = pytest.mark.gpu

TOL_TF32 = 1e-3
TWIN_FACTOR = 1.5


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _lib():
    from unet_distributed_b200 import _lib as L
    return L


def _tf32(a):
    return O._tf32_round(torch.tensor(np.asarray(a, np.float32))).numpy()


def _rand(shape, rng, relu_like=False, scale=1.0):
    a = rng.standard_normal(shape).astype(np.float32) * scale
    if relu_like:
        a = np.maximum(a, 0)
    return _tf32(a)


def _is_tf32(t):
    return bool(((t.contiguous().view(torch.int32) & 0x1FFF) == 0).all())


def _pack(name, w, *dims):
    L = _lib()
    out = torch.empty(w.numel(), dtype=torch.float32, device=w.device)
    L.call(name, _p(w), _p(out), *dims, L.TF32, None)
    assert _is_tf32(out)
    return out


CONV_CASES = [
    # N, H, W, C0, C1, Cout
    (1, 16, 16, 32, 0, 32),
    (2, 32, 32, 64, 0, 64),
    (2, 16, 16, 32, 32, 32),       # concat of two sources
    (3, 8, 8, 256, 0, 512),        # two images per tile, odd batch, two N tiles of 256
    (1, 15, 15, 64, 0, 128),       # ragged grid
    (5, 64, 64, 32, 0, 64),        # more tiles than SMs
]


@pytest.mark.parametrize("N,H,W,C0,C1,Co", CONV_CASES)
def test_conv3x3_tf32(cuda_device, N, H, W, C0, C1, Co):
    L = _lib()
    rng = np.random.default_rng(N * 1000 + H + C0 + Co)
    C = C0 + C1
    x0 = _rand((N, H, W, C0), rng, True)
    x1 = _rand((N, H, W, C1), rng, True) if C1 else None
    w = _rand((3, 3, C, Co), rng, scale=1.0 / np.sqrt(9 * C))
    b = rng.standard_normal(Co).astype(np.float32) * 0.1
    dz = _rand((N, H, W, Co), rng)
    dev = cuda_device
    d_x0, d_x1 = torch.tensor(x0, device=dev), (torch.tensor(x1, device=dev) if C1 else None)
    d_w, d_b = torch.tensor(w, device=dev), torch.tensor(b, device=dev)
    pk_f = _pack("unetb200_pack_conv3x3_fwd_dt", d_w, C, Co)
    pk_d = _pack("unetb200_pack_conv3x3_dgrad_dt", d_w, C, Co)
    y = torch.empty((N, H, W, Co), dtype=torch.float32, device=dev)
    L.call("unetb200_conv3x3_fwd_dt", _p(d_x0), C0, _p(d_x1), C1, _p(pk_f), _p(d_b), _p(y), N, H, W, Co, 1,
           ctypes.c_float(1.0), ctypes.c_uint64(0), ctypes.c_uint32(0), L.TF32, None)
    xc = np.concatenate([x0, x1], axis=3) if C1 else x0
    xt = torch.tensor(xc, dtype=torch.float64, requires_grad=True)
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    z = O._conv3x3(xt.permute(0, 3, 1, 2), wt, torch.tensor(b, dtype=torch.float64))
    yr = F.relu(z).permute(0, 2, 3, 1).detach().numpy()
    assert _is_tf32(y), "activation outputs of the tf32 mode are rounded to tf32"
    assert rel_err(y.cpu().numpy(), yr) < TOL_TF32
    assert rel_err(y.cpu().numpy(), _tf32(yr)) < 1e-4       # only the final rounding (and its ties) separates the two
    z.backward(torch.tensor(dz, dtype=torch.float64).permute(0, 3, 1, 2))
    d_dz = torch.tensor(dz, device=dev)
    dx0 = torch.empty_like(d_x0)
    dx1 = torch.empty_like(d_x1) if C1 else None
    scale = 1.25
    if C1:
        L.call("unetb200_conv3x3_dgrad_dt", _p(d_dz), Co, _p(pk_d), _p(dx0), C0, None, _p(dx1), C1, _p(d_x1),
               ctypes.c_float(scale), N, H, W, L.TF32, None)
        r0 = xt.grad.numpy()[..., :C0]
        r1 = xt.grad.numpy()[..., C0:] * (x1 > 0) * scale
        assert rel_err(dx1.cpu().numpy(), r1) < TOL_TF32
    else:
        L.call("unetb200_conv3x3_dgrad_dt", _p(d_dz), Co, _p(pk_d), _p(dx0), C0, _p(d_x0), None, 0, None,
               ctypes.c_float(scale), N, H, W, L.TF32, None)
        r0 = xt.grad.numpy() * (x0 > 0) * scale
    assert rel_err(dx0.cpu().numpy(), r0) < TOL_TF32
    nws = L.lib().unetb200_conv3x3_wgrad_ws_bytes(N, H, W, C, Co)
    ws = torch.empty(max(int(nws), 16), dtype=torch.uint8, device=dev)
    dw = torch.full((3, 3, C, Co), float("nan"), dtype=torch.float32, device=dev)
    L.call("unetb200_conv3x3_wgrad_dt", _p(d_dz), Co, _p(d_x0), C0, _p(d_x1), C1, _p(dw), _p(ws),
           ctypes.c_size_t(int(nws)), N, H, W, L.TF32, None)
    assert rel_err(dw.cpu().numpy(), wt.grad.numpy()) < 1e-5
    bws = torch.empty(int(L.lib().unetb200_bias_grad_ws_bytes(Co)), dtype=torch.uint8, device=dev)
    db = torch.empty(Co, dtype=torch.float32, device=dev)
    L.call("unetb200_bias_grad_dt", _p(d_dz), _p(db), _p(bws), ctypes.c_longlong(N * H * W), Co, L.TF32, None)
    assert rel_err(db.cpu().numpy(), dz.astype(np.float64).reshape(-1, Co).sum(0)) < 1e-5
    torch.cuda.synchronize()


@pytest.mark.parametrize("N,H,W,Ci,Co", [(2, 8, 8, 512, 256), (2, 16, 16, 64, 32), (1, 15, 15, 128, 64)])
def test_tconv2x2_tf32(cuda_device, N, H, W, Ci, Co):
    L = _lib()
    rng = np.random.default_rng(H + Ci)
    x = _rand((N, H, W, Ci), rng, True)
    w = _rand((2, 2, Co, Ci), rng, scale=1.0 / np.sqrt(Ci))
    b = rng.standard_normal(Co).astype(np.float32) * 0.1
    dy = _rand((N, 2 * H, 2 * W, Co), rng)
    dev = cuda_device
    d_x, d_w, d_b, d_dy = (torch.tensor(a, device=dev) for a in (x, w, b, dy))
    pk_f = _pack("unetb200_pack_tconv2x2_fwd_dt", d_w, Ci, Co)
    pk_d = _pack("unetb200_pack_tconv2x2_dgrad_dt", d_w, Ci, Co)
    y = torch.empty((N, 2 * H, 2 * W, Co), dtype=torch.float32, device=dev)
    L.call("unetb200_tconv2x2_fwd_dt", _p(d_x), Ci, _p(pk_f), _p(d_b), _p(y), Co, N, H, W, L.TF32, None)
    xt = torch.tensor(x, dtype=torch.float64, requires_grad=True)
    wt = torch.tensor(w, dtype=torch.float64, requires_grad=True)
    yr = O._tconv2x2(xt.permute(0, 3, 1, 2), wt, torch.tensor(b, dtype=torch.float64))
    assert rel_err(y.cpu().numpy(), yr.permute(0, 2, 3, 1).detach().numpy()) < TOL_TF32
    yr.backward(torch.tensor(dy, dtype=torch.float64).permute(0, 3, 1, 2))
    dx = torch.empty_like(d_x)
    L.call("unetb200_tconv2x2_dgrad_dt", _p(d_dy), Co, _p(pk_d), _p(dx), Ci, _p(d_x), ctypes.c_float(1.0), N, H, W,
           L.TF32, None)
    assert rel_err(dx.cpu().numpy(), xt.grad.numpy() * (x > 0)) < TOL_TF32
    nws = L.lib().unetb200_tconv2x2_wgrad_ws_bytes(N, H, W, Ci, Co)
    ws = torch.empty(max(int(nws), 16), dtype=torch.uint8, device=dev)
    dw = torch.full((2, 2, Co, Ci), float("nan"), dtype=torch.float32, device=dev)
    L.call("unetb200_tconv2x2_wgrad_dt", _p(d_dy), Co, _p(d_x), Ci, _p(dw), _p(ws), ctypes.c_size_t(int(nws)), N, H, W,
           L.TF32, None)
    assert rel_err(dw.cpu().numpy(), wt.grad.numpy()) < 1e-5
    torch.cuda.synchronize()


def test_bandwidth_kernels_fp32(cuda_device):
    """MaxPool / MaxPoolGrad / nearest upsampling and its gradient on fp32 activations: exact against NumPy (sums of
    tf32 values are re-rounded to tf32 on store)."""
    L = _lib()
    rng = np.random.default_rng(3)
    N, H, W, C = 2, 8, 12, 32
    dev = cuda_device
    x = _rand((N, H, W, C), rng, True)
    x[0, 0:2, 0:2, :4] = 0.5          # forced ties: the first maximum in window order takes the gradient
    d_x = torch.tensor(x, device=dev)
    y = torch.empty((N, H // 2, W // 2, C), dtype=torch.float32, device=dev)
    L.call("unetb200_maxpool2x2_fwd_dt", _p(d_x), _p(y), N, H, W, C, L.TF32, None)
    np.testing.assert_array_equal(y.cpu().numpy(), O.np_maxpool2x2(x))
    dp = _rand((N, H // 2, W // 2, C), rng)
    skip = _rand((N, H, W, C), rng)
    dz = torch.tensor(skip, device=dev)
    L.call("unetb200_maxpool2x2_bwd_dt", _p(d_x), _p(torch.tensor(dp, device=dev)), _p(dz), _p(dz), N, H, W, C, L.TF32,
           None)
    want = skip + O.np_maxpool2x2_grad(x, dp) * (x > 0)
    np.testing.assert_array_equal(dz.cpu().numpy(), _tf32(want))
    up = torch.empty((N, 2 * H, 2 * W, C), dtype=torch.float32, device=dev)
    L.call("unetb200_upsample2x2_fwd_dt", _p(d_x), _p(up), N, H, W, C, L.TF32, None)
    np.testing.assert_array_equal(up.cpu().numpy(), O.np_upsample2x2(x))
    dy = _rand((N, 2 * H, 2 * W, C), rng)
    dx = torch.empty_like(d_x)
    L.call("unetb200_upsample2x2_bwd_dt", _p(torch.tensor(dy, device=dev)), _p(d_x), _p(dx), N, H, W, C, 1, L.TF32, None)
    s = dy.reshape(N, H, 2, W, 2, C)
    want = ((s[:, :, 0, :, 0] + s[:, :, 0, :, 1]) + (s[:, :, 1, :, 0] + s[:, :, 1, :, 1])) * (x > 0)
    np.testing.assert_array_equal(dx.cpu().numpy(), _tf32(want))
    rc = L.lib().unetb200_maxpool2x2_fwd_dt(_p(d_x), _p(y), N, H, W, C, 7, None)
    assert rc < 0 and "dtype" in L.last_error()
    torch.cuda.synchronize()


def _run_model(cuda_device, N, H, W, in_ch, ups, dropout):
    from unet_distributed_b200.plan import UNetPlan
    imgs, msks = synthetic_batch(N, H, W, in_ch)
    weights = O.init_weights(1234, ups, in_ch, 1)
    rng = np.random.default_rng(5)
    for k in weights:
        if k.endswith("/bias"):
            weights[k] = rng.uniform(-0.05, 0.05, weights[k].shape).astype(np.float32)
    plan = UNetPlan(N, H, W, in_ch, 1, ups, dropout, precision="tf32")
    assert plan.precision == "tf32" and plan.tensor("a3").dtype == torch.float32
    plan.set_weights(weights)
    d_imgs, d_msks = torch.tensor(imgs, device=cuda_device), torch.tensor(msks, device=cuda_device)
    plan.forward(d_imgs, d_msks, seed=DROPOUT_SEED)
    plan.backward(d_imgs, d_msks)
    torch.cuda.synchronize()
    masks = plan_dropout_masks(plan, DROPOUT_SEED) if dropout > 0 else None
    names = act_names(ups)
    keep_scale = 1.0 / (1.0 - dropout) if dropout > 0 else 1.0

    def masked(o, oname):
        g = o["act_grads"][oname]
        if oname.startswith("transConv"):
            return g
        g = g * (o["acts"][oname] > 0)
        return g * keep_scale if oname in ("conv3a", "conv4a") else g

    gpu_acts = {o: plan.tensor(p).cpu().numpy() for p, o in names.items()}
    gpu_dacts = {o: plan.tensor("g" + p).cpu().numpy() for p, o in names.items()}
    preds, met, grads = plan.preds.cpu().numpy(), plan.metrics.cpu().numpy(), plan.get_grads()
    ref = O.loss_and_grads(weights, imgs, msks, ups, masks, dropout, torch.float32, keep_act_grads=True)
    twin = O.loss_and_grads(weights, imgs, msks, ups, masks, dropout, torch.float32, emulate_bf16="tf32",
                            keep_act_grads=True)
    emu = O.loss_and_grads(weights, imgs, msks, ups, None, dropout, torch.float32, emulate_bf16="tf32",
                           keep_act_grads=True, forced_acts=gpu_acts)
    rows = []   # (name, vs fp32 oracle, vs oracle at the GPU's forward state, twin vs fp32 oracle)
    for o, got in gpu_acts.items():
        rows.append((f"act {o}", rel_err(got, ref["acts"][o]), rel_err(got, emu["acts"][o]),
                     rel_err(twin["acts"][o], ref["acts"][o])))
    rows.append(("act preds", rel_err(preds, ref["preds"]), rel_err(preds, emu["preds"]),
                 rel_err(twin["preds"], ref["preds"])))
    for o, got in gpu_dacts.items():
        rows.append((f"dact {o}", rel_err(got, masked(ref, o)), rel_err(got, masked(emu, o)),
                     rel_err(masked(twin, o), masked(ref, o))))
    for k in ref["grads"]:
        rows.append((f"grad {k}", rel_err(grads[k], ref["grads"][k]), rel_err(grads[k], emu["grads"][k]),
                     rel_err(twin["grads"][k], ref["grads"][k])))
    out = os.environ.get("UNETB200_TEST_REPORT")
    if out:
        with open(out, "a") as f:
            f.write(f"# TF32 N{N} {H}x{W} cin{in_ch} ups={ups} dropout={dropout}   [vs fp32 oracle | vs oracle "
                    f"backward at the GPU's forward state | oracle tf32 twin vs fp32 oracle]\n")
            for name, a, b, t in rows:
                f.write(f"{name:<28s} {a:.3e} {b:.3e} {t:.3e}\n")
    for i, key in enumerate(("loss", "dice", "sensitivity", "specificity")):
        assert abs(met[i] - ref[key]) < 1e-3, (key, met[i], ref[key])
    bad = []
    for name, a, b, t in rows:
        if not b < TOL_TF32:
            bad.append(f"{name}: {b:.3e} vs the oracle at the same forward state")
        if name.startswith("act ") and not a < max(TOL_TF32, TWIN_FACTOR * t):
            bad.append(f"{name}: {a:.3e} vs fp32 oracle (tf32 twin {t:.3e})")
        if not name.startswith("act ") and not a <= TWIN_FACTOR * t + TOL_TF32:
            bad.append(f"{name}: {a:.3e} vs fp32 oracle free-running (tf32 twin {t:.3e})")
    assert not bad, "; ".join(bad)


def test_model_tf32_tconv(cuda_device):
    _run_model(cuda_device, 2, 64, 64, 1, False, 0.2)


def test_model_tf32_upsampling(cuda_device):
    _run_model(cuda_device, 2, 64, 64, 1, True, 0.0)


def test_model_tf32_four_modalities_240(cuda_device):
    _run_model(cuda_device, 1, 240, 240, 4, False, 0.0)


def test_trainer_tf32_first_steps(cuda_device):
    """Three optimizer steps in the tf32 mode track the fp32 oracle's loss and dice to 1e-3 (bf16: 1e-2)."""
    from unet_distributed_b200.trainer import Trainer
    N, H, W, steps = 4, 64, 64, 3
    weights = O.init_weights(1234)
    batches = [synthetic_batch(N, H, W, seed=816 + i) for i in range(4)]
    tr = Trainer(N, H, W, dropout=0.2, learning_rate=0.0001, weights=weights, dropout_seed=DROPOUT_SEED,
                 precision="tf32", use_cuda_graph=True)
    masks = [plan_dropout_masks(tr.plan, DROPOUT_SEED + s) for s in range(steps)]
    got = []
    for s in range(steps):
        a, b = batches[s % 4]
        got.append(tr.train_step(torch.tensor(a, device=cuda_device), torch.tensor(b, device=cuda_device)).clone())
    got = torch.stack(got).cpu().numpy()
    _, hist = O.train(weights, batches, steps, False, dropout_masks_fn=lambda s: masks[s], dropout=0.2, lr0=0.0001)
    ref = np.array(hist)
    assert np.abs(got[:, 0] - ref[:, 0]).max() < 2e-3, (got[:, 0], ref[:, 0])
    assert np.abs(got[:, 1] - ref[:, 1]).max() < 2e-3, (got[:, 1], ref[:, 1])
