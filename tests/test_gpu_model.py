"""Whole-network parity on the GPU: the C-ABI plan against the CPU oracle on identical weights and inputs.

Two oracles, two bars (all errors are ||got - ref||_F / ||ref||_F per tensor):

* fp32 oracle ("the reference's precision"): forward activations, predictions within 2e-2; loss / dice / sensitivity /
  specificity within 1e-2 absolute.  Gradients are compared too but only bounded loosely (GRAD_VS_FP32): storing
  activations in bf16 flips a fraction of ReLU masks and 2x2 max-pool winners (near-ties), and every flip re-routes a
  whole gradient element, so ANY bf16-storage implementation deviates from fp32 by 10-20 % on the deep layers of
  this network.  tests/test_oracle.py::test_bf16_storage_gradient_deviation shows the oracle does exactly that to
  itself, which is why this cannot be a 2e-2 bar.
* the oracle's backward evaluated AT THE GPU'S FORWARD STATE (oracle Net.forward(forced_acts=...): layer outputs,
  ReLU/dropout gates and pool routing are the ones the GPU stored; gradients flow through the oracle's fp32
  operators with the bf16-rounded GEMM weights): every activation gradient and every weight gradient within 2e-2.
  Together with the forward bar this is the chain rule: same forward (to 2e-2), same backward given the forward.

Dropout is compared with the exact keep masks the kernels use (regenerated through unetb200_dropout_mask and
injected into the oracle)."""
import os

import numpy as np
import pytest
import torch

from util import O, DROPOUT_SEED, act_names, plan_dropout_masks, rel_err, synthetic_batch

pytestmark = pytest.mark.gpu

TOL_BF16 = 2e-2          # north_star bf16 tolerance
GRAD_VS_FP32 = 0.35      # sanity bound only, see the module docstring


def _run_case(cuda_device, N, H, W, in_ch, ups, dropout, n_cl=1):
    from unet_distributed_b200.plan import UNetPlan
    imgs, msks = synthetic_batch(N, H, W, in_ch)
    if n_cl > 1:
        msks = np.concatenate([msks] + [np.roll(msks, 3 * k, axis=1) for k in range(1, n_cl)], axis=3)
    weights = O.init_weights(1234, ups, in_ch, n_cl)
    rng = np.random.default_rng(5)
    for k in weights:                      # non-zero biases so that the bias paths are exercised
        if k.endswith("/bias"):
            weights[k] = rng.uniform(-0.05, 0.05, weights[k].shape).astype(np.float32)
    plan = UNetPlan(N, H, W, in_ch, n_cl, ups, dropout)
    assert list(plan.variables) == list(weights), "variable order / names differ from the TF creation order"
    for k, (_, shape) in plan.variables.items():
        assert tuple(shape) == tuple(weights[k].shape), k
    plan.set_weights(weights)
    d_imgs = torch.tensor(imgs, device=cuda_device)
    d_msks = torch.tensor(msks, device=cuda_device)
    plan.forward(d_imgs, d_msks, seed=DROPOUT_SEED)
    plan.backward(d_imgs, d_msks)
    torch.cuda.synchronize()
    masks = plan_dropout_masks(plan, DROPOUT_SEED) if dropout > 0 else None
    ref = O.loss_and_grads(weights, imgs, msks, ups, masks, dropout, torch.float32, keep_act_grads=True)
    names = act_names(ups)
    forced = {oname: plan.tensor(pname).float().cpu().numpy() for pname, oname in names.items()}
    emu = O.loss_and_grads(weights, imgs, msks, ups, None, dropout, torch.float32, emulate_bf16=True,
                           keep_act_grads=True, forced_acts=forced)

    report = []          # (what, err vs fp32 oracle, err vs oracle evaluated at the GPU's forward state)
    for pname, oname in names.items():
        got = plan.tensor(pname).float().cpu().numpy()
        report.append((f"act {oname}", rel_err(got, ref["acts"][oname]), rel_err(got, emu["acts"][oname])))
    preds = plan.preds.cpu().numpy()
    report.append(("act preds", rel_err(preds, ref["preds"]), rel_err(preds, emu["preds"])))
    met = plan.metrics.cpu().numpy()
    for i, key in enumerate(("loss", "dice", "sensitivity", "specificity")):
        assert abs(met[i] - ref[key]) < 1e-2, (key, met[i], ref[key])

    keep_scale = 1.0 / (1.0 - dropout) if dropout > 0 else 1.0

    def masked(o, oname):   # the plan stores d/d(pre-activation): oracle's d/d(output) times ReLU (and dropout) mask
        g = o["act_grads"][oname]
        if oname.startswith("transConv"):
            return g
        g = g * (o["acts"][oname] > 0)
        return g * keep_scale if oname in ("conv3a", "conv4a") else g

    for pname, oname in names.items():
        got = plan.tensor("g" + pname).float().cpu().numpy()
        report.append((f"dact {oname}", rel_err(got, masked(ref, oname)), rel_err(got, masked(emu, oname))))
    grads = plan.get_grads()
    for k in ref["grads"]:
        report.append((f"grad {k}", rel_err(grads[k], ref["grads"][k]), rel_err(grads[k], emu["grads"][k])))

    out = os.environ.get("UNETB200_TEST_REPORT")
    if out:
        with open(out, "a") as f:
            f.write(f"# N{N} {H}x{W} cin{in_ch} ups={ups} dropout={dropout} ncl={n_cl}   "
                    f"[vs fp32 oracle | vs oracle backward at the GPU's forward state]\n")
            for name, e32, eb in report:
                f.write(f"{name:<28s} {e32:.3e} {eb:.3e}\n")
    bad = []
    for name, e32, eb in report:
        if name.startswith("act ") and not e32 < TOL_BF16:
            bad.append(f"{name}: {e32:.3e} vs fp32 oracle")
        if not eb < TOL_BF16:
            bad.append(f"{name}: {eb:.3e} vs the oracle at the same forward state")
        if not e32 < GRAD_VS_FP32:
            bad.append(f"{name}: {e32:.3e} vs fp32 oracle (sanity bound)")
    assert not bad, "; ".join(bad)
    return report


def test_tconv_no_dropout(cuda_device):
    _run_case(cuda_device, 4, 64, 64, 1, False, 0.0)


def test_tconv_dropout(cuda_device):
    _run_case(cuda_device, 4, 64, 64, 1, False, 0.2)


def test_upsampling_dropout(cuda_device):
    _run_case(cuda_device, 4, 64, 64, 1, True, 0.2)


def test_full_resolution_128(cuda_device):
    _run_case(cuda_device, 2, 128, 128, 1, False, 0.2)


def test_odd_batch_and_two_classes(cuda_device):
    _run_case(cuda_device, 3, 32, 32, 1, False, 0.0, n_cl=2)


def test_four_modalities_240(cuda_device):
    """Config 4 of BASELINE.json: 240x240x4 input (15x15 at the bottom level: ragged tiles everywhere)."""
    _run_case(cuda_device, 1, 240, 240, 4, False, 0.2)


def test_four_modalities_240_upsampling(cuda_device):
    _run_case(cuda_device, 1, 240, 240, 4, True, 0.0)


def test_deterministic_repeat(cuda_device):
    """Two identical steps give bit-identical gradients (every reduction runs in a fixed order, no float atomics)."""
    from unet_distributed_b200.plan import UNetPlan
    imgs, msks = synthetic_batch(2, 64, 64)
    plan = UNetPlan(2, 64, 64, 1, 1, False, 0.2)
    plan.set_weights(O.init_weights(7))
    d_imgs, d_msks = torch.tensor(imgs, device=cuda_device), torch.tensor(msks, device=cuda_device)
    outs = []
    for _ in range(2):
        plan.grads.zero_()
        plan.forward(d_imgs, d_msks, seed=3)
        plan.backward(d_imgs, d_msks)
        torch.cuda.synchronize()
        outs.append((plan.grads.clone(), plan.metrics.clone()))
    assert torch.equal(outs[0][1], outs[1][1])
    assert torch.equal(outs[0][0], outs[1][0])


@pytest.mark.parametrize("ups", [False, True])
def test_side_stream_on_off_identical(cuda_device, ups):
    """Weight gradients on the plan's second stream (default) and in program order give bit-identical gradient buffers:
    the fork/join only changes when kernels run, never what they read."""
    from unet_distributed_b200.plan import UNetPlan
    imgs, msks = synthetic_batch(2, 64, 64)
    plan = UNetPlan(2, 64, 64, 1, 1, ups, 0.2)
    plan.set_weights(O.init_weights(11, use_upsampling=ups))
    d_imgs, d_msks = torch.tensor(imgs, device=cuda_device), torch.tensor(msks, device=cuda_device)
    outs = []
    for on in (True, False, True):
        old = plan.side_stream(on)
        assert isinstance(old, bool)
        plan.grads.zero_()   # (the flat buffer has never-written padding between variables)
        plan.forward(d_imgs, d_msks, seed=5)
        plan.backward(d_imgs, d_msks)
        torch.cuda.synchronize()
        assert torch.isfinite(plan.grads).all() and float(plan.grads.abs().sum()) > 0.0
        outs.append(plan.grads.clone())
    assert torch.equal(outs[0], outs[1])
    assert torch.equal(outs[0], outs[2])
