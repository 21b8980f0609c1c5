"""Whole-network parity on the GPU: the C-ABI plan against the CPU oracle on identical weights and inputs.

Two oracles, two bars (all errors are ||got - ref||_F / ||ref||_F per tensor):

* fp32 oracle ("the reference's precision"): forward activations, predictions within 2e-2; loss / dice / sensitivity /
  specificity within 1e-2 absolute.  Free-running gradients are bounded PER TENSOR by the oracle's own bf16-storage
  twin (the same fp32 oracle with activations and GEMM weights rounded to bf16, run on the same inputs and dropout
  masks): err(GPU, fp32 oracle) <= TWIN_FACTOR * err(twin, fp32 oracle) + TWIN_FLOOR.  Storing activations in bf16
  flips a fraction of ReLU masks and 2x2 max-pool winners (near-ties) and every flip re-routes a whole gradient
  element, so ANY bf16-storage implementation deviates from fp32 by 5-20 % on the deep layers of this network -- which
  is why this cannot be a flat 2e-2 bar -- but it must not deviate by more than the reference's own arithmetic does
  when it is given the same storage format.  TWIN_FLOOR (= the 2e-2 bf16 budget) covers what the twin does not model:
  gradient tensors stored in bf16, a different accumulation order, and -- on tensors of a few thousand elements -- the
  noise of WHICH near-zero gates flip (measured GPU/twin ratios: <= 1.5 on every tensor of every case except the
  3 x 32 x 32 two-class case, 1.8-2.3 on conv9b there; profiles/r02_parity_report.txt).
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
TWIN_FACTOR = 1.5        # free-running gradient error allowed relative to the oracle's own bf16-storage twin
TWIN_FLOOR = TOL_BF16      # the bf16 budget itself: what the GPU's own arithmetic may add on top of the storage effect


def _weights_with_biases(ups, in_ch, n_cl):
    weights = O.init_weights(1234, ups, in_ch, n_cl)
    rng = np.random.default_rng(5)
    for k in weights:                      # non-zero biases so that the bias paths are exercised
        if k.endswith("/bias"):
            weights[k] = rng.uniform(-0.05, 0.05, weights[k].shape).astype(np.float32)
    return weights


def _run_case(cuda_device, N, H, W, in_ch, ups, dropout, n_cl=1):
    from unet_distributed_b200.plan import UNetPlan
    imgs, msks = synthetic_batch(N, H, W, in_ch)
    if n_cl > 1:
        msks = np.concatenate([msks] + [np.roll(msks, 3 * k, axis=1) for k in range(1, n_cl)], axis=3)
    weights = _weights_with_biases(ups, in_ch, n_cl)
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
    names = act_names(ups)
    keep_scale = 1.0 / (1.0 - dropout) if dropout > 0 else 1.0

    def masked(o, oname):   # the plan stores d/d(pre-activation): oracle's d/d(output) times ReLU (and dropout) mask
        g = o["act_grads"][oname]
        if oname.startswith("transConv"):
            return g
        g = g * (o["acts"][oname] > 0)
        return g * keep_scale if oname in ("conv3a", "conv4a") else g

    gpu_acts = {oname: plan.tensor(pname).float().cpu().numpy() for pname, oname in names.items()}
    gpu_dacts = {oname: plan.tensor("g" + pname).float().cpu().numpy() for pname, oname in names.items()}
    preds = plan.preds.cpu().numpy()
    met = plan.metrics.cpu().numpy()
    grads = plan.get_grads()

    # one oracle pass at a time (the bs128 cases hold ~3 GB of fp32 activations per pass)
    ref = O.loss_and_grads(weights, imgs, msks, ups, masks, dropout, torch.float32, keep_act_grads=True)
    for i, key in enumerate(("loss", "dice", "sensitivity", "specificity")):
        assert abs(met[i] - ref[key]) < 1e-2, (key, met[i], ref[key])
    e32 = {}
    for oname, got in gpu_acts.items():
        e32[f"act {oname}"] = rel_err(got, ref["acts"][oname])
    e32["act preds"] = rel_err(preds, ref["preds"])
    ref_dacts = {oname: masked(ref, oname) for oname in gpu_dacts}
    for oname, got in gpu_dacts.items():
        e32[f"dact {oname}"] = rel_err(got, ref_dacts[oname])
    for k in ref["grads"]:
        e32[f"grad {k}"] = rel_err(grads[k], ref["grads"][k])
    ref_grads = ref["grads"]
    del ref

    # the oracle's bf16-storage twin, free-running on the same inputs and masks, against the fp32 oracle
    twin = O.loss_and_grads(weights, imgs, msks, ups, masks, dropout, torch.float32, emulate_bf16=True,
                            keep_act_grads=True)
    etw = {}
    for oname in gpu_dacts:
        etw[f"dact {oname}"] = rel_err(masked(twin, oname), ref_dacts[oname])
    for k in ref_grads:
        etw[f"grad {k}"] = rel_err(twin["grads"][k], ref_grads[k])
    del twin, ref_dacts

    # the oracle's backward at the GPU's forward state
    emu = O.loss_and_grads(weights, imgs, msks, ups, None, dropout, torch.float32, emulate_bf16=True,
                           keep_act_grads=True, forced_acts=gpu_acts)
    eb = {}
    for oname, got in gpu_acts.items():
        eb[f"act {oname}"] = rel_err(got, emu["acts"][oname])
    eb["act preds"] = rel_err(preds, emu["preds"])
    for oname, got in gpu_dacts.items():
        eb[f"dact {oname}"] = rel_err(got, masked(emu, oname))
    for k in ref_grads:
        eb[f"grad {k}"] = rel_err(grads[k], emu["grads"][k])
    del emu

    report = [(name, e32[name], eb[name], etw.get(name)) for name in e32]
    out = os.environ.get("UNETB200_TEST_REPORT")
    if out:
        with open(out, "a") as f:
            f.write(f"# N{N} {H}x{W} cin{in_ch} ups={ups} dropout={dropout} ncl={n_cl}   "
                    f"[vs fp32 oracle | vs oracle backward at the GPU's forward state | oracle bf16 twin vs fp32 oracle]\n")
            for name, a, b, t in report:
                f.write(f"{name:<28s} {a:.3e} {b:.3e} {'' if t is None else format(t, '.3e')}\n")
    bad = []
    for name, a, b, t in report:
        if name.startswith("act ") and not a < TOL_BF16:
            bad.append(f"{name}: {a:.3e} vs fp32 oracle")
        if not b < TOL_BF16:
            bad.append(f"{name}: {b:.3e} vs the oracle at the same forward state")
        if t is not None and not a <= TWIN_FACTOR * t + TWIN_FLOOR:
            bad.append(f"{name}: {a:.3e} vs fp32 oracle, the oracle's own bf16 twin is at {t:.3e}")
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


@pytest.mark.parametrize("ups", [False, True])
def test_benchmarked_shape_bs128(cuda_device, ups):
    """The shape bench.py times (BASELINE configs[0] / [1]: 128x128x1, bs128 per GPU): kernel variants, split-K factors
    and wave sizing depend on the batch, so parity is asserted at this batch too, not only at N <= 4."""
    _run_case(cuda_device, 128, 128, 128, 1, ups, 0.2)


def test_bs512_one_step(cuda_device):
    """BASELINE configs[4] shape (bs512 per GPU): loss / dice against the fp32 oracle, every weight gradient against
    the oracle's backward at the GPU's forward state (activation gradients are not pulled back: 23 GB)."""
    from unet_distributed_b200.plan import UNetPlan
    N, H, W, ups, dropout = 512, 128, 128, True, 0.2
    imgs, msks = synthetic_batch(N, H, W, 1)
    weights = _weights_with_biases(ups, 1, 1)
    plan = UNetPlan(N, H, W, 1, 1, ups, dropout)
    plan.set_weights(weights)
    d_imgs, d_msks = torch.tensor(imgs, device=cuda_device), torch.tensor(msks, device=cuda_device)
    plan.forward(d_imgs, d_msks, seed=DROPOUT_SEED)
    plan.backward(d_imgs, d_msks)
    torch.cuda.synchronize()
    masks = plan_dropout_masks(plan, DROPOUT_SEED)
    met = plan.metrics.cpu().numpy()
    grads = plan.get_grads()
    names = act_names(ups)
    ref = O.loss_and_grads(weights, imgs, msks, ups, masks, dropout, torch.float32)
    for i, key in enumerate(("loss", "dice", "sensitivity", "specificity")):
        assert abs(met[i] - ref[key]) < 1e-2, (key, met[i], ref[key])
    assert rel_err(plan.preds.cpu().numpy(), ref["preds"]) < TOL_BF16
    assert rel_err(plan.tensor("c5").float().cpu().numpy(), ref["acts"]["conv5b"]) < TOL_BF16
    del ref
    forced = {oname: plan.tensor(pname).float().cpu().numpy() for pname, oname in names.items()}
    emu = O.loss_and_grads(weights, imgs, msks, ups, None, dropout, torch.float32, emulate_bf16=True,
                           forced_acts=forced)
    bad = [f"{k}: {rel_err(grads[k], g):.3e}" for k, g in emu["grads"].items() if not rel_err(grads[k], g) < TOL_BF16]
    assert not bad, "; ".join(bad)


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
