"""Drop-in surface of the reference's model.py on B200.

    define_model(input_tensor, use_upsampling=False, n_cl_out=1, dropout=0.2, print_summary=False)   model.py:36
    dice_coef / dice_coef_loss / sensitivity / specificity (y_true, y_pred, smooth=1.)              model.py:4,13,133,139

Tensors are CUDA float32 NHWC (`channels_last`, model.py:25-34).  PyTorch owns the tensors and wires autograd; every
FLOP runs in libunetb200 (hand-written sm_100a kernels) through the C ABI.  Like the TF graph, the network's variables
live in a module-level registry: `get_model()` returns the UNet the last `define_model` call built, whose `.params`
leaf tensor holds all variables (TF names/layouts via `.plan.variables`)."""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .init import initial_weights
from .plan import UNetPlan, _ptr, _stream

CHANNEL_LAST = True
concat_axis = -1
data_format = "channels_last"

_MODELS = {}
_LAST = None


class UNet:
    """The variables and launch plan behind define_model for one input shape."""

    def __init__(self, N, H, W, in_channels, n_cl_out, use_upsampling, dropout, weight_seed=1234):
        self.plan = UNetPlan(N, H, W, in_channels, n_cl_out, use_upsampling, dropout)
        self.plan.set_weights(initial_weights(self.plan.variables, weight_seed))
        self.params = self.plan.params.requires_grad_(True)        # leaf holding every trainable variable
        self.dropout_seed = 816
        self.calls = 0

    def __call__(self, input_tensor):
        return _UNetFn.apply(input_tensor, self.params, self)

    def summary(self):
        rows = [f"{'variable':<22}{'shape':<22}{'params':>10}"]
        for name, (_, shape) in self.plan.variables.items():
            n = 1
            for s in shape:
                n *= s
            rows.append(f"{name:<22}{str(shape):<22}{n:>10}")
        rows.append(f"Total params: {self.plan.num_params:,}")
        return "\n".join(rows)


class _UNetFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, imgs, params, net):
        plan = net.plan
        if params.data_ptr() != plan.params.data_ptr():
            raise _lib.UnetB200Error("define_model: the parameter tensor must be the plan's flat buffer")
        imgs = imgs.contiguous()
        plan.pack_weights()                      # variables may have been updated by an optimizer since last call
        net.calls += 1
        plan.forward(imgs, None, seed=net.dropout_seed + net.calls)
        ctx.net, ctx.imgs = net, imgs
        return plan.preds.clone()

    @staticmethod
    def backward(ctx, dpreds):
        plan = ctx.net.plan
        plan.backward(ctx.imgs, None, dpreds=dpreds.contiguous(), grad_scale=1.0)
        return None, plan.grads.clone(), None


def define_model(input_tensor, use_upsampling=False, n_cl_out=1, dropout=0.2, print_summary=False):
    """model.py:36-130.  Builds the UNet for input_tensor's static shape on first use (variables are shared by later
    calls with the same configuration, as re-running the TF graph would) and returns the sigmoid prediction tensor
    [N,H,W,n_cl_out], differentiable with respect to the network's variables.  The learning phase is always
    "train": dropout stays active (model.py:40)."""
    global _LAST
    if not (torch.is_tensor(input_tensor) and input_tensor.is_cuda and input_tensor.dim() == 4):
        raise ValueError("define_model expects a CUDA float32 NHWC tensor")
    N, H, W, C = (int(s) for s in input_tensor.shape)
    key = (input_tensor.device.index, N, H, W, C, bool(use_upsampling), int(n_cl_out), float(dropout))
    net = _MODELS.get(key)
    if net is None:
        with torch.cuda.device(input_tensor.device):
            net = UNet(N, H, W, C, n_cl_out, use_upsampling, dropout)
        _MODELS[key] = net
    _LAST = net
    if print_summary:
        print(net.summary())
    return net(input_tensor.float())


def get_model():
    """The UNet built by the most recent define_model call."""
    if _LAST is None:
        raise _lib.UnetB200Error("define_model has not been called")
    return _LAST


def reset_models():
    global _LAST
    _MODELS.clear()
    _LAST = None


class _DiceMetricFn(torch.autograd.Function):
    """One of the four batch-global dice metrics; the three reductions and the closed-form gradient are CUDA kernels."""

    @staticmethod
    def forward(ctx, y_true, y_pred, smooth, which):
        y_true = y_true.contiguous().float()
        y_pred = y_pred.contiguous().float()
        if not (y_true.is_cuda and y_pred.is_cuda) or y_true.numel() != y_pred.numel():
            raise ValueError("dice metrics expect CUDA tensors of equal size")
        L = _lib.lib()
        dev = y_pred.device
        sums = torch.empty(3, dtype=torch.float32, device=dev)
        out4 = torch.empty(4, dtype=torch.float32, device=dev)
        ws = torch.empty(int(L.unetb200_dice_sums_ws_bytes()), dtype=torch.uint8, device=dev)
        _lib.call("unetb200_dice_sums", _ptr(y_true), _ptr(y_pred), ctypes.c_longlong(y_pred.numel()), _ptr(sums),
                  _ptr(ws), _stream())
        _lib.call("unetb200_dice_finalize", _ptr(sums), float(smooth), _ptr(out4), _stream())
        ctx.save_for_backward(y_true, sums)
        ctx.smooth, ctx.which, ctx.shape = float(smooth), int(which), y_pred.shape
        return out4[which].clone()

    @staticmethod
    def backward(ctx, grad_out):
        y_true, sums = ctx.saved_tensors
        dp = torch.empty(ctx.shape, dtype=torch.float32, device=y_true.device)
        up = grad_out.contiguous().float().reshape(1)
        _lib.call("unetb200_dice_grad", _ptr(y_true), _ptr(sums), ctx.smooth, ctx.which, _ptr(up), _ptr(dp),
                  ctypes.c_longlong(dp.numel()), _stream())
        return None, dp, None, None


def dice_coef_loss(y_true, y_pred, smooth=1.0):
    """model.py:13-21: -log(2*sum(y*p)+smooth) + log(sum(y)+sum(p)+smooth), sums over the whole batch tensor."""
    return _DiceMetricFn.apply(y_true, y_pred, smooth, 0)


def dice_coef(y_true, y_pred, smooth=1.0):
    """model.py:4-10."""
    return _DiceMetricFn.apply(y_true, y_pred, smooth, 1)


def sensitivity(y_true, y_pred, smooth=1.0):
    """model.py:133-137."""
    return _DiceMetricFn.apply(y_true, y_pred, smooth, 2)


def specificity(y_true, y_pred, smooth=1.0):
    """model.py:139-143 (the reference's formula is precision; kept as is)."""
    return _DiceMetricFn.apply(y_true, y_pred, smooth, 3)
