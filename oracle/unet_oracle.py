"""CPU oracle for the UNet training hot path of dmsuehir/unet_distributed.

TEST INFRASTRUCTURE ONLY.  Nothing in the product (unet_distributed_b200/) imports this module; only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline / --impl reference legs may.

PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path, and it cannot run here
(Python 2.7 + Intel TensorFlow 1.4.0, neither installable offline).  The arithmetic lives in that un-vendored
dependency (TensorFlow 1.4.0 / Keras 2.0.8, named at README.md:22,32 and Distributed.sh:21), so this file restates
the published semantics of the ops the reference's call sites invoke:

    model.py:36-130   define_model  (layer graph, initialisers, [up, skip] concat order, dropout always on)
    model.py:4-21     dice_coef / dice_coef_loss        model.py:133-143  sensitivity / specificity
    test_dist.py:228-248  AdamOptimizer + exponential_decay(staircase=False)
    test_dist.py:251-262  SyncReplicasOptimizer (mean of per-replica gradients), :388-394 shard split

Two independent restatements live here so that one can check the other (tests/test_oracle.py):
  * torch-CPU functional ops (fp32 = "the reference's precision", fp64 = truth for tolerances), with autograd;
  * direct NumPy loops for tiny shapes that pin the TF conventions (HWIO kernels, Conv2DTranspose [kh,kw,out,in]
    layout, first-max pooling gradient, nearest upsampling).
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

# ------------------------------------------------------------------------------------------------ layer table


def layer_table(use_upsampling=False, in_channels=1, n_cl_out=1):
    """Ordered (name, kind, Cin, Cout) rows in graph order -- model.py:51-120.

    kind: 'conv' = Conv2D 3x3 same relu he_uniform; 'tconv' = Conv2DTranspose 2x2 s2 glorot_uniform no activation;
    'mask' = Conv2D 1x1 sigmoid glorot_uniform.
    """
    up = use_upsampling
    rows = [
        ("conv1a", "conv", in_channels, 32), ("conv1b", "conv", 32, 32),
        ("conv2a", "conv", 32, 64), ("conv2b", "conv", 64, 64),
        ("conv3a", "conv", 64, 128), ("conv3b", "conv", 128, 128),
        ("conv4a", "conv", 128, 256), ("conv4b", "conv", 256, 256),
        ("conv5a", "conv", 256, 512), ("conv5b", "conv", 512, 256 if up else 512),   # model.py:75 / :78
    ]
    if not up:
        rows.append(("transConv6", "tconv", 512, 256))                               # model.py:79-80
    rows += [("conv6a", "conv", 512, 256), ("conv6b", "conv", 256, 128 if up else 256)]  # model.py:82, :86/:89
    if not up:
        rows.append(("transConv7", "tconv", 256, 128))
    rows += [("conv7a", "conv", 256, 128), ("conv7b", "conv", 128, 64 if up else 128)]
    if not up:
        rows.append(("transConv8", "tconv", 128, 64))
    rows += [("conv8a", "conv", 128, 64), ("conv8b", "conv", 64, 32 if up else 64)]
    if not up:
        rows.append(("transConv9", "tconv", 64, 32))
    rows += [("conv9a", "conv", 64, 32), ("conv9b", "conv", 32, 32), ("Mask", "mask", 32, n_cl_out)]
    return rows


def variable_shapes(use_upsampling=False, in_channels=1, n_cl_out=1):
    """TF variable name -> shape, in creation order (kernel then bias per layer)."""
    out = OrderedDict()
    for name, kind, cin, cout in layer_table(use_upsampling, in_channels, n_cl_out):
        if kind == "conv":
            out[name + "/kernel"] = (3, 3, cin, cout)          # HWIO [TF-1.4 Conv2D]
        elif kind == "tconv":
            out[name + "/kernel"] = (2, 2, cout, cin)          # [kh, kw, out, in] [TF-1.4 Conv2DTranspose]
        else:
            out[name + "/kernel"] = (1, 1, cin, cout)
        out[name + "/bias"] = (cout,)
    return out


def init_weights(seed=1234, use_upsampling=False, in_channels=1, n_cl_out=1):
    """he_uniform for the 3x3 convs (model.py:47-49), glorot_uniform for tconv / Mask (Keras default), zero biases."""
    rng = np.random.default_rng(seed)
    w = OrderedDict()
    for name, kind, cin, cout in layer_table(use_upsampling, in_channels, n_cl_out):
        shape = variable_shapes(use_upsampling, in_channels, n_cl_out)[name + "/kernel"]
        if kind == "conv":
            limit = math.sqrt(6.0 / (9 * cin))                 # VarianceScaling(2, fan_in, uniform)
        elif kind == "tconv":
            limit = math.sqrt(6.0 / (4 * cout + 4 * cin))      # fans of a [2,2,out,in] kernel
        else:
            limit = math.sqrt(6.0 / (cin + cout))
        w[name + "/kernel"] = rng.uniform(-limit, limit, size=shape).astype(np.float32)
        w[name + "/bias"] = np.zeros((cout,), np.float32)
    return w


def num_params(weights):
    return int(sum(int(np.prod(v.shape)) for v in weights.values()))


# ------------------------------------------------------------------------------------------------ torch restatement


def _bf16_round(t):
    return t.to(torch.bfloat16).to(t.dtype)


def _tf32_round(t):
    """Nearest tf32 value (10-bit mantissa, ties away from zero -- PTX cvt.rna.tf32.f32) of an fp32/fp64 tensor."""
    bits = t.detach().to(torch.float32).contiguous().view(torch.int32)
    bits = (bits + 0x1000) & ~0x1FFF
    return bits.view(torch.float32).to(t.dtype)


_ROUNDERS = {True: _bf16_round, "bf16": _bf16_round, "tf32": _tf32_round}


def _conv3x3(x, k_hwio, b):
    # NHWC semantics expressed in NCHW storage: kernel HWIO -> OIHW, zero pad 1 ('same')
    return F.conv2d(x, k_hwio.permute(3, 2, 0, 1), b, padding=1)


def _tconv2x2(x, k, b):
    # TF kernel [a, b, out, in] -> torch conv_transpose2d weight [in, out, a, b]
    return F.conv_transpose2d(x, k.permute(3, 2, 0, 1), b, stride=2)


class Net:
    """Functional UNet over a dict of torch leaf tensors named like the TF variables."""

    def __init__(self, weights, use_upsampling=False, dtype=torch.float32, emulate_bf16=False):
        self.up = use_upsampling
        self.dtype = dtype
        self.emulate_bf16 = emulate_bf16
        self.params = OrderedDict(
            (k, torch.tensor(np.asarray(v), dtype=dtype, requires_grad=True)) for k, v in weights.items())

    def _w(self, name):
        k = self.params[name + "/kernel"]
        if self.emulate_bf16 and not name.startswith(("conv1a", "Mask")):
            rnd = _ROUNDERS[self.emulate_bf16]                 # True / "bf16", or "tf32"
            k = k + (rnd(k) - k).detach()                      # reduced-precision GEMM weights, straight-through gradient
        return k, self.params[name + "/bias"]

    def _act(self, t):
        if self.emulate_bf16:
            t = t + (_ROUNDERS[self.emulate_bf16](t) - t).detach()   # activations stored in bf16 / rounded to tf32
        return t

    def forward(self, imgs_nhwc, dropout_masks=None, dropout=0.2, keep_intermediates=False, forced_acts=None):
        """imgs_nhwc: [N,H,W,C] array/tensor.  dropout_masks: {'conv3a': keep[N,H,W,C], 'conv4a': ...} with 0/1
        entries (None and dropout>0 draws fresh masks; dropout == 0 disables).  Returns sigmoid map [N,H,W,n_cl].

        forced_acts (name -> NHWC array of the layer's stored output, e.g. read back from the GPU) turns the pass into
        the reference's backward EVALUATED AT A GIVEN FORWARD STATE: every layer output takes the forced value, ReLU /
        dropout gates are the forced tensor's non-zero pattern and max-pool routing follows the forced values, while
        gradients flow through the oracle's own fp32 operators.  (Gradients of a ReLU network are discontinuous in
        the forward state -- a gate that flips re-routes a whole gradient element -- so backward parity between two
        precisions is only meaningful at a common forward state; the forward itself is compared separately.)"""
        x = torch.as_tensor(np.asarray(imgs_nhwc), dtype=self.dtype).permute(0, 3, 1, 2)
        acts = OrderedDict()

        def forced(name):
            return torch.as_tensor(np.asarray(forced_acts[name]), dtype=self.dtype).permute(0, 3, 1, 2)

        def conv(name, t):
            k, b = self._w(name)
            dropped = name in ("conv3a", "conv4a") and dropout > 0
            if forced_acts is not None:
                yf = forced(name)
                y = (_conv3x3(t, k, b)) * (yf > 0).to(self.dtype)
                if dropped:
                    y = y / (1.0 - dropout)
                y = y + (yf - y).detach()
                acts[name] = y
                return y
            y = F.relu(_conv3x3(t, k, b))
            if dropped:                                        # model.py:60,66 (learning phase forced, model.py:40)
                if dropout_masks is not None:
                    keep = torch.as_tensor(np.asarray(dropout_masks[name]), dtype=self.dtype).permute(0, 3, 1, 2)
                else:
                    keep = torch.floor((1.0 - dropout) + torch.rand_like(y))   # floor(keep_prob + U[0,1))
                y = y * keep / (1.0 - dropout)
            y = self._act(y)
            acts[name] = y
            return y

        def up_path(name_t, t, skip):
            if self.up:
                u = F.interpolate(t, scale_factor=2, mode="nearest")          # UpSampling2D (model.py:76)
            else:
                k, b = self._w(name_t)
                u = _tconv2x2(t, k, b)
                if forced_acts is not None:
                    u = u + (forced(name_t) - u).detach()
                else:
                    u = self._act(u)
                acts[name_t] = u
            return torch.cat([u, skip], dim=1)                                # up first, skip second (model.py:76,79)

        c1 = conv("conv1b", conv("conv1a", x))
        p1 = F.max_pool2d(c1, 2)
        c2 = conv("conv2b", conv("conv2a", p1))
        p2 = F.max_pool2d(c2, 2)
        c3 = conv("conv3b", conv("conv3a", p2))
        p3 = F.max_pool2d(c3, 2)
        c4 = conv("conv4b", conv("conv4a", p3))
        p4 = F.max_pool2d(c4, 2)
        c5 = conv("conv5b", conv("conv5a", p4))
        c6 = conv("conv6b", conv("conv6a", up_path("transConv6", c5, c4)))
        c7 = conv("conv7b", conv("conv7a", up_path("transConv7", c6, c3)))
        c8 = conv("conv8b", conv("conv8a", up_path("transConv8", c7, c2)))
        c9 = conv("conv9b", conv("conv9a", up_path("transConv9", c8, c1)))
        k, b = self.params["Mask/kernel"], self.params["Mask/bias"]
        logits = F.conv2d(c9, k.permute(3, 2, 0, 1), b)
        preds = torch.sigmoid(logits).permute(0, 2, 3, 1)
        if keep_intermediates:
            self.acts = acts
        return preds


def dice_coef(y_true, y_pred, smooth=1.0):
    """model.py:4-10 -- sums run over the whole batch tensor."""
    inter = (y_true * y_pred).sum()
    return (2.0 * inter + smooth) / (y_true.sum() + y_pred.sum() + smooth)


def dice_coef_loss(y_true, y_pred, smooth=1.0):
    """model.py:13-21."""
    inter = (y_true * y_pred).sum()
    return -torch.log(2.0 * inter + smooth) + torch.log(y_true.sum() + y_pred.sum() + smooth)


def sensitivity(y_true, y_pred, smooth=1.0):
    """model.py:133-137."""
    inter = (y_true * y_pred).sum()
    return (inter + smooth) / (y_true.sum() + smooth)


def specificity(y_true, y_pred, smooth=1.0):
    """model.py:139-143 (really precision; formula preserved)."""
    inter = (y_true * y_pred).sum()
    return (inter + smooth) / (y_pred.sum() + smooth)


def loss_and_grads(weights, imgs, msks, use_upsampling=False, dropout_masks=None, dropout=0.2,
                   dtype=torch.float32, emulate_bf16=False, keep_act_grads=False, forced_acts=None):
    """One forward/backward (test_dist.py:204-212,246-248).  Returns dict with loss, dice, preds, grads (TF layouts,
    numpy) and optionally the activations and their gradients keyed by layer name (NHWC numpy).
    forced_acts: see Net.forward (backward evaluated at a given forward state)."""
    net = Net(weights, use_upsampling, dtype, emulate_bf16)
    preds = net.forward(imgs, dropout_masks, dropout, keep_intermediates=True, forced_acts=forced_acts)
    y = torch.as_tensor(np.asarray(msks), dtype=dtype)
    if keep_act_grads:
        for a in net.acts.values():
            a.retain_grad()
    loss = dice_coef_loss(y, preds)
    loss.backward()
    out = {
        "loss": float(loss.detach()), "dice": float(dice_coef(y, preds).detach()),
        "sensitivity": float(sensitivity(y, preds).detach()), "specificity": float(specificity(y, preds).detach()),
        "preds": preds.detach().numpy(),
        "grads": OrderedDict((k, p.grad.detach().numpy().copy()) for k, p in net.params.items()),
        "acts": OrderedDict((k, a.detach().permute(0, 2, 3, 1).numpy()) for k, a in net.acts.items()),
    }
    if keep_act_grads:
        out["act_grads"] = OrderedDict((k, a.grad.permute(0, 2, 3, 1).numpy().copy()) for k, a in net.acts.items())
    return out


# ------------------------------------------------------------------------------------------------ optimizer


def learning_rate(step, lr0=0.0005, const_lr=True, lr_fraction=0.2, decay_steps=100):
    """test_dist.py:228-232: exponential_decay(lr0, global_step, decay_steps, lr_fraction, staircase=False),
    evaluated with global_step before this update's increment."""
    if const_lr:
        return lr0
    return lr0 * lr_fraction ** (step / float(decay_steps))


def adam_update(weights, m, v, grads, step, lr0=0.0005, const_lr=True, lr_fraction=0.2, decay_steps=100,
                beta1=0.9, beta2=0.999, eps=1e-8):
    """TF-1.4 AdamOptimizer (test_dist.py:246): epsilon is added to sqrt(v) *before* bias correction is folded
    into lr_t -- this differs from torch.optim.Adam.  In place on float64/float32 numpy dicts; step = updates so far."""
    lr = learning_rate(step, lr0, const_lr, lr_fraction, decay_steps)
    t = step + 1
    lr_t = lr * math.sqrt(1.0 - beta2 ** t) / (1.0 - beta1 ** t)
    for k in weights:
        g = grads[k].astype(np.float64)
        m[k] = beta1 * m[k] + (1.0 - beta1) * g
        v[k] = beta2 * v[k] + (1.0 - beta2) * g * g
        weights[k] = (weights[k].astype(np.float64) - lr_t * m[k] / (np.sqrt(v[k]) + eps)).astype(weights[k].dtype)
    return lr


def sgd_update(weights, grads, step, lr0=0.0005, const_lr=True, lr_fraction=0.2, decay_steps=100):
    """GradientDescentOptimizer, the commented alternative at test_dist.py:239-244."""
    lr = learning_rate(step, lr0, const_lr, lr_fraction, decay_steps)
    for k in weights:
        weights[k] = (weights[k].astype(np.float64) - lr * grads[k].astype(np.float64)).astype(weights[k].dtype)
    return lr


def zeros_like_weights(weights):
    return OrderedDict((k, np.zeros(v.shape, np.float64)) for k, v in weights.items())


def sync_replica_grads(weights, imgs, msks, n_replicas, **kw):
    """SyncReplicasOptimizer semantics (test_dist.py:251-258, :388-394): replica r takes samples
    [r*B/n, (r+1)*B/n) of the global batch, computes the dice loss ON ITS SHARD, and the applied gradient is the
    MEAN of the replica gradients.  Returns (mean grads, per-replica losses)."""
    B = imgs.shape[0]
    per = int(B / n_replicas)
    acc, losses = None, []
    masks = kw.pop("dropout_masks", None)
    for r in range(n_replicas):
        sl = slice(r * per, (r + 1) * per)
        dm = None if masks is None else {k: m[sl] for k, m in masks.items()}
        res = loss_and_grads(weights, imgs[sl], msks[sl], dropout_masks=dm, **kw)
        losses.append(res["loss"])
        if acc is None:
            acc = OrderedDict((k, g.astype(np.float64)) for k, g in res["grads"].items())
        else:
            for k, g in res["grads"].items():
                acc[k] += g
    return OrderedDict((k, (g / n_replicas)) for k, g in acc.items()), losses


def train(weights, batches, steps, use_upsampling=False, dropout_masks_fn=None, dropout=0.2, lr0=0.0005,
          const_lr=True, lr_fraction=0.2, decay_steps=100, dtype=torch.float32, n_replicas=1):
    """The loop of test_dist.py:380-398 on pre-built batches [(imgs, msks), ...]; returns per-step loss/dice."""
    w = OrderedDict((k, v.copy()) for k, v in weights.items())
    m, v = zeros_like_weights(w), zeros_like_weights(w)
    hist = []
    for step in range(steps):
        imgs, msks = batches[step % len(batches)]
        dm = dropout_masks_fn(step) if dropout_masks_fn else None
        if n_replicas == 1:
            res = loss_and_grads(w, imgs, msks, use_upsampling, dm, dropout, dtype)
            grads, loss, dice = res["grads"], res["loss"], res["dice"]
        else:
            grads, losses = sync_replica_grads(w, imgs, msks, n_replicas, use_upsampling=use_upsampling,
                                               dropout_masks=dm, dropout=dropout, dtype=dtype)
            loss, dice = losses[0], float("nan")
        adam_update(w, m, v, grads, step, lr0, const_lr, lr_fraction, decay_steps)
        hist.append((loss, dice))
    return w, hist


# ------------------------------------------------------------------------------------------------ NumPy loop restatement
# Deliberately naive: these pin index conventions, not performance.  Tiny shapes only.


def np_conv3x3_same(x, k, b, relu=True):
    """x [N,H,W,Ci], k [3,3,Ci,Co] (HWIO), zero padding 1 -- Conv2D(padding='same') [TF-1.4]."""
    N, H, W, Ci = x.shape
    Co = k.shape[3]
    y = np.zeros((N, H, W, Co), np.float64)
    for n in range(N):
        for h in range(H):
            for w in range(W):
                acc = b.astype(np.float64).copy()
                for r in range(3):
                    for s in range(3):
                        hh, ww = h + r - 1, w + s - 1
                        if 0 <= hh < H and 0 <= ww < W:
                            acc += x[n, hh, ww, :].astype(np.float64) @ k[r, s].astype(np.float64)
                y[n, h, w] = np.maximum(acc, 0) if relu else acc
    return y


def np_tconv2x2(x, k, b):
    """Conv2DTranspose(2x2, stride 2, 'same'): out[n,2i+a,2j+b,o] = sum_c x[n,i,j,c] k[a,b,o,c] + bias[o]."""
    N, H, W, Ci = x.shape
    Co = k.shape[2]
    y = np.zeros((N, 2 * H, 2 * W, Co), np.float64)
    for n in range(N):
        for i in range(H):
            for j in range(W):
                for a in range(2):
                    for bb in range(2):
                        y[n, 2 * i + a, 2 * j + bb] = k[a, bb].astype(np.float64) @ x[n, i, j].astype(np.float64) + b
    return y


def np_maxpool2x2(x):
    N, H, W, C = x.shape
    return x.reshape(N, H // 2, 2, W // 2, 2, C).max(axis=(2, 4))


def np_maxpool2x2_grad(x, dy):
    """Gradient routed to the FIRST maximal element in row-major window order."""
    N, H, W, C = x.shape
    dx = np.zeros_like(x, dtype=np.float64)
    for n in range(N):
        for i in range(H // 2):
            for j in range(W // 2):
                for c in range(C):
                    win = x[n, 2 * i:2 * i + 2, 2 * j:2 * j + 2, c].reshape(-1)
                    a = int(np.argmax(win))            # argmax returns the first maximum
                    dx[n, 2 * i + a // 2, 2 * j + a % 2, c] += dy[n, i, j, c]
    return dx


def np_upsample2x2(x):
    return x.repeat(2, axis=1).repeat(2, axis=2)


def np_dice(y, p, smooth=1.0):
    inter = float((y.astype(np.float64) * p.astype(np.float64)).sum())
    sy, sp = float(y.astype(np.float64).sum()), float(p.astype(np.float64).sum())
    return {
        "loss": -math.log(2 * inter + smooth) + math.log(sy + sp + smooth),
        "dice": (2 * inter + smooth) / (sy + sp + smooth),
        "sensitivity": (inter + smooth) / (sy + smooth),
        "specificity": (inter + smooth) / (sp + smooth),
        "dloss_dp": -2.0 * y / (2 * inter + smooth) + 1.0 / (sy + sp + smooth),
    }
