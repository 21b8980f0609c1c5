"""Optimizers for the eager API: tf.train.AdamOptimizer + tf.train.exponential_decay as the reference wires them
(test_dist.py:228-248), and plain gradient descent (the commented alternative at test_dist.py:239-244).
The update itself is the fused CUDA kernel behind unetb200_adam_step / unetb200_sgd_step."""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .plan import _ptr, _stream


class _FlatOptimizer:
    def __init__(self, net, learning_rate=0.0005, const_learningrate=True, lr_fraction=0.2, decay_steps=100,
                 num_replicas=1):
        self.net = net
        self.lr0, self.const_lr = float(learning_rate), bool(const_learningrate)
        self.lr_fraction, self.decay_steps = float(lr_fraction), float(decay_steps)
        self.grad_scale = 1.0 / num_replicas
        self.global_step = torch.zeros(1, dtype=torch.int64, device=net.params.device)

    def zero_grad(self):
        self.net.params.grad = None

    def _grads(self):
        g = self.net.params.grad
        if g is None:
            raise _lib.UnetB200Error("optimizer.step(): no gradient -- call loss.backward() first")
        return g.contiguous()

    def learning_rate(self):
        """Value the next update will use (global_step before the increment, staircase=False)."""
        step = int(self.global_step.item())
        return self.lr0 if self.const_lr else self.lr0 * self.lr_fraction ** (step / self.decay_steps)


class AdamOptimizer(_FlatOptimizer):
    def __init__(self, net, learning_rate=0.0005, beta1=0.9, beta2=0.999, epsilon=1e-8, **kw):
        super().__init__(net, learning_rate, **kw)
        self.beta1, self.beta2, self.eps = float(beta1), float(beta2), float(epsilon)
        self.m = torch.zeros_like(net.params.detach())
        self.v = torch.zeros_like(net.params.detach())

    @torch.no_grad()
    def step(self):
        p = self.net.params
        _lib.call("unetb200_adam_step", _ptr(p), _ptr(self.m), _ptr(self.v), _ptr(self._grads()),
                  ctypes.c_longlong(p.numel()), _ptr(self.global_step), self.lr0, int(self.const_lr), self.lr_fraction,
                  self.decay_steps, self.beta1, self.beta2, self.eps, self.grad_scale, _stream())
        _lib.call("unetb200_step_increment", _ptr(self.global_step), _stream())
        self.net.plan.pack_weights()


class GradientDescentOptimizer(_FlatOptimizer):
    @torch.no_grad()
    def step(self):
        p = self.net.params
        _lib.call("unetb200_sgd_step", _ptr(p), _ptr(self._grads()), ctypes.c_longlong(p.numel()),
                  _ptr(self.global_step), self.lr0, int(self.const_lr), self.lr_fraction, self.decay_steps,
                  self.grad_scale, _stream())
        _lib.call("unetb200_step_increment", _ptr(self.global_step), _stream())
        self.net.plan.pack_weights()
