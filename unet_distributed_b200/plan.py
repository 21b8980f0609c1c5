"""Python owner of a libunetb200 network plan: allocates the arena and the flat fp32 buffers with PyTorch (tensor
ownership only) and drives the native launch sequence.  Mirrors the graph model.py:36-130 of the reference builds."""
from __future__ import annotations

import ctypes
from collections import OrderedDict

import numpy as np
import torch

from . import _lib


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


PRECISIONS = {"bf16": _lib.BF16, "tf32": _lib.TF32, "fp32": _lib.TF32}   # "fp32" = the fp32/tf32 mode


def _dtype_code(precision):
    try:
        return PRECISIONS[str(precision).lower()]
    except KeyError:
        raise ValueError(f"precision must be one of {sorted(PRECISIONS)}, got {precision!r}") from None


def describe_plan(N, H, W, in_channels=1, n_cl_out=1, use_upsampling=False, dropout=0.2, precision="bf16"):
    """Host-only description of a plan (no GPU needed): variable table (TF names -> (offset, shape)), padded flat
    size, arena bytes and the gradient buckets (one per backward segment).  This is the native planner's own answer, not a Python copy."""
    L = _lib.lib()
    h = ctypes.c_void_p()
    _lib.check(L.unetb200_plan_create_dt(ctypes.byref(h), int(N), int(H), int(W), int(in_channels), int(n_cl_out),
                                         int(bool(use_upsampling)), float(dropout), _dtype_code(precision)),
               "unetb200_plan_create_dt")
    try:
        variables = OrderedDict()
        name = ctypes.create_string_buffer(64)
        off, rank = ctypes.c_longlong(), ctypes.c_int()
        shape = (ctypes.c_int * 4)()
        for i in range(L.unetb200_plan_num_variables(h)):
            _lib.check(L.unetb200_plan_variable(h, i, name, 64, ctypes.byref(off), ctypes.byref(rank), shape),
                       "unetb200_plan_variable")
            variables[name.value.decode()] = (int(off.value), tuple(shape[k] for k in range(rank.value)))
        buckets = []
        b, e = ctypes.c_longlong(), ctypes.c_longlong()
        for seg in range(L.unetb200_plan_num_segments(h)):
            _lib.check(L.unetb200_plan_bucket(h, seg, ctypes.byref(b), ctypes.byref(e)), "unetb200_plan_bucket")
            buckets.append((int(b.value), int(e.value)))
        return {"variables": variables, "num_params_padded": int(L.unetb200_plan_num_params(h)),
                "num_params": sum(int(np.prod(s)) for _, s in variables.values()),
                "arena_bytes": int(L.unetb200_plan_arena_bytes(h)), "buckets": buckets}
    finally:
        L.unetb200_plan_destroy(h)


class UNetPlan:
    """One UNet instance for a fixed (N, H, W, in_channels) input shape on the current CUDA device.

    precision: "bf16" (default: bf16 activations and GEMM operands, fp32 accumulation) or "tf32" / "fp32" (the
    reference's precision class, test_dist.py:196-202: fp32 activations rounded to tf32, kind::tf32 tensor-core MMAs,
    fp32 accumulation and fp32 bandwidth kernels; arena tensors are then float32)."""

    def __init__(self, N, H, W, in_channels=1, n_cl_out=1, use_upsampling=False, dropout=0.2, device=None,
                 precision="bf16"):
        if not torch.cuda.is_available():
            raise _lib.UnetB200Error("UNetPlan needs a CUDA device (sm_100a); there is no CPU path")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.N, self.H, self.W = int(N), int(H), int(W)
        self.in_channels, self.n_cl_out = int(in_channels), int(n_cl_out)
        self.use_upsampling, self.dropout = bool(use_upsampling), float(dropout)
        self.precision = "tf32" if _dtype_code(precision) == _lib.TF32 else "bf16"
        L = _lib.lib()
        h = ctypes.c_void_p()
        _lib.check(L.unetb200_plan_create_dt(ctypes.byref(h), self.N, self.H, self.W, self.in_channels, self.n_cl_out,
                                             int(self.use_upsampling), self.dropout, _dtype_code(precision)),
                   "unetb200_plan_create_dt")
        self._h = h
        self.num_params_padded = int(L.unetb200_plan_num_params(h))
        self.variables = OrderedDict()
        name = ctypes.create_string_buffer(64)
        off, rank = ctypes.c_longlong(), ctypes.c_int()
        shape = (ctypes.c_int * 4)()
        for i in range(L.unetb200_plan_num_variables(h)):
            _lib.check(L.unetb200_plan_variable(h, i, name, 64, ctypes.byref(off), ctypes.byref(rank), shape),
                       "unetb200_plan_variable")
            self.variables[name.value.decode()] = (int(off.value), tuple(shape[k] for k in range(rank.value)))
        self.num_params = sum(int(np.prod(s)) for _, s in self.variables.values())
        with torch.cuda.device(self.device):
            nbytes = int(L.unetb200_plan_arena_bytes(h))
            self._arena_raw = torch.empty(nbytes + 1024, dtype=torch.uint8, device=self.device)
            shift = (-self._arena_raw.data_ptr()) % 1024
            self.arena = self._arena_raw[shift:shift + nbytes]
            self.params = torch.zeros(self.num_params_padded, dtype=torch.float32, device=self.device)
            self.grads = torch.zeros(self.num_params_padded, dtype=torch.float32, device=self.device)
        self.arena_bytes = nbytes
        self.bind(self.params, self.grads)
        self.buckets = []
        b, e = ctypes.c_longlong(), ctypes.c_longlong()
        self.num_segments = int(L.unetb200_plan_num_segments(h))
        for seg in range(self.num_segments):
            _lib.check(L.unetb200_plan_bucket(h, seg, ctypes.byref(b), ctypes.byref(e)), "unetb200_plan_bucket")
            self.buckets.append((int(b.value), int(e.value)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                _lib.lib().unetb200_plan_destroy(h)
            except Exception:
                pass
            self._h = None

    # ------------------------------------------------------------------ buffers
    def bind(self, params, grads):
        """Use caller-provided flat fp32 buffers (e.g. views of a trainer-wide allocation)."""
        assert params.numel() == self.num_params_padded and grads.numel() == self.num_params_padded
        self.params, self.grads = params, grads
        _lib.call("unetb200_plan_bind", self._h, _ptr(self.arena), _ptr(params), _ptr(grads))

    def _view(self, flat, name):
        off, shape = self.variables[name]
        return flat[off:off + int(np.prod(shape))].view(*shape)

    def param(self, name):
        return self._view(self.params, name)

    def grad(self, name):
        return self._view(self.grads, name)

    def set_weights(self, weights):
        """weights: {TF variable name: array in TF layout}. Refreshes the packed bf16 GEMM operands."""
        with torch.no_grad():
            for name in self.variables:
                w = torch.as_tensor(np.asarray(weights[name], dtype=np.float32))
                self.param(name).copy_(w.to(self.device))
        self.pack_weights()

    def get_weights(self):
        return OrderedDict((n, self.param(n).detach().cpu().numpy().copy()) for n in self.variables)

    def get_grads(self):
        return OrderedDict((n, self.grad(n).detach().cpu().numpy().copy()) for n in self.variables)

    def tensor(self, name):
        """View of a named arena tensor: activations 'a1'..'c9','p1'..'p4','u6'..'u9' (bf16 NHWC), their gradients
        ('g'+name), 'preds' (fp32 NHWC) or 'metrics' (fp32 [loss, dice, sensitivity, specificity])."""
        off = ctypes.c_longlong()
        n, h, w, c, eb = (ctypes.c_int() for _ in range(5))
        _lib.call("unetb200_plan_tensor", self._h, name.encode(), ctypes.byref(off), ctypes.byref(n), ctypes.byref(h),
                  ctypes.byref(w), ctypes.byref(c), ctypes.byref(eb))
        count = n.value * h.value * w.value * c.value
        raw = self.arena[off.value:off.value + count * eb.value]
        dt = torch.bfloat16 if eb.value == 2 else torch.float32
        t = raw.view(dt)
        if name in ("metrics", "sums"):
            return t.view(count)
        return t.view(n.value, h.value, w.value, c.value)

    @property
    def preds(self):
        return self.tensor("preds")

    @property
    def metrics(self):
        return self.tensor("metrics")

    @property
    def sums(self):
        """[sum(y*p), sum(y), sum(p)] of the last forward with labels (fp32, device)."""
        return self.tensor("sums")

    # ------------------------------------------------------------------ launches (asynchronous on the current stream)
    def side_stream(self, enable):
        """Run weight gradients on the plan's second stream (default) or in program order; returns the old setting."""
        return bool(_lib.lib().unetb200_plan_side_stream(self._h, 1 if enable else 0))

    def defer_join(self, enable):
        """Segments of backward() stop ending with a wait for their weight gradients; call join() on the stream that
        consumes the gradients instead (see unetb200_plan_defer_join). Returns the old setting."""
        return bool(_lib.lib().unetb200_plan_defer_join(self._h, 1 if enable else 0))

    def join(self):
        """Make the CURRENT stream wait for every weight gradient enqueued so far."""
        _lib.call("unetb200_plan_join", self._h, _stream())

    def pack_weights(self, begin=None, end=None):
        """Refresh the packed bf16 GEMM operands from the fp32 masters: all of them, or only the variables inside the
        flat range [begin, end) (one gradient bucket)."""
        if begin is None:
            _lib.call("unetb200_plan_pack_weights", self._h, _stream())
        else:
            _lib.call("unetb200_plan_pack_range", self._h, ctypes.c_longlong(begin), ctypes.c_longlong(end), _stream())

    def forward(self, imgs, msks=None, seed=0, seed_add=None):
        """imgs fp32 [N,H,W,Cin] and msks fp32 [N,H,W,n_cl] on this device, contiguous."""
        self._check_input(imgs, self.in_channels, "imgs")
        if msks is not None:
            self._check_input(msks, self.n_cl_out, "msks")
        _lib.call("unetb200_plan_forward", self._h, _ptr(imgs), _ptr(msks), ctypes.c_uint64(int(seed)),
                  _ptr(seed_add), _stream())

    def backward(self, imgs, msks=None, dpreds=None, grad_scale=1.0, segment=None):
        segs = range(self.num_segments) if segment is None else (segment,)
        for s in segs:
            _lib.call("unetb200_plan_backward", self._h, _ptr(imgs), _ptr(msks), _ptr(dpreds),
                      ctypes.c_float(grad_scale), int(s), _stream())

    def _check_input(self, t, C, what):
        want = (self.N, self.H, self.W, C)
        if tuple(t.shape) != want or t.dtype != torch.float32 or not t.is_cuda or not t.is_contiguous():
            raise ValueError(f"{what}: expected contiguous CUDA float32 {want}, got {tuple(t.shape)} {t.dtype} "
                             f"{t.device}")
