"""The training step of test_dist.py:380-398 on B200: forward -> fused dice loss -> backward (with the gradient
buckets all-reduced as they complete) -> TF-style Adam with the decaying learning rate -> weight re-pack.
No host synchronisation happens inside a step; loss / dice stay on the device until the caller reads them."""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib, parallel
from .init import initial_weights
from .plan import UNetPlan, _ptr, _stream


class Trainer:
    """Owns one UNetPlan (per-GPU batch N), the Adam slots and the device-side global_step.

    Flags follow test_dist.py:64-85 / settings_dist.py:17-22: learning_rate, const_learningrate, lr_fraction,
    decay_steps, use_upsampling.  `optimizer` may be "adam" (the reference, test_dist.py:246) or "sgd" (the
    commented alternative, test_dist.py:239-244).  `precision`: "bf16" or "tf32" / "fp32" (see UNetPlan)."""

    def __init__(self, N, H=128, W=128, in_channels=1, n_cl_out=1, use_upsampling=False, dropout=0.2,
                 learning_rate=0.0005, const_learningrate=True, lr_fraction=0.2, decay_steps=100, optimizer="adam",
                 beta1=0.9, beta2=0.999, epsilon=1e-8, weight_seed=1234, dropout_seed=816, weights=None,
                 use_cuda_graph=False, shared_dropout_masks=False, precision="bf16"):
        self.world, self.rank = parallel.world()
        self.plan = UNetPlan(N, H, W, in_channels, n_cl_out, use_upsampling, dropout, precision=precision)
        p = self.plan
        dev = p.device
        self.m = torch.zeros_like(p.params)
        self.v = torch.zeros_like(p.params)
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)   # global_step (test_dist.py:185)
        self.lr0, self.const_lr = float(learning_rate), bool(const_learningrate)
        self.lr_fraction, self.decay_steps = float(lr_fraction), float(decay_steps)
        self.beta1, self.beta2, self.eps = float(beta1), float(beta2), float(epsilon)
        self.optimizer = optimizer
        if optimizer not in ("adam", "sgd"):
            raise ValueError("optimizer must be 'adam' or 'sgd'")
        # Every replica draws its own dropout masks, as the reference's workers do (each sess.run samples anew,
        # model.py:60,66); shared_dropout_masks=True gives all ranks the same Philox seed (parity tests).
        self.dropout_seed = parallel.replica_seed(dropout_seed, 0 if shared_dropout_masks else self.rank)
        self._eval_calls = 0
        p.set_weights(weights if weights is not None else initial_weights(p.variables, weight_seed))
        parallel.broadcast_flat(p.params)
        if self.world > 1:
            p.pack_weights()
        self.reducer = parallel.BucketedAllReduce(p.grads, p.buckets)
        # bucket updates (all-reduce wait, Adam, re-pack) run here, beside the rest of backward, at high priority: they
        # are short and the next step's forward waits for the last of them (UNETB200_UPD_PRIORITY=same turns it off)
        self._upd_stream = torch.cuda.Stream(device=dev,
                                             priority=0 if os.environ.get("UNETB200_UPD_PRIORITY") == "same" else -1)
        # SMs left to the NCCL kernels that run beside backward segments 1..4 (see unetb200_set_reserved_sms)
        self.reserved_sms = int(os.environ.get("UNETB200_RESERVE_SMS", str(parallel.RESERVED_SMS)))
        self.use_cuda_graph = bool(use_cuda_graph)
        # True: the caller's stream waits for every bucket update before the next segment (bench.py's per-kernel timing,
        # where a kernel must not share the GPU with another stream's work); eager steps only
        self.serial_update = False
        self._graph = None
        self._g_imgs = self._g_msks = None
        self.steps_done = 0
        # host-facing step: two device input slots + pinned staging, created on first use (_pipeline)
        self._slots = None
        self._pending = None
        self._d_imgs = torch.empty((N, H, W, in_channels), dtype=torch.float32, device=dev)
        self._d_msks = torch.empty((N, H, W, n_cl_out), dtype=torch.float32, device=dev)

    # ------------------------------------------------------------------ one optimizer step on device-resident data
    def _enqueue_step(self, imgs, msks):
        """The whole step as one launch sequence (eager, or captured into ONE CUDA graph at any world size):
        forward, then per backward segment: the segment's kernels -> all-reduce of the bucket it completes (NCCL on
        its own stream) -> optimizer update + operand re-pack of that bucket on the update stream.  Only the last,
        smallest bucket (conv1a..conv2b, 0.26 MB) has no backward work left to hide behind."""
        p = self.plan
        cur = torch.cuda.current_stream()
        upd = self._upd_stream
        p.forward(imgs, msks, seed=self.dropout_seed, seed_add=self.step_dev)
        lib = _lib.lib()
        # the dgrad chain does not wait for the weight gradients between segments: the update stream does
        old_defer = p.defer_join(True)
        for seg in range(p.num_segments):
            # from the second segment on, an all-reduce runs beside the backward kernels: leave it its SMs
            lib.unetb200_set_reserved_sms(self.reserved_sms if (self.world > 1 and seg > 0) else 0)
            p.backward(imgs, msks, grad_scale=1.0, segment=seg)
            upd.wait_stream(cur)
            with torch.cuda.stream(upd):
                p.join()                                       # ... and behind the segment's weight gradients
                # the all-reduce is ordered behind the segment through the update stream; `cur` does not wait for it
                self.reducer.reduce_now(seg)
                self._update_range(*p.buckets[seg])
            if self.serial_update:
                cur.wait_stream(upd)
        lib.unetb200_set_reserved_sms(0)
        p.defer_join(old_defer)
        with torch.cuda.stream(upd):
            _lib.call("unetb200_step_increment", _ptr(self.step_dev), _stream())
        cur.wait_stream(upd)

    def _update_range(self, b, e):
        """Optimizer update of the flat range [b, e) and re-pack of the GEMM operands inside it (current stream)."""
        p = self.plan
        if e <= b:
            return
        scale = 1.0 / self.world            # SyncReplicasOptimizer averages the replica gradients
        n = ctypes.c_longlong(e - b)
        if self.optimizer == "adam":
            _lib.call("unetb200_adam_step", _ptr(p.params[b:e]), _ptr(self.m[b:e]), _ptr(self.v[b:e]),
                      _ptr(p.grads[b:e]), n, _ptr(self.step_dev), self.lr0, int(self.const_lr), self.lr_fraction,
                      self.decay_steps, self.beta1, self.beta2, self.eps, scale, _stream())
        else:
            _lib.call("unetb200_sgd_step", _ptr(p.params[b:e]), _ptr(p.grads[b:e]), n, _ptr(self.step_dev), self.lr0,
                      int(self.const_lr), self.lr_fraction, self.decay_steps, scale, _stream())
        p.pack_weights(b, e)

    def train_step(self, imgs, msks, _after=None):
        """imgs/msks: contiguous CUDA float32 NHWC tensors for THIS replica's shard. Returns the device tensor
        [loss, dice, sensitivity, specificity] of this step (no synchronisation). `_after` (internal): a CUDA event the
        step's kernels wait for -- the read-back of the previous step's metrics, which this step overwrites."""
        if self.use_cuda_graph:
            if self._graph is None:
                self._capture(imgs, msks)
            self._g_imgs.copy_(imgs, non_blocking=True)
            self._g_msks.copy_(msks, non_blocking=True)
            if _after is not None:
                torch.cuda.current_stream().wait_event(_after)
            self._graph.replay()      # one replay per step at every world size (the all-reduces are inside)
        else:
            if _after is not None:
                torch.cuda.current_stream().wait_event(_after)
            self._enqueue_step(imgs, msks)
        self.steps_done += 1
        return self.plan.metrics

    def close(self):
        """Release the captured step. A CUDA graph that holds NCCL kernels keeps the communicator busy:
        torch.distributed.destroy_process_group() blocks until every such graph is gone, so call this first."""
        self._graph = None
        torch.cuda.synchronize(self.plan.device)

    def _capture(self, imgs, msks):
        self._g_imgs, self._g_msks = imgs.clone(), msks.clone()
        # warm-up outside capture (kernel attribute set-up, lazy module loading, NCCL channel set-up) on a side stream,
        # then restore state
        snap = [t.clone() for t in (self.plan.params, self.m, self.v, self.step_dev)]
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            self._enqueue_step(self._g_imgs, self._g_msks)
        torch.cuda.current_stream().wait_stream(s)
        for dst, src in zip((self.plan.params, self.m, self.v, self.step_dev), snap):
            dst.copy_(src)
        self.plan.pack_weights()
        torch.cuda.synchronize()
        before = _lib.lib().unetb200_launch_count()
        self._graph = torch.cuda.CUDAGraph()
        # thread_local: the NCCL watchdog thread may query events while this thread captures
        with torch.cuda.graph(self._graph, capture_error_mode="thread_local"):
            self._enqueue_step(self._g_imgs, self._g_msks)
        # kernels of this library inside one replay (the library's launch counter only sees the capture)
        self.graph_launches_per_step = int(_lib.lib().unetb200_launch_count() - before)

    # ------------------------------------------------------------------ host-facing step (what a user calls)
    def _pipeline(self):
        """Two input slots + a copy stream: the host->device copy of step k+1 runs while step k computes."""
        if self._slots is None:
            dev = self.plan.device
            self._copy_stream = torch.cuda.Stream(device=dev)
            # the 16-byte read-back of a step's metrics runs on its own stream: a copy-engine operation in the compute
            # stream would sit between two graph launches (UNETB200_D2H_INLINE=1 puts it back there)
            self._d2h_stream = None if os.environ.get("UNETB200_D2H_INLINE") else torch.cuda.Stream(device=dev)
            self._slots = []
            for _ in range(2):
                self._slots.append({
                    "imgs": torch.empty_like(self._d_imgs), "msks": torch.empty_like(self._d_msks),
                    "ready": torch.cuda.Event(), "free": torch.cuda.Event(), "done": torch.cuda.Event(),
                    "met_ready": torch.cuda.Event(),
                    "h_metrics": torch.empty(4, dtype=torch.float32).pin_memory(),
                    "h_imgs": None, "h_msks": None})
            self._k = 0
            self._pending = None
        return self._slots

    def train_step_host(self, imgs_np, msks_np, wait=True):
        """One step fed from HOST arrays (the feed_dict of test_dist.py:394-398): stages this replica's shard in pinned
        memory, copies it to the device, runs the step and reads the step's [loss, dice, sens, spec] back.
        See train_step_pinned for `wait`."""
        slot = self._pipeline()[self._k % 2]
        if slot["h_imgs"] is None:
            slot["h_imgs"] = torch.empty(self._d_imgs.shape, dtype=torch.float32).pin_memory()
            slot["h_msks"] = torch.empty(self._d_msks.shape, dtype=torch.float32).pin_memory()
        slot["ready"].synchronize()   # the copy that last read this staging buffer (two steps ago) has finished
        slot["h_imgs"].numpy()[...] = imgs_np
        slot["h_msks"].numpy()[...] = msks_np
        return self.train_step_pinned(slot["h_imgs"], slot["h_msks"], wait=wait)

    def train_step_pinned(self, h_imgs, h_msks, wait=True):
        """One step fed from pinned host tensors. wait=True (default): returns this step's metrics as a numpy array
        (one synchronisation per step). wait=False: software-pipelined -- the step is enqueued, the PREVIOUS
        step's metrics are returned (None on the first call) and the copy of the next batch overlaps this step's
        kernels; call finish() for the last one. The caller must not overwrite h_imgs / h_msks before the following
        call returns."""
        slots = self._pipeline()
        slot = slots[self._k % 2]
        self._k += 1
        cur = torch.cuda.current_stream()
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(slot["free"])      # step k-2 has consumed this slot (no-op the first time)
            slot["imgs"].copy_(h_imgs, non_blocking=True)
            slot["msks"].copy_(h_msks, non_blocking=True)
            slot["ready"].record(self._copy_stream)
        cur.wait_event(slot["ready"])
        prev = self._pending
        if self._d2h_stream is None:
            met = self.train_step(slot["imgs"], slot["msks"])
            slot["free"].record(cur)
            slot["h_metrics"].copy_(met, non_blocking=True)
            slot["done"].record(cur)
        else:
            # plan.metrics is one buffer: this step may only overwrite it once the previous read-back has finished
            met = self.train_step(slot["imgs"], slot["msks"], _after=prev["done"] if prev is not None else None)
            slot["free"].record(cur)
            slot["met_ready"].record(cur)
            with torch.cuda.stream(self._d2h_stream):
                self._d2h_stream.wait_event(slot["met_ready"])
                slot["h_metrics"].copy_(met, non_blocking=True)
                slot["done"].record(self._d2h_stream)
        self._pending = slot
        if wait:
            self._pending = None
            slot["done"].synchronize()
            return slot["h_metrics"].numpy().copy()
        if prev is None:
            return None
        prev["done"].synchronize()
        return prev["h_metrics"].numpy().copy()

    def finish(self):
        """Metrics of the last pipelined step (train_step_pinned(..., wait=False)), or None if nothing is pending."""
        prev, self._pending = self._pending, None
        if prev is None:
            return None
        prev["done"].synchronize()
        return prev["h_metrics"].numpy().copy()

    def h2d_bytes_per_step(self):
        return (self._d_imgs.numel() + self._d_msks.numel()) * 4

    def d2h_bytes_per_step(self):
        return 16

    # ------------------------------------------------------------------ evaluation / state
    def predict(self, imgs, msks=None):
        """Forward only (dropout stays on, as in the reference: model.py:40). Returns (preds, metrics)."""
        self.plan.forward(imgs, msks, seed=self._eval_seed(), seed_add=self.step_dev)
        return self.plan.preds, (self.plan.metrics if msks is not None else None)

    def dice_sums(self, imgs_np, msks_np):
        """Forward only on one per-device batch of HOST arrays; returns numpy [sum(y*p), sum(y), sum(p)].
        Dropout stays on, as in the reference's test pass (model.py:40)."""
        self._d_imgs.copy_(torch.as_tensor(np.ascontiguousarray(imgs_np, dtype=np.float32)))
        self._d_msks.copy_(torch.as_tensor(np.ascontiguousarray(msks_np, dtype=np.float32)))
        self.plan.forward(self._d_imgs, self._d_msks, seed=self._eval_seed(), seed_add=self.step_dev)
        return self.plan.sums.cpu().numpy().astype(np.float64)

    def _eval_seed(self):
        """Every evaluation forward draws fresh dropout masks (the reference samples anew on each sess.run of the
        test pass, test_dist.py:421): a per-call counter is folded into the Philox seed."""
        self._eval_calls += 1
        return (self.dropout_seed + 0xD1B54A32D192ED03 * self._eval_calls) & 0xFFFFFFFFFFFFFFFF

    def evaluate(self, imgs_test, msks_test, batch_size):
        """Test-set metrics with the reference's batching and averaging (test_dist.py:406-424); see evaluate.py."""
        from .evaluate import evaluate
        return evaluate(self.dice_sums, imgs_test, msks_test, batch_size, self.plan.N)

    def global_step(self):
        return int(self.step_dev.item())

    def state_dict(self):
        p = self.plan
        sd = {"global_step": self.global_step()}
        for name in p.variables:
            sd[name] = p.param(name).detach().cpu().numpy().copy()
            sd[name + "/Adam"] = p._view(self.m, name).detach().cpu().numpy().copy()      # TF slot names
            sd[name + "/Adam_1"] = p._view(self.v, name).detach().cpu().numpy().copy()
        return sd

    def load_state_dict(self, sd):
        p = self.plan
        for name in p.variables:
            p.param(name).copy_(torch.as_tensor(np.asarray(sd[name], np.float32)).to(p.device))
            if name + "/Adam" in sd:
                p._view(self.m, name).copy_(torch.as_tensor(np.asarray(sd[name + "/Adam"], np.float32)).to(p.device))
                p._view(self.v, name).copy_(torch.as_tensor(np.asarray(sd[name + "/Adam_1"], np.float32)).to(p.device))
        self.step_dev.fill_(int(sd.get("global_step", 0)))
        p.pack_weights()
