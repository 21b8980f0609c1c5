#!/usr/bin/env python
"""Benchmark of the UNet training hot path (BASELINE.json metric: train images/sec, 128x128 slices, bs128 per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--decoder upsampling|tconv] [--impl reference]

One "step" = one optimizer step over one synthetic batch: forward, fused dice loss, backward, gradient all-reduce
(N > 1), TF-style Adam, weight re-pack.  N > 1 is launched by torchrun (one rank per GPU, weak scaling: bs128 per GPU).
Prints ONE JSON line on rank 0.  `--impl reference` times the CPU restatement of the reference (oracle/) on the host
cores instead -- the reference itself (TensorFlow 1.4, Python 2.7) cannot be installed here."""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# algorithmic training GFLOP per image: fwd + dgrad + wgrad of every conv / tconv / 1x1
TRAIN_GFLOP_PER_IMG = {("tconv", 128, 1): 18.0408, ("upsampling", 128, 1): 15.4235,
                       ("tconv", 240, 4): 63.6236, ("upsampling", 240, 4): 54.4223}
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=30)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    p.add_argument("--decoder", type=str, default="upsampling", choices=["upsampling", "tconv"],
                   help="configs[1] of BASELINE.json (--use_upsampling) is the default single-GPU workload")
    p.add_argument("--batch", type=int, default=128, help="per-GPU batch")
    p.add_argument("--size", type=int, default=128)
    p.add_argument("--channels", type=int, default=1)
    p.add_argument("--no-graph", action="store_true", help="do not replay the step from a CUDA graph (1 GPU)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-tconv-block", dest="no_tconv_block", action="store_true",
                   help="skip the side measurement of the Conv2DTranspose decoder")
    p.add_argument("--const_learningrate", type=int, default=1)
    p.add_argument("--decay_steps", type=int, default=100)
    p.add_argument("--lr_fraction", type=float, default=0.2)
    p.add_argument("--precision", type=str, default="bf16", choices=["bf16", "tf32"],
                   help="bf16 (the headline) or the fp32/tf32 mode; for tf32 the tensor peak is a cuBLAS tf32 GEMM "
                        "measured in this process (MEASURED_PEAKS.json has no tf32 number)")
    return p.parse_args()


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        d["_source"] = "measured (MEASURED_PEAKS.json)"
        return d
    d = dict(FALLBACK_PEAKS)
    d["_source"] = "fallback (B200_PROFILING.md)"
    return d


_SAMPLER_SRC = r"""
import sys, time
import pynvml as nv
nv.nvmlInit()
uuid, idx = sys.argv[1], int(sys.argv[2])
try:
    try:
        h = nv.nvmlDeviceGetHandleByUUID(uuid)
    except TypeError:
        h = nv.nvmlDeviceGetHandleByUUID(uuid.encode())
except Exception:
    h = nv.nvmlDeviceGetHandleByIndex(idx)
print("max", nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM), flush=True)
while True:
    try:
        try:
            mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
        except Exception:
            mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
        print(time.time(), nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM), mask, flush=True)
    except Exception:
        pass
    time.sleep(0.004)
"""


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region by a CHILD PROCESS polling NVML every ~4 ms
    (started before the warm-up; samples are filtered to the timed region by wall-clock time). A polling thread in
    this process took the GIL from the launch loop and cost 8 % at N = 8, where a step is four graph launches plus
    three NCCL calls; nvidia-smi's own loop starts too slowly for an 80 ms timed region."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
               (0x4, "sw_power_cap"))

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.proc = None
        self.t0 = self.t1 = None

    def start(self):
        """Launch the child (call well before the timed region: importing NVML takes a few hundred ms)."""
        try:
            import torch
            uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            parts = vis.split(",") if vis else []
            idx = int(parts[self.gpu]) if self.gpu < len(parts) and parts[self.gpu].isdigit() else self.gpu
            self.proc = subprocess.Popen([sys.executable, "-c", _SAMPLER_SRC, uuid, str(idx)], stdout=subprocess.PIPE,
                                         stderr=open(os.environ.get("UNETB200_SAMPLER_ERR", os.devnull), "w"), text=True)
        except Exception:
            self.proc = None

    def begin(self):
        self.t0 = time.time()

    def end(self):
        self.t1 = time.time()

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            text, _ = self.proc.communicate(timeout=5)
        except Exception:
            try:
                self.proc.kill()
            except Exception:
                pass
            return out
        sm, late, reasons, mx = [], [], set(), None
        for line in text.splitlines():
            parts = line.split()
            try:
                if parts[0] == "max":
                    mx = float(parts[1])
                    continue
                t, clk, mask = float(parts[0]), float(parts[1]), int(parts[2])
            except (ValueError, IndexError):
                continue
            inside = self.t0 is None or self.t1 is None or (self.t0 <= t <= self.t1)
            late.append(clk)
            if not inside:
                continue
            sm.append(clk)
            for bit, name in self.REASONS:
                if mask & bit:
                    reasons.add(name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm),
                       source="nvml child process, 4 ms polling, samples inside the timed region")
        elif late:   # the child came up too late for the timed region: report what it saw, and say so
            out.update(sm_mhz=statistics.median(late), sm_max_mhz=mx, samples=len(late),
                       source="nvml child process, samples OUTSIDE the timed region (child started late)")
        return out


# ---------------------------------------------------------------------------------------------------- shared
def workload_config(args, world):
    """The `config` object of the JSON line -- the same for the GPU arm and for --impl reference (which times the CPU
    restatement ON the GPU arm's configuration; cuda_graph / l2 describe the GPU arm)."""
    N, S, C = args.batch, args.size, args.channels
    ups = args.decoder == "upsampling"
    cfg_index = 3 if (S, C) == (240, 4) else (4 if N == 512 else (2 if world > 1 else (1 if ups else 0)))
    return {
        "workload": f"UNet {S}x{S}x{C} {args.decoder} decoder (BASELINE configs[{cfg_index}] shape), "
                    f"bs{N} per GPU, Adam, dropout 0.2, train step fwd+dice+bwd+allreduce+Adam",
        "decoder": args.decoder, "per_gpu_batch": N, "global_batch": N * world, "image": [S, S, C],
        "parallelism": f"dp{world}", "const_learningrate": bool(args.const_learningrate), "precision": args.precision,
        "cuda_graph": bool(not args.no_graph),
        "l2": "no explicit flush: one step touches >2.5 GB of activations and gradients (L2 is 126 MB); "
              "input batches rotate"}


def cpu_model_name():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


# ---------------------------------------------------------------------------------------------------- CPU arms
CPU_WARMUP, CPU_STEPS = 2, 5      # 2 warm-up + 5 timed steps of the bs128 workload


def cpu_train_images_per_sec(decoder, size, channels, images_per_step, steps, warmup, threads):
    """The oracle's training step (fwd + dice loss + bwd + TF-Adam, fp32, dropout on) on the host cores."""
    import torch
    from oracle import unet_oracle as O
    from unet_distributed_b200.synthetic import make_slices
    torch.set_num_threads(threads)
    ups = decoder == "upsampling"
    w = O.init_weights(1234, ups, channels)
    m, v = O.zeros_like_weights(w), O.zeros_like_weights(w)
    imgs, msks = make_slices(images_per_step, size, size, channels)
    t_total, step = 0.0, 0
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        res = O.loss_and_grads(w, imgs, msks, ups, None, 0.2, torch.float32)
        O.adam_update(w, m, v, res["grads"], step)
        step += 1
        dt = time.perf_counter() - t0
        if it >= warmup:
            t_total += dt
    return images_per_step * steps / t_total, t_total / steps * 1e3


def cpu_baseline_block(args, steps, warmup):
    """One recipe for both arms: every step is the FULL per-GPU batch of the workload (bs128 unless --batch says
    otherwise) through the torch-CPU fp32 restatement on all host cores; only the number of timed steps differs
    (the GPU arm times CPU_STEPS after CPU_WARMUP, --impl reference times the --steps / --warmup it is given)."""
    cores = os.cpu_count() or 1
    value, ms = cpu_train_images_per_sec(args.decoder, args.size, args.channels, args.batch, steps, warmup, cores)
    return {"value": value, "unit": "images/s", "cores": cores, "kind": "port", "cpu_model": cpu_model_name(),
            "ms_per_step": ms,
            "sample": f"{steps} timed steps of {args.batch} images ({args.size}x{args.size}x{args.channels}, "
                      f"{args.decoder} decoder: the full per-GPU batch of the workload) after {warmup} warm-up steps; "
                      "torch-CPU fp32 (oneDNN) restatement of the reference's train step (oracle/unet_oracle.py), "
                      "1 process, no parameter server / gRPC; TF 1.4 is not installable"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    base = cpu_baseline_block(args, max(args.steps, 1), max(args.warmup, 1))
    value, ms = base["value"], base["ms_per_step"]
    line = {
        "impl": "reference", "metric": "train images/sec", "value": value, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, max(args.gpus, 1)),
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------------- GPU arm
def measure_decoder(args, decoder, world, rank, dev, sampler, with_roofline):
    """Device-resident and end-to-end timing of one decoder variant (+ the per-kernel-class roofline timing)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from unet_distributed_b200 import _lib
    from unet_distributed_b200.synthetic import make_slices
    from unet_distributed_b200.trainer import Trainer
    N, S, C = args.batch, args.size, args.channels
    ups = decoder == "upsampling"
    tr = Trainer(N, S, S, C, 1, ups, dropout=0.2, learning_rate=0.0005,
                 const_learningrate=bool(args.const_learningrate), lr_fraction=args.lr_fraction,
                 decay_steps=args.decay_steps, use_cuda_graph=(not args.no_graph), precision=args.precision)
    # synthetic batches: a few distinct ones per rank, rotated (inputs are far smaller than the per-step working set)
    nb = 4
    base_imgs, base_msks = make_slices(min(N, 64), S, S, C, seed=816 + rank)
    reps = -(-N // base_imgs.shape[0])
    h_batches = []
    for b in range(nb):
        im = np.roll(np.tile(base_imgs, (reps, 1, 1, 1))[:N], b, axis=0).copy()
        mk = np.roll(np.tile(base_msks, (reps, 1, 1, 1))[:N], b, axis=0).copy()
        h_batches.append((torch.from_numpy(im).pin_memory(), torch.from_numpy(mk).pin_memory()))
    d_batches = [(a.to(dev), b.to(dev)) for a, b in h_batches]
    L = _lib.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    out = {"decoder": decoder}
    # ---- device-resident timing
    for i in range(args.warmup):
        tr.train_step(*d_batches[i % nb])
    barrier()
    launches0 = L.unetb200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if sampler is not None:
        sampler.begin()
    e0.record()
    for i in range(args.steps):
        tr.train_step(*d_batches[i % nb])
    e1.record()
    barrier()
    if sampler is not None:
        sampler.end()
    launches = L.unetb200_launch_count() - launches0
    if tr.use_cuda_graph:   # replays bypass the library's counter: kernels per captured step x steps
        launches = tr.graph_launches_per_step * args.steps
    ms_total = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    out["ms_step"] = float(ms_total.item()) / args.steps
    out["launches"] = int(launches)
    out["metrics_last"] = tr.plan.metrics.cpu().numpy()

    # ---- end-to-end timing through the host-facing call (pinned host -> device copy and loss read-back per step)
    # software-pipelined: the pinned-host -> device copy of step k+1 is enqueued on a copy stream while step k computes;
    # every step's [loss, dice, sens, spec] is read back (the call for step k+1 returns step k's numbers)
    for i in range(max(3, args.warmup // 2)):
        tr.train_step_pinned(*h_batches[i % nb], wait=False)
    tr.finish()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(args.steps):
        tr.train_step_pinned(*h_batches[i % nb], wait=False)
    e2e_last = tr.finish()
    e3.record()
    barrier()
    ms_e2e = torch.tensor([e2.elapsed_time(e3)], device=dev)
    if world > 1:
        dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
    out["ms_e2e_step"] = float(ms_e2e.item()) / args.steps
    out["e2e_last_loss"] = float(e2e_last[0])
    out["h2d"], out["d2h"] = tr.h2d_bytes_per_step(), tr.d2h_bytes_per_step()

    # ---- per-kernel-class timing for the roofline (events around every tile-engine launch; graph replay off)
    # every rank runs these steps (they contain the gradient all-reduce); only rank 0's records are reported
    if with_roofline:
        if tr.use_cuda_graph:
            tr.use_cuda_graph = False
        # kernels are timed one at a time: in the timed region above the weight gradients share the GPU with the
        # dgrad chain (second stream), which would charge each kernel for the time it waits for SMs
        tr.plan.side_stream(False)
        tr.serial_update = True        # ... and so would the bucket updates (Adam, re-pack) of the update stream
        for i in range(2):
            tr.train_step(*d_batches[i % nb])
        torch.cuda.synchronize()
        L.unetb200_profile_enable(1)
        psteps = 3
        for i in range(psteps):
            tr.train_step(*d_batches[i % nb])
        ms = (ctypes.c_double * 3)()
        fl = (ctypes.c_double * 3)()
        by = (ctypes.c_double * 3)()
        ln = (ctypes.c_longlong * 3)()
        _lib.check(L.unetb200_profile_collect(ms, fl, by, ln), "profile_collect")
        L.unetb200_profile_enable(0)
        tr.plan.side_stream(True)
        tr.serial_update = False
        out["roof"] = {"igemm": (ms[0], fl[0], by[0], ln[0]), "wgrad": (ms[1], fl[1], by[1], ln[1]), "steps": psteps}
        barrier()

    # ---- the gradient all-reduce alone (the buckets back to back), for bus bandwidth against NVLink
    out["allreduce"] = None
    if world > 1 and with_roofline:
        nbk = len(tr.reducer.buckets)
        for _ in range(3):
            for seg in range(nbk):
                tr.reducer.launch(seg)
            tr.reducer.wait()
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        a0.record()
        for _ in range(reps):
            for seg in range(nbk):
                tr.reducer.launch(seg)
            tr.reducer.wait()
        a1.record()
        barrier()
        t_ar = torch.tensor([a0.elapsed_time(a1) / reps], device=dev)
        dist.all_reduce(t_ar, op=dist.ReduceOp.MAX)
        nbytes = tr.reducer.bytes_per_step()
        alg = nbytes / (float(t_ar.item()) * 1e-3) / 1e9
        out["allreduce"] = {"bytes_per_step": nbytes, "ms_back_to_back": float(t_ar.item()), "alg_gbs": alg,
                            "bus_gbs": alg * 2 * (world - 1) / world, "peak": 900, "unit": "GB/s",
                            "frac": alg * 2 * (world - 1) / world / 900.0,
                            "peak_source": "NVLink 5 nominal per direction per GPU", "buckets": nbk,
                            "note": "latency-bound: %.1f MB per step in %d buckets; inside the step the buckets "
                                    "overlap the backward segments that follow them" % (nbytes / 1e6, nbk)}
    tr.close()     # the captured step holds NCCL kernels: it has to go before the process group is destroyed
    del tr
    torch.cuda.empty_cache()
    return out


def measure_tf32_gemm_tflops(dev):
    """Denominator for the fp32/tf32 mode: torch.matmul (cuBLAS) on fp32 8192^3 with tf32 allowed, best of 10."""
    import torch
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(8192, 8192, device=dev)
        b = torch.randn(8192, 8192, device=dev)
        for _ in range(3):
            a @ b
        best = float("inf")
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * 8192 ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def run_b200(args):
    import torch
    import torch.distributed as dist
    from unet_distributed_b200 import parallel

    world, rank, local = parallel.init_from_env("nccl")
    if world == 1:
        torch.cuda.set_device(0)
    dev = torch.device("cuda", torch.cuda.current_device())
    sampler = ClockSampler(torch.cuda.current_device())
    if rank == 0:
        sampler.start()   # child process; it needs a moment to load NVML, so it starts before the model is built
    N, S, C = args.batch, args.size, args.channels
    main = measure_decoder(args, args.decoder, world, rank, dev, sampler, True)
    clocks = sampler.stop() if rank == 0 else None
    # the reference's DEFAULT decoder is Conv2DTranspose (settings_dist.py:24, USE_UPSAMPLING = False): its numbers ride
    # along in every line whose headline workload is the UpSampling2D decoder (BASELINE configs[1])
    other = None
    if args.decoder == "upsampling" and not args.no_tconv_block:
        other = measure_decoder(args, "tconv", world, rank, dev, None, False)

    if rank == 0:
        peaks = load_peaks()
        gb = N * world
        ms_step, ms_e2e_step = main["ms_step"], main["ms_e2e_step"]
        value = gb / (ms_step * 1e-3)
        e2e_value = gb / (ms_e2e_step * 1e-3)
        gflop_img = TRAIN_GFLOP_PER_IMG.get((args.decoder, S, C))
        roof = main["roof"]
        ig_ms, ig_fl, ig_by, ig_n = roof["igemm"]
        wg_ms, wg_fl, wg_by, wg_n = roof["wgrad"]
        peak_tf = float(peaks.get("bf16_tflops_sustained", FALLBACK_PEAKS["bf16_tflops_sustained"]))
        peak_src = "bf16_tflops_sustained, " + peaks["_source"]
        if args.precision == "tf32":
            peak_tf = measure_tf32_gemm_tflops(dev)
            peak_src = "cuBLAS tf32 GEMM 8192^3 measured in this process (best of 10), no MEASURED_PEAKS.json entry"
        achieved = ig_fl / (ig_ms * 1e-3) / 1e12 if ig_ms > 0 else 0.0
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get("igemm_dram_bytes_per_launch")
            except Exception:
                traffic = None
        cfg = workload_config(args, world)
        line = {
            "metric": "train images/sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": cfg,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "images/s", "ms_per_step": ms_e2e_step,
                    "h2d_bytes_per_step": main["h2d"], "d2h_bytes_per_step": main["d2h"],
                    "mode": "Trainer.train_step_pinned(wait=False): H2D of step k+1 overlaps step k on a copy "
                            "stream, every step's metrics are read back; last step's loss %.4f" % main["e2e_last_loss"]},
            "gpu_launches": int(main["launches"]),
            "roofline": {
                "kernel": "conv_halo_rp_kernel + conv_halo_kernel + igemm_kernel (3x3 conv / transposed conv forward + dgrad, tcgen05 "
                          "implicit GEMM; every launch of the class in a step)",
                "bound": "tensor", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": achieved / peak_tf if peak_tf else None, "traffic": traffic,
                "peak_source": peak_src,
                "launches_per_step": ig_n / roof["steps"], "ms_per_step": ig_ms / roof["steps"],
                "alg_gbs": (ig_by / (ig_ms * 1e-3) / 1e9 if ig_ms > 0 else 0.0),
                "hbm_peak_gbs": float(peaks.get("hbm_gbs", FALLBACK_PEAKS.get("hbm_gbs", 0.0))),
                "wgrad": {"achieved": (wg_fl / (wg_ms * 1e-3) / 1e12 if wg_ms > 0 else 0.0), "unit": "TFLOP/s",
                          "launches_per_step": wg_n / roof["steps"], "ms_per_step": wg_ms / roof["steps"]},
                "whole_step": ({"achieved": value * gflop_img / 1e3 / world, "unit": "TFLOP/s per GPU",
                                "frac": value * gflop_img / 1e3 / world / peak_tf} if gflop_img else None),
                "collective": main["allreduce"]},
            "loss": float(main["metrics_last"][0]), "dice": float(main["metrics_last"][1]),
        }
        if other is not None:
            gf = TRAIN_GFLOP_PER_IMG.get(("tconv", S, C))
            v = gb / (other["ms_step"] * 1e-3)
            line["tconv"] = {
                "what": "the same run with the reference's default decoder (Conv2DTranspose, settings_dist.py:24; "
                        "BASELINE configs[0] shape), measured in this process right after the headline workload",
                "value": v, "unit": "images/s", "ms_per_step": other["ms_step"],
                "whole_step_frac": (v * gf / 1e3 / world / peak_tf) if gf else None,
                "e2e": {"value": gb / (other["ms_e2e_step"] * 1e-3), "unit": "images/s",
                        "ms_per_step": other["ms_e2e_step"], "h2d_bytes_per_step": other["h2d"],
                        "d2h_bytes_per_step": other["d2h"]},
                "gpu_launches": int(other["launches"]), "loss": float(other["metrics_last"][0])}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_block(args, CPU_STEPS, CPU_WARMUP)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else this process (or NCCL) prints goes to stderr."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
