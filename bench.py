#!/usr/bin/env python
"""bench.py -- headline benchmark of the Rfft/Irfft hot path on B200.

Metric (BASELINE.json): rfft2+irfft2 HBM GB/s and transforms/s at the FourCastNet AFNO shape
[8,768,90,180] (configs[2]; fits one GPU).  One step = one rfft2 followed by one irfft2 over the
whole [8,768,90,180] fp32 batch, both through the C ABI (include/trt_dft.h).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

* `value`   : algorithmic GB/s of the pair (read input once + write output once, both directions:
              2 x 800,686,080 B per step), inputs resident in HBM, CUDA-event timed, max over ranks.
* `e2e`     : same metric through trt_dft_exec_host with pinned HOST buffers (H2D + kernels + D2H
              inside the timed region) -- the call a TensorRT-less user makes.
* `roofline`: slower of the two kernels; achieved = 800,686,080 B / its CUDA-event launch time,
              peak = MEASURED_PEAKS.json hbm_gbs (fallback 6650 GB/s, B200_PROFILING.md).
* `cpu_baseline`: the reference tests' oracle (torch.fft, tests/test_dft.py:40,:52-54) on this
              box's host cores, on a bounded batch slice.
* `--impl reference`: the reference's own CPU path for this op (same torch.fft expressions; the
              cuFFT-backed plugin cannot be built without TensorRT), all host threads.
Multi-GPU (torchrun): every rank transforms its own [8,768,90,180] shard of the batch x channel
dimension (weak scaling, no collective in the data path).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
sys.path.insert(0, str(ROOT))

SHAPE = (8, 768, 90, 180)
H, W = 90, 180
MO = W // 2 + 1
METRIC = "rfft2+irfft2 algorithmic HBM GB/s at AFNO [8,768,90,180] fp32"


def algo_bytes_one_direction(nimg):
    return nimg * (H * W * 4 + H * MO * 8)


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc = None
        self.lines = []
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_pair(x):
    """The reference tests' oracle expressions (tests/test_dft.py:40, :52-54) as one pair."""
    from oracle import oracle as O
    y = O.torch_rfftn(x, 2)
    return O.torch_irfftn(y, 2)


def time_cpu_pair(nimg_sample, repeats, threads):
    import torch
    torch.set_num_threads(threads)
    torch.manual_seed(1)
    x = torch.randn(nimg_sample, H, W)
    cpu_reference_pair(x)                      # warm-up (MKL plan creation)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        cpu_reference_pair(x)
        best = min(best, time.perf_counter() - t0)
    return best


def run_reference(args):
    """--impl reference: the reference's CPU path (torch.fft), rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    nimg = 768                                  # 1/8 of the batch per step: a bounded sample
    torch.set_num_threads(cores)
    torch.manual_seed(1)
    x = torch.randn(nimg, H, W)
    for _ in range(max(args.warmup, 1)):
        cpu_reference_pair(x)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_pair(x)
    dt = time.perf_counter() - t0
    gbs = 2 * algo_bytes_one_direction(nimg) * args.steps / dt / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "AFNO rfft2+irfft2 [8,768,90,180] fp32 (configs[2])", "sample": f"[{nimg},90,180] per step"},
        "pairs_per_s": nimg * args.steps / dt,
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port",
                         "sample": f"torch.fft rfft2+irfft2 on [{nimg},90,180] fp32 per step, {args.steps} steps; "
                                   "the reference tests' CPU oracle (its cuFFT plugin needs TensorRT, absent)"},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from trt_dft_plugins import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); run it under gpurun")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    capi.lib()

    nimg = SHAPE[0] * SHAPE[1]
    fwd = capi.Plan(capi.FORWARD, (H, W), nimg)
    inv = capi.Plan(capi.INVERSE, (H, W), nimg)
    torch.manual_seed(1 + rank)
    x = torch.randn(*SHAPE, device=dev)
    y = torch.empty(SHAPE[:-1] + (MO, 2), device=dev)
    xr = torch.empty(*SHAPE, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    launches_per_step = fwd.num_launches + inv.num_launches

    def step():
        fwd.exec_ptr(x.data_ptr(), y.data_ptr(), 0, stream)
        inv.exec_ptr(y.data_ptr(), xr.data_ptr(), 0, stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # sanity: the round trip must reproduce x (guards against timing a broken kernel)
    rt = ((xr - x).double().norm() / x.double().norm()).item()
    if not rt <= 1e-6:
        raise SystemExit(f"round-trip check failed before timing: rel-L2 {rt}")

    # nvidia-smi sampling covers the timed region and the per-kernel timing loops that follow it (same
    # kernels, same inputs); the sampler needs ~0.2 s to deliver its first line, so it starts a little early.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
        for _ in range(3):
            step()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    # per-kernel launch durations over the same kind of region (device events on the launch stream)
    kms = {}
    for name, fn in (("fused_rfft2", lambda: fwd.exec_ptr(x.data_ptr(), y.data_ptr(), 0, stream)),
                     ("fused_irfft2", lambda: inv.exec_ptr(y.data_ptr(), xr.data_ptr(), 0, stream))):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(args.steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        kms[name] = a.elapsed_time(b) / args.steps
    clocks = sampler.stop() if rank == 0 else None

    # end to end through the host-buffer C ABI call
    xh = torch.randn(*SHAPE).pin_memory()
    yh = torch.empty(SHAPE[:-1] + (MO, 2)).pin_memory()
    xrh = torch.empty(*SHAPE).pin_memory()
    e2e_steps = max(1, min(args.steps, 5))
    fwd.exec_host(xh, yh)
    inv.exec_host(yh, xrh)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        fwd.exec_host(xh, yh)
        inv.exec_host(yh, xrh)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0

    if world > 1:
        t = torch.tensor([ms, e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = t[0].item(), t[1].item()

    if rank == 0:
        pair_bytes = 2 * algo_bytes_one_direction(nimg)
        gbs = world * pair_bytes * args.steps / (ms * 1e-3) / 1e9
        peak, peak_src = measured_peak()
        dom = max(kms, key=kms.get)
        ach = algo_bytes_one_direction(nimg) / (kms[dom] * 1e-3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "traffic.json"
        if tf.exists():
            try:
                traffic = json.loads(tf.read_text()).get(dom)
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "AFNO rfft2+irfft2 [8,768,90,180] fp32 per GPU (configs[2])",
                       "parallelism": f"batch-shard x{world}, no collective",
                       "l2": "inputs (398 MB + 403 MB per direction) exceed the 126 MB L2; no flush needed"},
            "pairs_per_s": world * nimg * args.steps / (ms * 1e-3),
            "frac_of_8TBs_spec": gbs / world / 8000.0,
            "kernel_ms": kms,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo_bytes_one_direction(nimg)},
            "e2e": {"value": world * pair_bytes * e2e_steps / e2e_s / 1e9, "unit": "GB/s",
                    "h2d_bytes_per_step": (x.numel() + y.numel()) * 4, "d2h_bytes_per_step": (y.numel() + xr.numel()) * 4,
                    "steps": e2e_steps, "api": "trt_dft_exec_host (pinned host buffers)"},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks,
            "plan": [fwd.describe()[1], inv.describe()[1]],
        }
        if world == 1:
            cores = os.cpu_count() or 1
            sample = 768
            t = time_cpu_pair(sample, 3, cores)
            line["cpu_baseline"] = {"value": 2 * algo_bytes_one_direction(sample) / t / 1e9, "unit": "GB/s", "cores": cores,
                                    "kind": "port",
                                    "sample": f"torch.fft rfft2+irfft2 (reference tests' oracle) on [{sample},90,180] fp32, best of 3"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
