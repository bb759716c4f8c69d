"""Pencil-decomposed rfft2 over N GPUs (torchrun): correctness against a single-GPU transform of the
same images and device-timed throughput (max over ranks).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/pencil_bench.py --shape 4 8192 8192 --iters 10"""
import argparse, json, os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
import torch.distributed as dist
from trt_dft_plugins import capi
from trt_dft_plugins.pencil import PencilRfft2
from trt_dft_plugins.multi_gpu import max_over_ranks

ap = argparse.ArgumentParser()
ap.add_argument("--shape", type=int, nargs=3, default=[4, 8192, 8192])
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--check", action="store_true")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
B, H, W = a.shape
p = PencilRfft2(B, H, W, device=dev)
g = torch.Generator(device=dev); g.manual_seed(1)            # same full image on every rank, then slice
x_full = torch.randn(B, H, W, device=dev, generator=g) if a.check else None
x_loc = (x_full[:, p.h0:p.h0 + p.hn].contiguous() if a.check else torch.randn(B, p.hn, W, device=dev))
y = p.forward(x_loc)
torch.cuda.synchronize()
err = None
if a.check:
    ref = capi.rfftn(x_full, 2)[:, :, p.c0:p.c0 + p.cn]
    err = ((y - ref).double().norm() / ref.double().norm()).item()
# phases timed separately and together
def timed(fn, n):
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return max_over_ranks(e0.elapsed_time(e1) / n, dev)
yr = p.row_pass(x_loc); z = p.transpose(yr)
t_all = timed(lambda: p.forward(x_loc), a.iters)                                   # fused-pack, per-image all-to-all, overlapped
t_unfused = timed(lambda: p.col_pass(p.transpose(p.row_pass(x_loc))), a.iters)      # separate pack / unpack copies
t_rows = timed(lambda: p.row_pass(x_loc), a.iters)
t_a2a = timed(lambda: p.transpose(yr), a.iters)
t_cols = timed(lambda: p.col_pass(z), a.iters)
if rank == 0:
    algo = B * (H * W * 4 + H * (W // 2 + 1) * 8)
    print(json.dumps({"pencil_rfft2": [B, H, W], "n_gpus": world, "ms": t_all, "unfused_ms": t_unfused, "rows_ms": t_rows, "transpose_ms": t_a2a,
                      "cols_ms": t_cols, "algorithmic_GBs": algo / t_all / 1e6, "rel_l2_vs_single_gpu": err}))
if world > 1:
    dist.destroy_process_group()
