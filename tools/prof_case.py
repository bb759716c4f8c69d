"""Runs rfftN + irfftN a few times through the C ABI for one shape (target for ncu / timing).
    python tools/prof_case.py 8 768 90 180 --nd 2 --iters 3"""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

ap = argparse.ArgumentParser()
ap.add_argument("shape", type=int, nargs="+")
ap.add_argument("--nd", type=int, default=2)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--time", action="store_true")
a = ap.parse_args()
shape = tuple(a.shape)
nd = a.nd
batch = 1
for s in shape[:-nd]:
    batch *= s
dims = shape[-nd:]
torch.manual_seed(1)
x = torch.randn(*shape, device="cuda")
y = torch.empty(shape[:-1] + (dims[-1] // 2 + 1, 2), device="cuda")
xr = torch.empty_like(x)
f = capi.Plan(capi.FORWARD, dims, batch)
i = capi.Plan(capi.INVERSE, dims, batch)
for _ in range(a.iters):
    f.exec(x, y)
    i.exec(y, xr)
torch.cuda.synchronize()
print(f.describe(), i.describe())
print("round trip rel-L2", ((xr - x).double().norm() / x.double().norm()).item())
if a.time:
    algo = (x.numel() + y.numel()) * 4
    for name, fn in (("fwd", lambda: f.exec(x, y)), ("inv", lambda: i.exec(y, xr))):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 50
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(f"{name}: {ms:.4f} ms  {algo / ms / 1e6:.1f} GB/s algorithmic")
