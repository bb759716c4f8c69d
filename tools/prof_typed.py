"""Runs rfft2 + irfft2 over 16-bit tensors (trt_dft_plan_create_typed) a few times for one shape (target for ncu).
    python tools/prof_typed.py 8 768 90 180 --io f16 --iters 3"""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

ap = argparse.ArgumentParser()
ap.add_argument("shape", type=int, nargs="+")
ap.add_argument("--io", default="f16")
ap.add_argument("--iters", type=int, default=3)
a = ap.parse_args()
dt = {"f16": torch.float16, "bf16": torch.bfloat16}[a.io]
shape = tuple(a.shape)
batch = 1
for s in shape[:-2]:
    batch *= s
dims = shape[-2:]
torch.manual_seed(1)
x = torch.randn(*shape, device="cuda").to(dt)
y = torch.empty(shape[:-1] + (dims[-1] // 2 + 1, 2), device="cuda", dtype=dt)
xr = torch.empty_like(x)
f = capi.TypedPlan(capi.FORWARD, dims, batch, dt)
i = capi.TypedPlan(capi.INVERSE, dims, batch, dt)
for _ in range(a.iters):
    f.exec(x, y)
    i.exec(y, xr)
torch.cuda.synchronize()
print(f.describe(), i.describe())
print("round trip rel-L2", ((xr.double() - x.double()).norm() / x.double().norm()).item())
