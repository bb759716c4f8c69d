"""e2e (host buffers in/out) pair time at cfg3 for a few pipeline chunk counts (TRT_DFT_HOST_CHUNKS)."""
import os, subprocess, sys
CHILD = r'''
import sys, os, time
sys.path.insert(0, os.path.join(os.getcwd(), "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
x = torch.randn(8, 768, 90, 180).pin_memory(); y = torch.empty(8, 768, 90, 91, 2).pin_memory(); xr = torch.empty_like(x).pin_memory()
f = capi.Plan(capi.FORWARD, (90, 180), 6144); i = capi.Plan(capi.INVERSE, (90, 180), 6144)
for _ in range(2): f.exec_host(x, y); i.exec_host(y, xr)
t0 = time.perf_counter(); n = 6
for _ in range(n): f.exec_host(x, y); i.exec_host(y, xr)
dt = (time.perf_counter() - t0) / n
print(f"{dt*1e3:.2f} ms per pair, {2*800686080/dt/1e9:.1f} GB/s algorithmic, round trip {((xr-x).norm()/x.norm()).item():.2e}")
'''
for v in sys.argv[1:]:
    env = dict(os.environ); env["TRT_DFT_HOST_CHUNKS"] = v
    out = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    print(f"chunks={v}: {out.stdout.strip()} {out.stderr.strip()[-300:]}")
