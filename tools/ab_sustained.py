"""A/B of kernel variants under sustained load: alternates subprocesses (one per env setting), each running
the rfft2+irfft2 pair `steps` times at cfg3 and timing each direction alone as well.
    python tools/ab_sustained.py TRT_DFT_REG2D_VARIANT 0 1 2 3"""
import os
import subprocess
import sys

CHILD = r'''
import sys, os
sys.path.insert(0, os.path.join(os.getcwd(), "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
x = torch.randn(8, 768, 90, 180, device="cuda"); y = torch.empty(8, 768, 90, 91, 2, device="cuda"); xr = torch.empty_like(x)
f = capi.Plan(capi.FORWARD, (90, 180), 6144); i = capi.Plan(capi.INVERSE, (90, 180), 6144)
def t(fn, n):
    for _ in range(20): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
n = int(sys.argv[1])
def pair():
    f.exec(x, y); i.exec(y, xr)
print(f"pair {t(pair, n):.4f}  fwd {t(lambda: f.exec(x, y), n):.4f}  inv {t(lambda: i.exec(y, xr), n):.4f} ms")
'''
var, vals = sys.argv[1], sys.argv[2:]
for rep in range(2):
    for v in vals:
        env = dict(os.environ); env[var] = v
        out = subprocess.run([sys.executable, "-c", CHILD, "2000"], env=env, capture_output=True, text=True)
        print(f"{var}={v}: {out.stdout.strip()} {out.stderr.strip()[-200:]}")
