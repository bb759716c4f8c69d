"""Times the two passes of the un-fused 2-D path separately (rows R2C plan, strided-axis C2C plan).
    python tools/time_passes.py B H W"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
B, H, W = map(int, sys.argv[1:4])
Mo = W // 2 + 1
x = torch.randn(B, H, W, device="cuda")
y = torch.empty(B, H, Mo, 2, device="cuda")
rows = capi.Plan(capi.FORWARD, (W,), B * H)
cols = capi.C2CPlan(capi.FORWARD, H, B, Mo)
irows = capi.Plan(capi.INVERSE, (W,), B * H)
print(rows.describe()[1]); print(cols.describe()[1])
def t(fn, n=30):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
rb = (x.numel() + y.numel()) * 4
cb = 2 * y.numel() * 4
tr = t(lambda: rows.exec(x, y)); tc = t(lambda: cols.exec(y, y)); ti = t(lambda: irows.exec(y, x))
print(f"rows r2c : {tr:.4f} ms  {rb/tr/1e6:.0f} GB/s")
print(f"cols c2c : {tc:.4f} ms  {cb/tc/1e6:.0f} GB/s")
print(f"rows c2r : {ti:.4f} ms  {rb/ti/1e6:.0f} GB/s")
