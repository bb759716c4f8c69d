"""Truncated-mode kernels at the FNO shape (1024 images of 128 x 128) against the full kernels, per direction."""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
H, W, N = 128, 128, 1024
x = torch.randn(N, H, W, device="cuda"); xr = torch.empty_like(x)
st = torch.cuda.current_stream().cuda_stream
def t(fn, iters=300):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
f0, i0 = capi.Plan(capi.FORWARD, (H, W), N), capi.Plan(capi.INVERSE, (H, W), N)
y0 = torch.empty(N, H, 65, 2, device="cuda")
print(json.dumps({"full": True, "fwd_ms": t(lambda: f0.exec_ptr(x.data_ptr(), y0.data_ptr(), 0, st)), "inv_ms": t(lambda: i0.exec_ptr(y0.data_ptr(), xr.data_ptr(), 0, st))}))
for kh, kw in ((64, 65), (32, 32), (16, 16), (12, 12), (4, 4)):
    f, i = capi.TruncatedPlan(capi.FORWARD, (H, W), N, kh, kw), capi.TruncatedPlan(capi.INVERSE, (H, W), N, kh, kw)
    y = torch.randn(N, 2 * kh, kw, 2, device="cuda")
    print(json.dumps({"keep": [2 * kh, kw], "fwd_ms": t(lambda: f.exec_ptr(x.data_ptr(), y.data_ptr(), 0, st)),
                      "inv_ms": t(lambda: i.exec_ptr(y.data_ptr(), xr.data_ptr(), 0, st)), "route": f.describe()[1][-60:]}), flush=True)
