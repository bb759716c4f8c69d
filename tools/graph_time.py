"""GPU time per launch of a tiny problem without host launch overhead: capture N pairs in a CUDA graph."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
shape, nd = (4, 32, 64), 1
x = torch.randn(*shape, device="cuda"); y = torch.empty(4, 32, 33, 2, device="cuda"); xr = torch.empty_like(x)
f = capi.Plan(capi.FORWARD, (64,), 128); i = capi.Plan(capi.INVERSE, (64,), 128)
f.exec(x, y); i.exec(y, xr); torch.cuda.synchronize()
N = 200
for name, fn in (("fwd", lambda: f.exec(x, y)), ("inv", lambda: i.exec(y, xr))):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(N):
            fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f"cfg1 {name}: {e0.elapsed_time(e1) / N * 1e3:.2f} us per launch inside a CUDA graph ({f.describe()[1] if name == 'fwd' else i.describe()[1]})")
