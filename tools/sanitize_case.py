"""Small run that touches every kernel family once (for compute-sanitizer on the GPU box)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
CASES = [((3, 90, 180), 2),        # pipelined specialised fused kernels
         ((5, 128, 128), 2),       # FusedCfg kernels
         ((4, 24, 40), 2),         # generic fused tile kernel
         ((150, 3, 180, 360), 2),  # rows + cols, pipelined generic shell (>= 148 tiles)
         ((300, 2048), 1),         # skewed power-of-two rows, pipelined
         ((7, 30), 1), ((5, 9), 1), ((2, 17, 6), 2), ((2, 4, 6, 8), 3),
         ((1, 4096, 96), 2)]       # four-step column pass
for shape, nd in CASES:
    torch.manual_seed(1)
    x = torch.randn(*shape, device="cuda")
    y = capi.rfftn(x, nd)
    xr = capi.irfftn(y, nd, last=shape[-1])
    torch.cuda.synchronize()
    err = ((xr - x).double().norm() / x.double().norm()).item()
    print(shape, nd, f"round trip {err:.2e}")
    assert err < 1e-6
print("ok")
