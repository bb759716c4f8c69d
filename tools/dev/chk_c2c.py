import sys, os
sys.path.insert(0, os.path.join(os.getcwd(), "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
torch.manual_seed(0)
for (outer, n, inner) in [(600, 128, 64), (300, 64, 128), (2, 8192, 300), (2, 8192, 64)]:
    for d in (0, 1):
        z = torch.randn(outer, n, inner, 2, device="cuda")
        zc = torch.view_as_complex(z.double())
        want = torch.fft.fft(zc, dim=1) if d == 0 else torch.fft.ifft(zc, dim=1, norm="forward")
        plan = capi.C2CPlan(d, n, outer, inner, scale=1.0)
        out = torch.empty_like(z)
        plan.exec(z, out); torch.cuda.synchronize()
        got = torch.view_as_complex(out.double())
        err = (got - want).abs().amax(dim=(0, 2)) / want.abs().mean()
        bad = (err > 1e-3).nonzero().flatten()
        print(outer, n, inner, "dir", d, "rel", ((got - want).norm() / want.norm()).item(), "bad rows", bad[:12].tolist(), len(bad), plan.describe()[1][:150])
