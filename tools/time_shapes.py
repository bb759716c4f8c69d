"""CUDA-event times of rfftN / irfftN for a list of shapes through the C ABI (A/B checks of kernel changes).
    python tools/time_shapes.py "2048x60x120:2" "4x8192x8192:2" ..."""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

for spec in sys.argv[1:]:
    dims_s, nd = spec.split(":")
    shape = tuple(int(v) for v in dims_s.split("x"))
    nd = int(nd)
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
    ws = torch.empty(max(f.workspace_bytes, i.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    fw = lambda: f.exec_ptr(x.data_ptr(), y.data_ptr(), ws.data_ptr(), st)
    iv = lambda: i.exec_ptr(y.data_ptr(), xr.data_ptr(), ws.data_ptr(), st)
    for _ in range(3):
        fw(); iv()
    torch.cuda.synchronize()
    rt = ((xr - x).double().norm() / x.double().norm()).item()
    n = 200 if x.numel() < 3e8 else 20
    res = {}
    for key, fn in (("fwd", fw), ("inv", iv)):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        res[key] = e0.elapsed_time(e1) / n
    algo = (x.numel() + y.numel()) * 4
    print(json.dumps({"shape": spec, "fwd_ms": round(res["fwd"], 5), "inv_ms": round(res["inv"], 5), "fwd_GBs": round(algo / res["fwd"] / 1e6, 1),
                      "inv_GBs": round(algo / res["inv"] / 1e6, 1), "rt": rt, "plan_fwd": f.describe()[1][:160], "plan_inv": i.describe()[1][:160]}), flush=True)
    del x, y, xr, ws, f, i
    torch.cuda.empty_cache()
