"""All five BASELINE configs through the C ABI on one GPU: CUDA-event times, algorithmic GB/s, parity
spot checks.  Writes one JSON line per config.
    python tools/measure_all.py > gpurun_out/all_configs.jsonl"""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

CFGS = [("cfg1", (4, 32, 64), 1), ("cfg2", (16, 64, 128, 128), 2), ("cfg3", (8, 768, 90, 180), 2),
        ("cfg4", (64, 20, 720, 1440), 2), ("cfg5", (4, 8192, 8192), 2)]
for name, shape, nd in CFGS:
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
    # forward parity on the first image(s) against fp64 torch.fft on the CPU (the reference tests' oracle)
    sl = tuple([0] * (len(shape) - nd))
    want = torch.view_as_real(torch.fft.rfftn(x[sl].double().cpu(), dim=tuple(range(-nd, 0))))
    fe = ((y[sl].cpu().double() - want).norm() / want.norm()).item()
    n = 200 if x.numel() < 5e8 else 20
    res = {}
    for key, fn in (("fwd", fw), ("inv", iv), ("pair", lambda: (fw(), iv()))):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        res[key] = e0.elapsed_time(e1) / n
    algo = (x.numel() + y.numel()) * 4
    print(json.dumps({"config": name, "shape": list(shape), "signal_ndim": nd, "fwd_ms": res["fwd"], "inv_ms": res["inv"],
                      "pair_ms": res["pair"], "fwd_GBs": algo / res["fwd"] / 1e6, "inv_GBs": algo / res["inv"] / 1e6,
                      "pair_GBs": 2 * algo / res["pair"] / 1e6, "round_trip_rel_l2": rt, "fwd_rel_l2_vs_fp64": fe,
                      "plan_fwd": f.describe()[1], "plan_inv": i.describe()[1]}), flush=True)
    del x, y, xr, f, i, ws
    torch.cuda.empty_cache()
