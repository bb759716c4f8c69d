"""CUDA-event time of the passes that make up rfft2 / irfft2 on [B, 8192, 8192], each as a plan of its own:
the row pass (1-D R2C / C2R over B*8192 rows) and the column pass (C2C of length 8192 over 4097 columns: two launches,
four-step), next to a device copy of the same bytes.
    python tools/cfg5_passes.py [B]"""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
Mo = N // 2 + 1
x = torch.randn(B, N, N, device="cuda")
y = torch.empty(B, N, Mo, 2, device="cuda")
y2 = torch.empty_like(y)
xr = torch.empty_like(x)
st = torch.cuda.current_stream().cuda_stream


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out = {"shape": [B, N, N]}
rows = capi.Plan(capi.FORWARD, (N,), B * N)
irows = capi.Plan(capi.INVERSE, (N,), B * N)
cols = capi.C2CPlan(capi.FORWARD, N, B, Mo)
icols = capi.C2CPlan(capi.INVERSE, N, B, Mo)
full_f = capi.Plan(capi.FORWARD, (N, N), B)
full_i = capi.Plan(capi.INVERSE, (N, N), B)
wsb = max(p.workspace_bytes for p in (rows, irows, cols, icols, full_f, full_i))
ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device="cuda")
out["rows_r2c_ms"] = timed(lambda: rows.exec_ptr(x.data_ptr(), y.data_ptr(), ws.data_ptr(), st))
out["cols_c2c_ms"] = timed(lambda: cols.exec_ptr(y.data_ptr(), y2.data_ptr(), ws.data_ptr(), st))
out["cols_c2c_inplace_ms"] = timed(lambda: cols.exec_ptr(y.data_ptr(), y.data_ptr(), ws.data_ptr(), st))
out["icols_c2c_ms"] = timed(lambda: icols.exec_ptr(y.data_ptr(), y2.data_ptr(), ws.data_ptr(), st))
out["rows_c2r_ms"] = timed(lambda: irows.exec_ptr(y.data_ptr(), xr.data_ptr(), ws.data_ptr(), st))
out["rfft2_ms"] = timed(lambda: full_f.exec_ptr(x.data_ptr(), y.data_ptr(), ws.data_ptr(), st))
out["irfft2_ms"] = timed(lambda: full_i.exec_ptr(y.data_ptr(), xr.data_ptr(), ws.data_ptr(), st))
out["copy_complex_ms"] = timed(lambda: y2.copy_(y))
out["plans"] = {"rows": rows.describe()[1][:200], "cols": cols.describe()[1][:300], "full_f": full_f.describe()[1][:400]}
print(json.dumps(out))
