"""How fast are the 720-point passes when their data is L2-resident?  Times the column pass (C2C, in place) and the
row pass (R2C) on batches small enough to stay in the 126 MB L2, replayed from a CUDA graph, against the full batch.
    python tools/l2_resident_probe.py > gpurun_out/l2_resident_probe.jsonl"""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

H, W = 720, 1440
Mo = W // 2 + 1
st_main = torch.cuda.current_stream().cuda_stream


def graph_time(fn, reps, n=5):
    fn(st_main)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s2 = torch.cuda.Stream()
    with torch.cuda.stream(s2):
        with torch.cuda.graph(g, stream=s2):
            for _ in range(reps):
                fn(s2.cuda_stream)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n / reps


for nimg in (4, 8, 12, 16, 24, 1280):
    z = torch.randn(nimg, H, Mo, 2, device="cuda")
    x = torch.randn(nimg, H, W, device="cuda")
    cols = capi.C2CPlan(capi.FORWARD, H, nimg, Mo)
    rows = capi.Plan(capi.FORWARD, (W,), nimg * H)
    irows = capi.Plan(capi.INVERSE, (W,), nimg * H)
    reps = 40 if nimg < 100 else 2
    tc = graph_time(lambda s: cols.exec_ptr(z.data_ptr(), z.data_ptr(), 0, s), reps)
    tr = graph_time(lambda s: rows.exec_ptr(x.data_ptr(), z.data_ptr(), 0, s), reps)
    ti = graph_time(lambda s: irows.exec_ptr(z.data_ptr(), x.data_ptr(), 0, s), reps)
    print(json.dumps({"images": nimg, "MB_complex": nimg * H * Mo * 8 / 1e6, "cols_us_per_image": tc * 1e3 / nimg,
                      "rows_r2c_us_per_image": tr * 1e3 / nimg, "rows_c2r_us_per_image": ti * 1e3 / nimg,
                      "cols_ms": tc, "rows_ms": tr, "plan": cols.describe()[1][-60:]}), flush=True)
