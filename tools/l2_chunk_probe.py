"""Does cutting cfg4's two passes into batch chunks keep the intermediate in L2?  Times rows+columns per chunk
(stream-ordered launches of small plans over slices of one big tensor) for several chunk sizes.
    python tools/l2_chunk_probe.py > gpurun_out/l2_chunk_probe.jsonl"""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi

H, W, N = 720, 1440, 1280
Mo = W // 2 + 1
torch.manual_seed(1)
x = torch.randn(N, H, W, device="cuda")
y = torch.empty(N, H, Mo, 2, device="cuda")
xr = torch.empty_like(x)
st = torch.cuda.current_stream().cuda_stream
algo = (x.numel() + y.numel()) * 4


def timeit(fn, n=5):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for chunk in (1280, 64, 32, 16, 12, 8, 6, 4, 2):
    f = capi.Plan(capi.FORWARD, (H, W), chunk)
    i = capi.Plan(capi.INVERSE, (H, W), chunk)
    ws = torch.empty(max(i.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
    xs, ys, rs = H * W * 4 * chunk, H * Mo * 8 * chunk, H * W * 4 * chunk

    def fwd():
        for c in range(N // chunk):
            f.exec_ptr(x.data_ptr() + c * xs, y.data_ptr() + c * ys, 0, st)

    def inv():
        for c in range(N // chunk):
            i.exec_ptr(y.data_ptr() + c * ys, xr.data_ptr() + c * rs, ws.data_ptr(), st)

    fwd(); inv()
    torch.cuda.synchronize()
    # replay from CUDA graphs so that the host's launch rate does not cap the small-chunk runs
    gf, gi = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    s2 = torch.cuda.Stream()
    with torch.cuda.stream(s2):
        st2 = s2.cuda_stream
        with torch.cuda.graph(gf, stream=s2):
            for c in range(N // chunk):
                f.exec_ptr(x.data_ptr() + c * xs, y.data_ptr() + c * ys, 0, st2)
        with torch.cuda.graph(gi, stream=s2):
            for c in range(N // chunk):
                i.exec_ptr(y.data_ptr() + c * ys, xr.data_ptr() + c * rs, ws.data_ptr(), st2)
    tf, ti = timeit(gf.replay), timeit(gi.replay)
    rt = ((xr - x).double().norm() / x.double().norm()).item()
    print(json.dumps({"chunk_images": chunk, "chunk_MB": ys / 1e6, "fwd_ms": tf, "inv_ms": ti, "fwd_GBs": algo / tf / 1e6,
                      "inv_GBs": algo / ti / 1e6, "round_trip": rt, "plan_fwd": f.describe()[1]}), flush=True)
