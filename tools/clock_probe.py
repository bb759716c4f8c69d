"""Times the cfg3 kernels (a) back to back for ~1 s with nvidia-smi sampling and (b) one launch at a
time with idle gaps, to separate kernel quality from power-cap clock behaviour."""
import subprocess, sys, time, threading
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import torch
from trt_dft_plugins import capi
shape = (8, 768, 90, 180)
x = torch.randn(*shape, device="cuda"); y = torch.empty(8, 768, 90, 91, 2, device="cuda"); xr = torch.empty_like(x)
f = capi.Plan(capi.FORWARD, (90, 180), 6144); i = capi.Plan(capi.INVERSE, (90, 180), 6144)
st = torch.cuda.current_stream().cuda_stream
def pair():
    f.exec_ptr(x.data_ptr(), y.data_ptr(), 0, st); i.exec_ptr(y.data_ptr(), xr.data_ptr(), 0, st)
for _ in range(5): pair()
torch.cuda.synchronize()
q = "clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.hw_thermal_slowdown"
p = subprocess.Popen(["nvidia-smi", "-i", "0", f"--query-gpu={q}", "--format=csv,noheader", "-lms", "50"], stdout=subprocess.PIPE, text=True)
lines = []
t = threading.Thread(target=lambda: [lines.append((time.time(), l.strip())) for l in p.stdout], daemon=True); t.start()
time.sleep(0.3)
for n in (50, 200, 1000, 3000):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time(); e0.record()
    for _ in range(n): pair()
    e1.record(); torch.cuda.synchronize(); t1 = time.time()
    ms = e0.elapsed_time(e1) / n
    during = [l for (ts, l) in lines if t0 <= ts <= t1]
    print(f"back-to-back n={n}: {ms:.4f} ms/pair  ({1.601372160/ms*1e3/1e3:.0f} GB/s)  samples: {during[len(during)//2] if during else None}")
    time.sleep(0.5)
# isolated launches
for name, fn in (("fwd", lambda: f.exec_ptr(x.data_ptr(), y.data_ptr(), 0, st)), ("inv", lambda: i.exec_ptr(y.data_ptr(), xr.data_ptr(), 0, st))):
    ts = []
    for _ in range(20):
        time.sleep(0.02)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort(); print(f"isolated {name}: median {ts[10]:.4f} ms, min {ts[0]:.4f}")
# single direction back to back
for name, fn in (("fwd", lambda: f.exec_ptr(x.data_ptr(), y.data_ptr(), 0, st)), ("inv", lambda: i.exec_ptr(y.data_ptr(), xr.data_ptr(), 0, st))):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(500): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"back-to-back {name} x500: {e0.elapsed_time(e1)/500:.4f} ms")
p.terminate()
print("all samples (every 4th):"); [print(l) for (ts, l) in lines[::4]]
