"""Raw pinned-host <-> device copy rates on this box (what bounds bench.py's e2e leg): H2D alone, D2H alone, both at once."""
import time, torch
n = 800 * 1024 * 1024 // 4
h_in = torch.randn(n).pin_memory(); h_out = torch.empty(n).pin_memory()
d_a = torch.empty(n, device="cuda"); d_b = torch.randn(n, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
def h2d():
    with torch.cuda.stream(s1): d_a.copy_(h_in, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h_out.copy_(d_b, non_blocking=True)
gb = n * 4 / 1e9
print(f"H2D alone   {gb / t(h2d):6.1f} GB/s")
print(f"D2H alone   {gb / t(d2h):6.1f} GB/s")
tb = t(lambda: (h2d(), d2h()))
print(f"both at once {gb / tb:6.1f} GB/s each direction ({2 * gb / tb:.1f} total)")
import subprocess
print(subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current", "--format=csv"], capture_output=True, text=True).stdout)
