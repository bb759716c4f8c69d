"""Randomised sweep of the plan kinds beyond plain rfftN / irfftN -- truncated modes, 16-bit tensors, strided-axis C2C
(out of place and in place) -- against fp64 numpy FFTs on the CPU.  One line per failing case, exit code 1 if any.
(numpy, not torch.fft, for the reason given in tools/fuzz_shapes.py.)
    python tools/fuzz_plans.py [cases] [seed] [v]"""
import random
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
import numpy as np
import torch
from trt_dft_plugins import capi

ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
verbose = len(sys.argv) > 3
rng = random.Random(seed)

LENS = [2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 15, 16, 17, 18, 20, 22, 24, 26, 30, 31, 32, 34, 36, 37, 40, 45, 48, 60, 64, 74, 90, 96, 100, 120, 128, 180, 200,
        256, 360, 512, 720, 1024, 1440]
FUSED = [(90, 180), (128, 128), (64, 64)]


def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def truncated(case):
    H, W = rng.choice(FUSED) if rng.random() < 0.6 else (rng.choice([v for v in LENS if v % 2 == 0 and v <= 360]), rng.choice([v for v in LENS if v <= 720]))
    batch = rng.choice([1, 2, 3, 7, 33, 149, 300, 445])
    batch = max(1, min(batch, 3_000_000 // (H * W)))
    kh, kw = rng.randint(1, H // 2), rng.randint(1, W // 2 + 1)
    tag = f"truncated batch {batch} {H}x{W} keep {kh}x{kw}"
    g = torch.Generator().manual_seed(case)
    x = torch.randn(batch, H, W, generator=g)
    rows = capi.kept_rows(H, kh)
    want = np.fft.rfft2(x.double().numpy())[:, rows, :kw]
    out = torch.full((batch, 2 * kh, kw, 2), float("nan"), device="cuda")
    capi.TruncatedPlan(capi.FORWARD, (H, W), batch, kh, kw).exec(x.cuda(), out)
    torch.cuda.synchronize()
    got = torch.view_as_complex(out.cpu().double().contiguous()).numpy()
    ef = rel(got, want)
    s = torch.randn(batch, 2 * kh, kw, 2, generator=g)
    full = np.zeros((batch, H, W // 2 + 1), dtype=np.complex128)
    full[:, rows, :kw] = torch.view_as_complex(s.double().contiguous()).numpy()
    want_x = np.fft.irfft2(full, s=(H, W))
    back = torch.full((batch, H, W), float("nan"), device="cuda")
    capi.TruncatedPlan(capi.INVERSE, (H, W), batch, kh, kw).exec(s.cuda(), back)
    torch.cuda.synchronize()
    ei = rel(back.cpu().double().numpy(), want_x)
    return tag, ef, ei, 1e-5


def typed(case):
    dt = rng.choice([torch.float16, torch.bfloat16])
    nd = rng.choice([1, 2, 2, 2, 3])
    if nd == 2 and rng.random() < 0.5:
        dims = (90, 180)
    else:
        dims = tuple(rng.choice([v for v in LENS if v <= (1440 if nd < 3 else 40)]) for _ in range(nd))
    n = int(np.prod(dims))
    batch = max(1, min(rng.choice([1, 2, 3, 8, 149, 300, 445, 1001]), 3_000_000 // n))
    tag = f"typed {str(dt).split('.')[1]} batch {batch} dims {dims}"
    g = torch.Generator().manual_seed(case)
    x = torch.randn(batch, *dims, generator=g).to(dt)
    axes = tuple(range(-nd, 0))
    want = np.fft.rfftn(x.double().numpy(), axes=axes)
    out = torch.full((batch,) + dims[:-1] + (dims[-1] // 2 + 1, 2), float("nan"), dtype=dt, device="cuda")
    capi.TypedPlan(capi.FORWARD, dims, batch, dt).exec(x.cuda(), out)
    torch.cuda.synchronize()
    ef = rel(torch.view_as_complex(out.cpu().double().contiguous()).numpy(), want)
    s = torch.randn(*out.shape, generator=g).to(dt)
    want_x = np.fft.irfftn(torch.view_as_complex(s.double().contiguous()).numpy(), s=dims, axes=axes)
    back = torch.full((batch,) + dims, float("nan"), dtype=dt, device="cuda")
    capi.TypedPlan(capi.INVERSE, dims, batch, dt).exec(s.cuda(), back)
    torch.cuda.synchronize()
    ei = rel(back.cpu().double().numpy(), want_x)
    return tag, ef, ei, (3e-4 if dt == torch.float16 else 2.5e-3)


def c2c(case):
    n = rng.choice(LENS + [2048, 4096, 8192, 41, 53, 97, 127, 8209])
    outer = rng.choice([1, 2, 3, 5])
    inner = rng.choice([1, 2, 3, 7, 16, 33, 64, 100, 257, 1000])
    while outer * n * inner > 3_000_000 and inner > 1:
        inner = max(1, inner // 2)
    inplace = rng.random() < 0.5
    tag = f"c2c n {n} outer {outer} inner {inner} {'in place' if inplace else 'out of place'}"
    g = torch.Generator().manual_seed(case)
    z = torch.randn(outer, n, inner, 2, generator=g)
    zc = torch.view_as_complex(z.double().contiguous()).numpy()
    res = []
    for direction, ref in ((capi.FORWARD, np.fft.fft(zc, axis=1)), (capi.INVERSE, np.fft.ifft(zc, axis=1) * n)):
        p = capi.C2CPlan(direction, n, outer, inner)
        src = z.cuda()
        dst = src if inplace else torch.full_like(src, float("nan"))
        p.exec(src, dst)
        torch.cuda.synchronize()
        res.append(rel(torch.view_as_complex(dst.cpu().double().contiguous()).numpy(), ref))
    return tag, res[0], res[1], 1e-5


fails = 0
for case in range(ncases):
    kind = rng.choice([truncated, typed, c2c])
    try:
        if verbose:
            print(f"case {case} {kind.__name__}", flush=True)
        tag, ef, ei, tol = kind(case)
        bad = not (ef <= tol and ei <= tol)
    except Exception as e:
        tag, ef, ei, bad = f"{kind.__name__} EXC {e!r}"[:240], float("nan"), float("nan"), True
    if bad:
        fails += 1
        print(f"FAIL case {case} {tag}: forward {ef:.2e} inverse {ei:.2e}", flush=True)
print(f"{ncases} cases, {fails} failed (seed {seed})")
sys.exit(1 if fails else 0)
