"""Randomised shape sweep of rfftN / irfftN through the C ABI against an fp64 FFT on the CPU.  Prints one line per failing
case and a summary; exit code 1 if anything failed.
The CPU side is numpy's pocketfft, not torch.fft as in the parity tests: after several hundred DISTINCT shapes in one process
torch's CPU FFT (MKL descriptor cache) aborts with "corrupted size vs. prev_size" -- reproducible with torch alone, none of
this library loaded (seed 2, case 745).
    python tools/fuzz_shapes.py [cases] [seed] [v]"""
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
rng = random.Random(seed)
verbose = len(sys.argv) > 3          # any third argument: print every case before it runs (to find a crashing one)

SMOOTH = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 21, 22, 24, 25, 26, 27, 28, 30, 32, 33, 35, 36, 39, 40, 42, 44, 45, 48, 49, 50,
          52, 54, 55, 56, 60, 63, 64, 65, 66, 70, 72, 75, 77, 78, 80, 81, 84, 88, 90, 91, 96, 98, 99, 100, 104, 105, 108, 110, 112, 117, 120,
          121, 125, 126, 128, 130, 132, 135, 140, 143, 144, 150, 160, 162, 165, 168, 169, 175, 176, 180, 182, 192, 195, 196, 198, 200, 208, 210,
          216, 220, 224, 225, 231, 234, 240, 243, 245, 250, 252, 256, 260, 264, 270, 273, 275, 280, 286, 288, 294, 297, 300, 308, 312, 315, 320,
          324, 325, 330, 336, 338, 343, 350, 351, 352, 360, 363, 364, 375, 378, 384, 385, 390, 392, 396, 400, 405, 416, 420, 429, 432, 440, 441,
          448, 450, 455, 462, 468, 480, 486, 490, 495, 500, 504, 507, 512, 520, 525, 528, 539, 540, 546, 550, 560, 567, 572, 576, 585, 588, 594,
          600, 605, 616, 624, 625, 630, 637, 640, 648, 650, 660, 672, 675, 676, 686, 693, 700, 702, 704, 715, 720, 726, 728, 729, 735, 750, 756,
          768, 770, 780, 784, 792, 800, 810, 819, 825, 832, 840, 845, 847, 858, 864, 875, 880, 882, 891, 896, 900, 910, 924, 936, 945, 960, 968,
          972, 975, 980, 990, 1000, 1001, 1008, 1014, 1024, 1280, 1440, 1536, 1800, 2048, 2187, 2400, 2560, 3072, 3125, 3600, 4096, 4100, 5000, 6144, 8192]
PRIMES = [17, 19, 23, 29, 31, 37, 41, 43, 47, 53, 59, 61, 67, 71, 73, 79, 83, 89, 97, 101, 127, 131, 251, 257, 509, 521, 1021, 1031, 2053, 4099, 8191, 8209]


def pick_len(limit):
    r = rng.random()
    if r < 0.6:
        c = [v for v in SMOOTH if v <= limit]
        return rng.choice(c)
    if r < 0.8:
        c = [v for v in PRIMES if v <= limit] or [2]
        return rng.choice(c)
    if r < 0.9:
        p = rng.choice([17, 19, 23, 29, 31, 37, 41, 53, 97])
        m = rng.choice([1, 2, 3, 4, 5, 6, 8, 10])
        return min(p * m, max(limit, 2))
    return rng.randint(1, min(limit, 700))


fails = 0
done = 0
for case in range(ncases):
    nd = rng.choice([1, 1, 2, 2, 2, 3])
    if nd == 1:
        dims = (pick_len(8209),)
    elif nd == 2:
        a = pick_len(1500)
        dims = (a, pick_len(max(2, min(8192, 600000 // a))))
    else:
        a = pick_len(64)
        b = pick_len(max(2, 4000 // a))
        dims = (a, b, pick_len(max(2, min(2048, 120000 // (a * b)))))
    n = 1
    for d in dims:
        n *= d
    batch = rng.choice([1, 1, 2, 3, 5, 8, 17, 64, 150, 301])
    batch = max(1, min(batch, 2_000_000 // n))
    g = torch.Generator().manual_seed(case)
    x = torch.randn(batch, *dims, generator=g)
    dim = tuple(range(-nd, 0))
    tag = f"case {case}: batch {batch} dims {dims}"
    if verbose:
        print(tag, flush=True)
    try:
        want = torch.view_as_real(torch.from_numpy(np.fft.rfftn(x.double().numpy(), axes=dim)))
        y = capi.rfftn(x.cuda(), nd).cpu()
        ef = float((y.double() - want).norm() / want.norm().clamp_min(1e-30))
        s = torch.randn(*want.shape, generator=g)
        want_x = torch.from_numpy(np.fft.irfftn(torch.view_as_complex(s.double().contiguous()).numpy(), s=dims, axes=dim))
        xb = capi.irfftn(s.cuda(), nd, last=dims[-1]).cpu()
        ei = float((xb.double() - want_x).norm() / want_x.norm().clamp_min(1e-30))
        bad = not (ef <= 1e-5 and ei <= 1e-5) or bool(torch.isnan(y).any()) or bool(torch.isnan(xb).any())
    except Exception as e:  # a refused shape is a finding too
        bad, ef, ei = True, float("nan"), float("nan")
        tag += f" EXC {e!r}"[:200]
    done += 1
    if bad:
        fails += 1
        print(f"FAIL {tag}: forward {ef:.2e} inverse {ei:.2e}", flush=True)
print(f"{done} cases, {fails} failed (seed {seed})")
sys.exit(1 if fails else 0)
