"""CPU check of the FFT engine's planner + tile programs through the test-only host emulator.

tests/emu/dft_emu.cpp compiles the same dft_planner.h / dft_programs.cuh the CUDA kernels use,
with one host "thread" per tile, and this file compares it with the oracle (fp64 torch.fft, i.e.
the reference tests' oracle expression, tests/test_dft.py:40,:52-54).  It validates index
arithmetic, permutations and the real-FFT split formulas without a GPU; it does not replace
the -m gpu parity tests, which run the real kernels.
Tolerances (north_star): rel-L2 <= 1e-5 vs fp64; round trip <= 1e-6.
"""
import ctypes
import subprocess
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import oracle as O

EMU_DIR = Path(__file__).parent / "emu"


@pytest.fixture(scope="module")
def emu():
    subprocess.check_call(["make", "-C", str(EMU_DIR), "-s"])
    lib = ctypes.CDLL(str(EMU_DIR / "libdft_emu.so"))
    lib.dft_emu_exec.restype = ctypes.c_int
    lib.dft_emu_exec.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int64,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t]
    return lib


def run(lib, direction, x, nd, last=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    if direction == 0:
        dims = x.shape[-nd:]
        batch = int(np.prod(x.shape[:-nd], dtype=np.int64))
        out = np.full(x.shape[:-1] + (dims[-1] // 2 + 1, 2), np.nan, dtype=np.float32)
    else:
        m = x.shape[-2]
        last = (m - 1) * 2 if last is None else last
        dims = x.shape[-nd - 1:-2] + (last,)
        batch = int(np.prod(x.shape[:-nd - 1], dtype=np.int64))
        out = np.full(x.shape[:-2] + (last,), np.nan, dtype=np.float32)
    d = (ctypes.c_int64 * nd)(*dims)
    desc = ctypes.create_string_buffer(512)
    rc = lib.dft_emu_exec(direction, nd, d, batch, x.ctypes.data, out.ctypes.data, desc, 512)
    assert rc == 0
    return out, desc.value.decode()


SHAPES = [
    # the reference tests' shapes (tests/test_dft.py:124-127)
    ((1, 1, 1, 4), 2), ((2, 3, 2, 4), 2), ((2, 1, 1, 4), 2), ((1, 3, 2, 4), 2),
    # BASELINE configs (cfg1 in full; cfg2/cfg3 at a reduced batch)
    ((4, 32, 64), 1), ((2, 3, 128, 128), 2), ((5, 90, 180), 2),
    # 1-D: even, odd, mixed radix, primes in and out of the radix set
    ((3, 2), 1), ((3, 6), 1), ((5, 30), 1), ((2, 9), 1), ((2, 15), 1), ((3, 1), 1), ((2, 360), 1),
    ((2, 1024), 1), ((2, 64), 1), ((3, 128), 1), ((1, 16384), 1), ((2, 1440), 1), ((2, 2 * 7 * 11 * 13), 1), ((2, 34), 1), ((3, 19), 1), ((1, 4096), 1),
    # 2-D: fused, multi-image tiles, odd widths, unsupported primes (naive), H == 1
    ((7, 8, 8), 2), ((3, 6, 10), 2), ((2, 5, 7), 2), ((2, 12, 18), 2), ((2, 17, 6), 2), ((2, 6, 38), 2),
    ((700, 4, 4), 2), ((1, 4096, 64), 2), ((2, 5120, 40), 2), ((1, 1, 16), 2), ((2, 16, 1), 2), ((2, 3, 3), 2),
    # prime factors 17..31 run as register butterflies, 37 and up as generic prime-radix steps (one output per work item)
    ((2, 2 * 17 * 19), 1), ((3, 46), 1), ((2, 29), 1), ((2, 62), 1), ((2, 34, 58), 2), ((2, 31, 38), 2), ((1, 74), 1), ((2, 37, 6), 2), ((2, 2 * 37 * 41), 1), ((3, 1009), 1), ((2, 202), 1), ((2, 43, 86), 2), ((1, 3, 74, 6), 3), ((2, 8 * 53), 1),
    # 3-D
    ((2, 4, 6, 8), 3), ((2, 3, 5, 6), 3), ((1, 2, 3, 5, 7), 3), ((2, 1, 4, 4), 3), ((1, 17, 4, 6), 3),
]


@pytest.mark.parametrize("shape,nd", SHAPES)
def test_emulated_engine_matches_oracle(emu, shape, nd):
    torch.manual_seed(1)
    x = torch.randn(*shape)
    want = O.torch_rfftn(x, nd, dtype=torch.float64).numpy()
    got, desc = run(emu, 0, x.numpy(), nd)
    assert not np.isnan(got).any(), desc
    assert O.rel_l2(got, want) <= 1e-5, desc
    assert O.rel_l2(got, want) <= 2e-6, desc      # expected fp32 quality

    # inverse of the fp32 spectrum (Hermitian-consistent input, as tests/test_dft.py:169)
    y32 = want.astype(np.float32)
    s = shape[-nd:]
    want_i = torch.fft.irfftn(torch.view_as_complex(torch.from_numpy(y32).double()), s=s,
                              dim=tuple(range(-nd, 0)), norm="backward").numpy()
    got_i, desc_i = run(emu, 1, y32, nd, last=shape[-1])
    assert not np.isnan(got_i).any(), desc_i
    assert O.rel_l2(got_i, want_i) <= 1e-5, desc_i
    assert O.rel_l2(got_i, want_i) <= 2e-6, desc_i

    # round trip through both engine directions
    rt, _ = run(emu, 1, got, nd, last=shape[-1])
    assert O.rel_l2(rt, x.numpy()) <= 1e-6

    # C2R semantics on a spectrum that is not Hermitian-consistent
    torch.manual_seed(2)
    z = torch.randn(*y32.shape)
    want_z = torch.fft.irfftn(torch.view_as_complex(z.double()), s=s, dim=tuple(range(-nd, 0)),
                              norm="backward").numpy()
    got_z, _ = run(emu, 1, z.numpy(), nd, last=shape[-1])
    assert O.rel_l2(got_z, want_z) <= 1e-5


def test_large_two_pass_plan(emu):
    """An image too large to fuse takes the rows + strided-columns path (with workspace for the inverse)."""
    torch.manual_seed(1)
    x = torch.randn(1, 180, 360)   # 180 x 181 complex = 260 KB > one CTA's shared memory
    got, desc = run(emu, 0, x.numpy(), 2)
    assert "rows_r2c" in desc and "cols_c2c" in desc
    want = O.torch_rfftn(x, 2, dtype=torch.float64).numpy()
    assert O.rel_l2(got, want) <= 2e-6
    inv, desc_i = run(emu, 1, got, 2)
    assert "cols_c2c" in desc_i and "rows_c2r" in desc_i
    assert O.rel_l2(inv, x.numpy()) <= 1e-6
