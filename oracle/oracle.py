"""CPU oracle for the Rfft / Irfft plugin path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product (tensorrt-dft-plugins_b200/) never does.

Two oracles live here:

* ``torch_rfftn`` / ``torch_irfftn`` -- the *reference test-suite's own oracle*: the
  exact expressions of /root/reference/tests/test_dft.py:40 (OnnxRfft2.forward) and
  :52-54 (OnnxIrfft2.forward), ``torch.fft.rfft2 / irfft2(..., norm="backward")`` on
  CPU, generalised from ``dim=(-2,-1)`` to the trailing ``signal_ndim`` dims
  (signal_ndim in {1,2,3}, /root/reference/src/dft_plugins/dft_plugins.cpp:57).
  This is also what bench.py times as the CPU baseline (north_star: "the reference's
  CPU oracle path (torch.fft on the GPU box's own host cores)").

* ``c_rfftn`` / ``c_irfftn`` -- the plain-C restatement in dft_oracle.c (naive
  double-precision DFT; independent of MKL/pocketfft and of the CUDA kernels).

Parity pinning: tests/test_oracle.py checks both against tests/golden/*.npz, which
were produced by tests/golden/make_golden.py from the reference tests' expressions at
the reference tests' 16 shapes and seed.
"""
import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np
import torch

_HERE = Path(__file__).resolve().parent
_LIB = None


def _dims(nd):
    return tuple(range(-nd, 0))


def torch_rfftn(x: torch.Tensor, signal_ndim: int, dtype=None) -> torch.Tensor:
    """tests/test_dft.py:40 generalised: view_as_real(rfftn(x, trailing dims, "backward"))."""
    if dtype is not None:
        x = x.to(dtype)
    return torch.view_as_real(torch.fft.rfftn(x, dim=_dims(signal_ndim), norm="backward"))


def torch_irfftn(y: torch.Tensor, signal_ndim: int, dtype=None) -> torch.Tensor:
    """tests/test_dft.py:52-54 generalised.  Output last dim = (m-1)*2, the plugin's rule
    (dft_plugins.cpp:415-436), which is also torch's default."""
    if dtype is not None:
        y = y.to(dtype)
    return torch.fft.irfftn(torch.view_as_complex(y.contiguous()), dim=_dims(signal_ndim), norm="backward")


def build_c_oracle(force: bool = False) -> Path:
    so = _HERE / "libdft_oracle.so"
    src = _HERE / "dft_oracle.c"
    if force or not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["make", "-C", str(_HERE), "-s"])
    return so


def _lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(str(build_c_oracle()))
        for name in ("oracle_rfftn_f64", "oracle_irfftn_f64"):
            fn = getattr(_LIB, name)
            fn.restype = ctypes.c_int
            fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                           ctypes.POINTER(ctypes.c_int64), ctypes.c_int64]
    return _LIB


def _split(shape, nd):
    shape = tuple(int(s) for s in shape)
    batch = int(np.prod(shape[:-nd], dtype=np.int64)) if len(shape) > nd else 1
    return batch, shape[-nd:]


def c_rfftn(x: np.ndarray, signal_ndim: int) -> np.ndarray:
    """float32 real [..., d1..dk] -> float64 [..., d1.., dk//2+1, 2] (unnormalised)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    batch, dims = _split(x.shape, signal_ndim)
    out_shape = x.shape[:-1] + (dims[-1] // 2 + 1, 2)
    out = np.empty(out_shape, dtype=np.float64)
    d = (ctypes.c_int64 * signal_ndim)(*dims)
    rc = _lib().oracle_rfftn_f64(x.ctypes.data, out.ctypes.data, signal_ndim, d, batch)
    assert rc == 0
    return out


def c_irfftn(y: np.ndarray, signal_ndim: int, last: int = None) -> np.ndarray:
    """float32 [..., d1.., m, 2] -> float64 real [..., d1.., last]; last defaults to (m-1)*2."""
    y = np.ascontiguousarray(y, dtype=np.float32)
    assert y.shape[-1] == 2
    m = y.shape[-2]
    last = (m - 1) * 2 if last is None else int(last)
    assert last // 2 + 1 == m
    cshape = y.shape[:-2] + (last,)
    batch, dims = _split(cshape, signal_ndim)
    out = np.empty(cshape, dtype=np.float64)
    d = (ctypes.c_int64 * signal_ndim)(*dims)
    rc = _lib().oracle_irfftn_f64(y.ctypes.data, out.ctypes.data, signal_ndim, d, batch)
    assert rc == 0
    return out


def rel_l2(a, b) -> float:
    """||a-b||_2 / ||b||_2 in float64 (b is the reference)."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / (den if den > 0 else 1.0))
