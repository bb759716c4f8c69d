"""Generates tests/golden/reference_cases.npz -- run once in the build container.

The reference's hot path (cuFFT inside a TensorRT plugin) cannot run here (no TensorRT,
no GPU), and /root/reference/tests/test_dft.py cannot be imported (``import tensorrt``).
What *can* be reproduced is the expected-value side of the reference's own tests: the
oracle expressions at tests/test_dft.py:40 and :52-54 evaluated on CPU at the 16
parametrised cases of tests/test_dft.py:124-184 (seed ``torch.manual_seed(1)``,
``x = randn(batch, num_c, dft_dim1, dft_dim2)``).  The two expressions below are those
lines, nothing else; they are evaluated in fp32 (what the reference asserts against) and in
fp64 (what north_star's rel-L2 tolerance is stated against).

A few larger shapes (cfg1 [4,32,64] nd=1, mixed radix, odd lengths, nd=3) are added with
the same seed so the C restatement is pinned beyond 1|2 x 4.

    python tests/golden/make_golden.py
"""
from pathlib import Path

import numpy as np
import torch


def ref_rfft2(x):  # tests/test_dft.py:40
    return torch.view_as_real(torch.fft.rfft2(x, dim=(-2, -1), norm="backward"))


def ref_irfft2(y):  # tests/test_dft.py:52-54
    return torch.fft.irfft2(torch.view_as_complex(y), dim=(-2, -1), norm="backward")


def main():
    out = {}
    names = []
    # the reference's 8 + 8 parametrised cases
    for batch_size in (1, 2):
        for num_c in (1, 3):
            for dft_dim2 in (4,):
                for dft_dim1 in (1, 2):
                    torch.manual_seed(1)
                    x = torch.randn(batch_size, num_c, dft_dim1, dft_dim2)
                    y = ref_rfft2(x)
                    xr = ref_irfft2(y)
                    key = f"ref_b{batch_size}_c{num_c}_h{dft_dim1}_w{dft_dim2}"
                    names.append(key)
                    out[key + "_x"] = x.numpy()
                    out[key + "_rfft2"] = y.numpy()
                    out[key + "_irfft2"] = xr.numpy()
                    out[key + "_rfft2_f64"] = ref_rfft2(x.double()).numpy()
                    out[key + "_irfft2_f64"] = ref_irfft2(y.double()).numpy()
    # extra shapes: (shape, signal_ndim)
    extra = [((4, 32, 64), 1), ((3, 30), 1), ((2, 9), 1), ((2, 3, 6, 10), 2), ((2, 5, 7), 2),
             ((2, 12, 18), 2), ((1, 2, 4, 6, 8), 3), ((2, 3, 5, 6), 3), ((1, 90, 180), 2)]
    for shape, nd in extra:
        torch.manual_seed(1)
        x = torch.randn(*shape)
        dims = tuple(range(-nd, 0))
        y64 = torch.view_as_real(torch.fft.rfftn(x.double(), dim=dims, norm="backward"))
        y32 = torch.view_as_real(torch.fft.rfftn(x, dim=dims, norm="backward"))
        key = "ext_" + "x".join(map(str, shape)) + f"_nd{nd}"
        names.append(key)
        out[key + "_x"] = x.numpy()
        out[key + "_rfftn"] = y32.numpy()
        out[key + "_rfftn_f64"] = y64.numpy()
        # inverse of the fp32 spectrum; s= pins the logical real size (odd sizes included)
        s = shape[-nd:]
        xi64 = torch.fft.irfftn(torch.view_as_complex(y32.double()), s=s, dim=dims, norm="backward")
        out[key + "_irfftn_f64"] = xi64.numpy()
        # non-Hermitian-consistent spectrum: C2R semantics (imag of DC / Nyquist ignored)
        torch.manual_seed(2)
        z = torch.randn(*y32.shape)
        out[key + "_z"] = z.numpy()
        out[key + "_z_irfftn_f64"] = torch.fft.irfftn(
            torch.view_as_complex(z.double()), s=s, dim=dims, norm="backward").numpy()
    out["names"] = np.array(names)
    path = Path(__file__).resolve().parent / "reference_cases.npz"
    np.savez_compressed(path, **out)
    print("wrote", path, path.stat().st_size, "bytes,", len(names), "cases")


if __name__ == "__main__":
    main()
