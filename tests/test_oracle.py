"""Pins the CPU oracle (oracle/) against the committed golden vectors.

The golden file holds the expected values of the reference's own tests
(/root/reference/tests/test_dft.py:124-184: CPU torch.fft, seed 1) -- see
tests/golden/make_golden.py.  Tolerances: the C restatement computes in double, so it must
agree with the fp64 golden values to ~1e-12 relative, and with the fp32 golden values under
the reference's own assertion (torch.allclose defaults, tests/test_dft.py:151,:184).
"""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import oracle as O

GOLD = np.load(Path(__file__).parent / "golden" / "reference_cases.npz")
NAMES = [str(n) for n in GOLD["names"]]
REF = [n for n in NAMES if n.startswith("ref_")]
EXT = [n for n in NAMES if n.startswith("ext_")]


def test_golden_has_the_reference_cases():
    # 2 batch sizes x 2 channel counts x 2 heights x 1 width  (tests/test_dft.py:124-127)
    assert len(REF) == 8
    for n in REF:
        assert GOLD[n + "_x"].shape[-1] == 4


@pytest.mark.parametrize("name", REF)
def test_c_oracle_matches_reference_rfft2(name):
    x = GOLD[name + "_x"]
    got = O.c_rfftn(x, 2)
    assert O.rel_l2(got, GOLD[name + "_rfft2_f64"]) < 1e-12
    # the reference's own assertion, fp32
    assert torch.allclose(torch.from_numpy(GOLD[name + "_rfft2"]), torch.from_numpy(got.astype(np.float32)))


@pytest.mark.parametrize("name", REF)
def test_c_oracle_matches_reference_irfft2(name):
    y = GOLD[name + "_rfft2"]
    got = O.c_irfftn(y, 2)
    assert O.rel_l2(got, GOLD[name + "_irfft2_f64"]) < 1e-12
    assert torch.allclose(torch.from_numpy(GOLD[name + "_irfft2"]), torch.from_numpy(got.astype(np.float32)),
                          rtol=1e-5, atol=1e-7)


def _nd(name):
    return int(name.rsplit("_nd", 1)[1])


@pytest.mark.parametrize("name", EXT)
def test_c_oracle_extended_shapes(name):
    nd = _nd(name)
    x = GOLD[name + "_x"]
    fwd = O.c_rfftn(x, nd)
    assert O.rel_l2(fwd, GOLD[name + "_rfftn_f64"]) < 1e-12
    inv = O.c_irfftn(GOLD[name + "_rfftn"], nd, last=x.shape[-1])
    assert O.rel_l2(inv, GOLD[name + "_irfftn_f64"]) < 1e-12
    # C2R semantics on a spectrum that is not Hermitian-consistent
    inv_z = O.c_irfftn(GOLD[name + "_z"], nd, last=x.shape[-1])
    assert O.rel_l2(inv_z, GOLD[name + "_z_irfftn_f64"]) < 1e-12


@pytest.mark.parametrize("name", REF + EXT)
def test_torch_oracle_matches_golden(name):
    """The torch expressions in oracle.py are the ones the golden file was made from."""
    x = torch.from_numpy(GOLD[name + "_x"])
    if name.startswith("ref_"):
        assert torch.equal(O.torch_rfftn(x, 2), torch.from_numpy(GOLD[name + "_rfft2"]))
        y = torch.from_numpy(GOLD[name + "_rfft2"])
        assert torch.equal(O.torch_irfftn(y, 2), torch.from_numpy(GOLD[name + "_irfft2"]))
    else:
        nd = _nd(name)
        got = O.torch_rfftn(x, nd, dtype=torch.float64).numpy()
        assert O.rel_l2(got, GOLD[name + "_rfftn_f64"]) < 1e-14


def test_seed_reproduces_golden_inputs():
    # tests/test_dft.py:135-136: torch.manual_seed(1); randn(B, C, H, W)
    torch.manual_seed(1)
    x = torch.randn(2, 3, 2, 4)
    assert np.array_equal(x.numpy(), GOLD["ref_b2_c3_h2_w4_x"])
