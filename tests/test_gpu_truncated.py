"""Truncated-mode rfft2 / irfft2 (trt_dft_plan_create_truncated) against the oracle: torch.fft.rfft2 + slicing, and
torch.fft.irfft2 of the zero-padded spectrum (fp64 on the CPU; tolerance as everywhere, rel-L2 <= 1e-5).  Covers the
fused kept-modes-only kernels (90 x 180) and the crop / zero-pad route of every other shape."""
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

CASES = [
    # (batch, H, W, keep_h, keep_w)
    (7, 90, 180, 12, 20), (300, 90, 180, 45, 91), (149, 90, 180, 1, 1), (5, 90, 180, 30, 33), (64, 90, 180, 20, 64),
    (6, 128, 128, 16, 16), (3, 64, 64, 8, 12), (700, 128, 128, 12, 12), (5, 128, 128, 64, 65), (9, 128, 128, 1, 40), (4, 64, 64, 32, 33), (900, 64, 128, 12, 20), (5, 128, 64, 64, 33), (2, 720, 1440, 32, 48), (4, 34, 74, 5, 7), (3, 60, 120, 30, 61),
]


def oracle_truncated(x, kh, kw):
    from trt_dft_plugins.capi import kept_rows
    H = x.shape[-2]
    full = torch.fft.rfft2(x.double(), dim=(-2, -1), norm="backward")
    return torch.view_as_real(full[..., kept_rows(H, kh), :kw].contiguous())


def oracle_padded_inverse(kept, H, W, kh, kw):
    from trt_dft_plugins.capi import kept_rows
    full = torch.zeros(kept.shape[0], H, W // 2 + 1, dtype=torch.complex128)
    full[:, kept_rows(H, kh), :kw] = torch.view_as_complex(kept.double().contiguous())
    return torch.fft.irfft2(full, s=(H, W), dim=(-2, -1), norm="backward")


@pytest.mark.parametrize("batch,H,W,kh,kw", CASES)
def test_truncated_forward_and_inverse(batch, H, W, kh, kw):
    from trt_dft_plugins import capi
    g = torch.Generator().manual_seed(batch * 1000 + kh)
    x = torch.randn(batch, H, W, generator=g)
    f = capi.TruncatedPlan(capi.FORWARD, (H, W), batch, kh, kw)
    out = torch.full((batch, 2 * kh, kw, 2), float("nan"), device="cuda")
    xd = x.cuda()
    f.exec(xd, out)
    torch.cuda.synchronize()
    want = oracle_truncated(x, kh, kw)
    assert not torch.isnan(out).any()
    err = (out.cpu().double() - want).flatten(1).norm(dim=1) / want.flatten(1).norm(dim=1)
    assert float(err.max()) <= 1e-5, f"worst image {int(err.argmax())}: {float(err.max()):.3e}"
    assert torch.equal(xd.cpu(), x)                                      # input never written
    # inverse on a random (not Hermitian-consistent) kept block
    s = torch.randn(batch, 2 * kh, kw, 2, generator=g)
    i = capi.TruncatedPlan(capi.INVERSE, (H, W), batch, kh, kw)
    back = torch.full((batch, H, W), float("nan"), device="cuda")
    i.exec(s.cuda(), back)
    torch.cuda.synchronize()
    want_x = oracle_padded_inverse(s, H, W, kh, kw)
    err = (back.cpu().double() - want_x).flatten(1).norm(dim=1) / want_x.flatten(1).norm(dim=1)
    assert float(err.max()) <= 1e-5, f"worst image {int(err.argmax())}: {float(err.max()):.3e}"


def test_truncated_plan_describes_its_route_and_rejects_bad_sizes():
    from trt_dft_plugins import capi
    p = capi.TruncatedPlan(capi.FORWARD, (90, 180), 8, 12, 20)
    assert "kept modes only" in p.describe()[1] and p.workspace_bytes == 0 and p.num_launches == 1
    assert p.input_floats == 8 * 90 * 180 and p.output_floats == 8 * 24 * 20 * 2
    r = capi.TruncatedPlan(capi.INVERSE, (128, 128), 8, 16, 16)            # the FNO shape has fused kernels too
    assert "kept modes only" in r.describe()[1] and r.workspace_bytes == 0
    q = capi.TruncatedPlan(capi.INVERSE, (60, 120), 8, 16, 16)
    assert "zero-pad" in q.describe()[1] and q.workspace_bytes >= 8 * 60 * 61 * 8
    for kh, kw in ((0, 4), (46, 4), (4, 0), (4, 92)):
        with pytest.raises(capi.DftError):
            capi.TruncatedPlan(capi.FORWARD, (90, 180), 8, kh, kw)
    e = capi.TruncatedPlan(capi.FORWARD, (90, 180), 0, 4, 4)
    assert e.num_launches == 0


def test_full_keep_equals_the_ordinary_transform():
    from trt_dft_plugins import capi
    torch.manual_seed(3)
    x = torch.randn(200, 90, 180, device="cuda")
    full = capi.rfftn(x, 2)
    t = torch.empty(200, 90, 91, 2, device="cuda")
    capi.TruncatedPlan(capi.FORWARD, (90, 180), 200, 45, 91).exec(x, t)
    torch.cuda.synchronize()
    assert torch.equal(t, full)                 # same arithmetic, only the way out of shared memory differs
    back = torch.empty_like(x)
    capi.TruncatedPlan(capi.INVERSE, (90, 180), 200, 45, 91).exec(full, back)
    torch.cuda.synchronize()
    assert torch.equal(back, capi.irfftn(full, 2))


@pytest.mark.parametrize("H,W,kh,kw", [(90, 180, 10, 16), (64, 64, 8, 12), (60, 120, 7, 9)])
def test_truncated_plans_through_the_host_buffer_call(H, W, kh, kw):
    """trt_dft_exec_host with a truncated plan (fused route and crop / pad route): host buffers in and out."""
    from trt_dft_plugins import capi
    torch.manual_seed(5)
    batch = 9
    x = torch.randn(batch, H, W).pin_memory()
    out = torch.empty(batch, 2 * kh, kw, 2).pin_memory()
    capi.TruncatedPlan(capi.FORWARD, (H, W), batch, kh, kw).exec_host(x, out)
    want = oracle_truncated(x, kh, kw)
    assert O.rel_l2(out.numpy(), want.numpy()) <= 1e-5
    back = torch.empty(batch, H, W).pin_memory()
    capi.TruncatedPlan(capi.INVERSE, (H, W), batch, kh, kw).exec_host(out, back)
    want_x = oracle_padded_inverse(out, H, W, kh, kw)
    assert O.rel_l2(back.numpy(), want_x.numpy()) <= 1e-5


@pytest.mark.parametrize("B,C,H,W,kh,kw", [(4, 16, 128, 128, 12, 12), (2, 24, 90, 180, 16, 24), (3, 8, 64, 64, 8, 9), (2, 6, 60, 120, 10, 14)])
def test_fno_spectral_layer_through_truncated_plans(B, C, H, W, kh, kw):
    """The layer these plans exist for (FNO's SpectralConv2d, the reference's stated use, README.md:3):
    rfft2 -> keep the low modes -> per-mode channel mixing -> zero-pad -> irfft2.  Ours: truncated forward plan, the
    mixing as an einsum on the kept block, truncated inverse plan.  Checked against the same layer in fp64 torch on the
    CPU with the full transforms."""
    from trt_dft_plugins import capi
    g = torch.Generator().manual_seed(B * 100 + C)
    x = torch.randn(B, C, H, W, generator=g)
    w = torch.randn(C, C, 2 * kh, kw, 2, generator=g) / C
    rows = capi.kept_rows(H, kh)
    # fp64 layer with full transforms
    wc = torch.view_as_complex(w.double().contiguous())
    X = torch.fft.rfft2(x.double())
    Y = torch.zeros(B, C, H, W // 2 + 1, dtype=torch.complex128)
    Y[:, :, rows, :kw] = torch.einsum("bixy,ioxy->boxy", X[:, :, rows, :kw], wc)
    want = torch.fft.irfft2(Y, s=(H, W))
    # the same layer on the GPU through the truncated plans
    f = capi.TruncatedPlan(capi.FORWARD, (H, W), B * C, kh, kw)
    i = capi.TruncatedPlan(capi.INVERSE, (H, W), B * C, kh, kw)
    kept = torch.empty(B, C, 2 * kh, kw, 2, device="cuda")
    f.exec(x.cuda(), kept)
    mixed = torch.einsum("bixy,ioxy->boxy", torch.view_as_complex(kept), torch.view_as_complex(w.cuda().contiguous()))
    out = torch.empty(B, C, H, W, device="cuda")
    i.exec(torch.view_as_real(mixed).contiguous(), out)
    torch.cuda.synchronize()
    err = float((out.cpu().double() - want).norm() / want.norm())
    assert err <= 1e-5, err
