"""fp16 / bf16 tensors at the plan's boundary (trt_dft_plan_create_typed), fp32 arithmetic inside.

Two checks per case.  (1) Against the oracle: fp64 torch.fft of the very same 16-bit input; what is left is the rounding of
the output to 16 bits (half an ulp per element: rel-L2 about 2e-4 for fp16, 1.6e-3 for bf16), so the tolerances are 3e-4
and 2.5e-3 -- written here because north_star's 1e-5 is a statement about fp32 tensors.  (2) Against this library's own
fp32 path on the widened input, rounded to 16 bits by torch: the typed plans run the same fp32 arithmetic, so the two must
agree to the last bit almost everywhere (a different FMA contraction may flip a rounding once in a while: <= 1 ulp, <= 0.01 %
of the elements).  That second check is what catches a misplaced element; it is exact where (1) is statistical.

90 x 180, 128 x 128 and 64 x 64 run kernels that convert as they load and store (odd and even image counts, more images than
the grid holds at once, a single image); the other shapes run the fp32 plan between two conversion passes."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = {torch.float16: 3e-4, torch.bfloat16: 2.5e-3}

CASES = [
    # (leading shape, signal dims)
    ((1,), (90, 180)), ((2,), (90, 180)), ((149,), (90, 180)), ((445,), (90, 180)), ((2, 500), (90, 180)), ((1333,), (90, 180)),
    ((3,), (128, 128)), ((1024,), (128, 128)), ((445,), (64, 64)), ((2000,), (64, 64)), ((1001,), (64, 128)), ((9,), (128, 64)),
    ((4, 32), (64,)), ((5,), (34, 74)), ((2,), (8, 12, 10)), ((3,), (720, 1440)), ((7,), (31,)), ((6,), (60, 120)),
]


def ulps_apart(a, b):
    """Distance in units of the last place between two 16-bit tensors of one dtype (same-sign values; zero for equal)."""
    ia = a.view(torch.int16).to(torch.int32)
    ib = b.view(torch.int16).to(torch.int32)
    # sign-magnitude -> monotone integers
    ia = torch.where(ia < 0, -(ia & 0x7FFF), ia)
    ib = torch.where(ib < 0, -(ib & 0x7FFF), ib)
    return (ia - ib).abs()


def close_to_own_fp32_path(got16, want32):
    want16 = want32.to(got16.dtype)
    d = ulps_apart(got16, want16)
    assert int(d.max()) <= 1, f"max distance {int(d.max())} ulp"
    frac = float((d > 0).float().mean())
    assert frac <= 1e-4, f"{frac:.2e} of the elements differ from the rounded fp32 result"


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("lead,dims", CASES)
def test_typed_forward_and_inverse(lead, dims, dtype):
    from trt_dft_plugins import capi
    nd = len(dims)
    batch = 1
    for v in lead:
        batch *= v
    g = torch.Generator().manual_seed(batch + 7 * nd)
    x16 = torch.randn(*lead, *dims, generator=g).to(dtype)
    cshape = tuple(lead) + tuple(dims[:-1]) + (dims[-1] // 2 + 1, 2)
    dim = tuple(range(-nd, 0))

    f = capi.TypedPlan(capi.FORWARD, dims, batch, dtype)
    assert f.io_type == {torch.float16: capi.IO_F16, torch.bfloat16: capi.IO_BF16}[dtype]
    out = torch.full(cshape, float("nan"), dtype=dtype, device="cuda")
    xd = x16.cuda()
    f.exec(xd, out)
    torch.cuda.synchronize()
    assert not torch.isnan(out).any()
    assert torch.equal(xd.cpu(), x16)                                      # input never written
    want = torch.view_as_real(torch.fft.rfftn(x16.double(), dim=dim, norm="backward"))
    flat = lambda t: t.reshape(batch, -1)
    err = flat(out.cpu().double() - want).norm(dim=1) / flat(want).norm(dim=1)
    assert float(err.max()) <= TOL[dtype], f"worst image {int(err.argmax())}: {float(err.max()):.3e}"
    close_to_own_fp32_path(out.cpu(), capi.rfftn(xd.float(), nd).cpu())

    # inverse of a random (not Hermitian-consistent) spectrum, as the reference's tests do (tests/test_dft.py:52-54)
    s16 = torch.randn(*cshape, generator=g).to(dtype)
    i = capi.TypedPlan(capi.INVERSE, dims, batch, dtype)
    back = torch.full(tuple(lead) + tuple(dims), float("nan"), dtype=dtype, device="cuda")
    sd = s16.cuda()
    i.exec(sd, back)
    torch.cuda.synchronize()
    assert not torch.isnan(back).any()
    assert torch.equal(sd.cpu(), s16)
    want_x = torch.fft.irfftn(torch.view_as_complex(s16.double().contiguous()), s=dims, dim=dim, norm="backward")
    err = flat(back.cpu().double() - want_x).norm(dim=1) / flat(want_x).norm(dim=1)
    assert float(err.max()) <= TOL[dtype], f"worst image {int(err.argmax())}: {float(err.max()):.3e}"
    close_to_own_fp32_path(back.cpu(), capi.irfftn(sd.float(), nd, last=dims[-1]).cpu())


def test_typed_plan_routes_and_argument_checks():
    from trt_dft_plugins import capi
    p = capi.TypedPlan(capi.FORWARD, (90, 180), 64, capi.IO_F16)
    n, text = p.describe()
    assert n == 1 and "converts as it loads and stores" in text and p.workspace_bytes == 0, text
    for dims in ((128, 128), (64, 64)):
        n, text = capi.TypedPlan(capi.INVERSE, dims, 8, capi.IO_BF16).describe()
        assert n == 1 and "converts as it loads and stores" in text, text
    q = capi.TypedPlan(capi.INVERSE, (60, 120), 8, capi.IO_BF16)
    n, text = q.describe()
    assert n == 3 and "bf16 -> fp32" in text and "fp32 -> bf16" in text, text
    # fp32 input + fp32 output + whatever the fp32 plan wants
    inner = capi.Plan(capi.INVERSE, (60, 120), 8)
    assert q.workspace_bytes >= (q.input_floats + q.output_floats) * 4 + inner.workspace_bytes
    # IO_F32 is the ordinary constructor
    r = capi.TypedPlan(capi.FORWARD, (90, 180), 4, capi.IO_F32)
    assert r.io_type == capi.IO_F32 and r.describe()[1] == capi.Plan(capi.FORWARD, (90, 180), 4).describe()[1]
    with pytest.raises(capi.DftError):
        capi.TypedPlan(capi.FORWARD, (90, 180), 4, 3)
    with pytest.raises(capi.DftError):
        capi.TypedPlan(capi.FORWARD, (0, 180), 4, capi.IO_F16)
    # an empty batch is a valid plan that does nothing
    e = capi.TypedPlan(capi.FORWARD, (90, 180), 0, capi.IO_F16)
    assert e.describe()[0] == 0
    e.exec_ptr(0, 0, 0, 0)
    # the converting kernels move whole images with bulk copies: a tensor that is not 16-byte aligned is refused
    x = torch.zeros(8 * 90 * 180 + 8, dtype=torch.float16, device="cuda")
    y = torch.zeros(8 * 90 * 91 * 2, dtype=torch.float16, device="cuda")
    p4 = capi.TypedPlan(capi.FORWARD, (90, 180), 4, capi.IO_F16)
    with pytest.raises(capi.DftError):
        p4.exec_ptr(x.data_ptr() + 2, y.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    # the conversion route needs its workspace
    with pytest.raises(capi.DftError):
        q.exec_ptr(y.data_ptr(), x.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dims,batch", [((90, 180), 37), ((60, 120), 5)])
def test_typed_exec_host(dims, batch):
    from trt_dft_plugins import capi
    g = torch.Generator().manual_seed(3)
    x16 = torch.randn(batch, *dims, generator=g).to(torch.float16)
    out = torch.empty(batch, dims[0], dims[1] // 2 + 1, 2, dtype=torch.float16)
    f = capi.TypedPlan(capi.FORWARD, dims, batch, capi.IO_F16)
    f.exec_host(x16, out)
    f.exec_host(x16, out)                                   # second call reuses the device buffers
    want = torch.view_as_real(torch.fft.rfft2(x16.double(), norm="backward"))
    err = (out.double() - want).norm() / want.norm()
    assert float(err) <= TOL[torch.float16]


def test_round_trip_through_16_bit_tensors():
    """rfft2 then irfft2 with fp16 at every boundary: the spectrum's rounding comes back as noise of about its own relative
    size (each output sample averages H*W/2 independent roundings, scaled by 1/(H*W))."""
    from trt_dft_plugins import capi
    x16 = torch.randn(300, 90, 180, generator=torch.Generator().manual_seed(5)).to(torch.float16).cuda()
    y = torch.empty(300, 90, 91, 2, dtype=torch.float16, device="cuda")
    back = torch.empty_like(x16)
    capi.TypedPlan(capi.FORWARD, (90, 180), 300, capi.IO_F16).exec(x16, y)
    capi.TypedPlan(capi.INVERSE, (90, 180), 300, capi.IO_F16).exec(y, back)
    torch.cuda.synchronize()
    err = (back.double() - x16.double()).norm() / x16.double().norm()
    assert float(err) <= 6e-4, float(err)
