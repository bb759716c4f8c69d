"""GPU parity: the CUDA path (through the C ABI of include/trt_dft.h) against the oracle.

Oracle = the reference tests' own expected-value expressions (CPU torch.fft,
/root/reference/tests/test_dft.py:40,:52-54) in fp64, the committed golden vectors, and the plain-C
restatement in oracle/.  Tolerances are north_star's: rel-L2 <= 1e-5 against fp64, round trip
<= 1e-6, plus the reference's own ``torch.allclose`` (rtol 1e-5, atol 1e-8) on its 16 cases.
"""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

GOLD = np.load(Path(__file__).parent / "golden" / "reference_cases.npz")
NAMES = [str(n) for n in GOLD["names"]]
REF = [n for n in NAMES if n.startswith("ref_")]
EXT = [n for n in NAMES if n.startswith("ext_")]

FWD_TOL = 1e-5      # rel-L2 vs fp64 torch.fft (north_star)
RT_TOL = 1e-6       # round trip irfft(rfft(x)) (north_star)


@pytest.fixture(scope="module")
def capi():
    from trt_dft_plugins import capi as c
    c.lib()
    return c


def dims_of(nd):
    return tuple(range(-nd, 0))


# ---- the reference's own 8 + 8 cases, with the reference's own assertion ---------------------
@pytest.mark.parametrize("name", REF)
def test_reference_rfft2_cases(capi, name):
    x = torch.from_numpy(GOLD[name + "_x"])
    y = capi.rfftn(x.cuda(), 2).cpu()
    y_expected = torch.from_numpy(GOLD[name + "_rfft2"])
    assert torch.allclose(y_expected, y)                       # tests/test_dft.py:151
    assert O.rel_l2(y.numpy(), GOLD[name + "_rfft2_f64"]) <= FWD_TOL


@pytest.mark.parametrize("name", REF)
def test_reference_irfft2_cases(capi, name):
    y = torch.from_numpy(GOLD[name + "_rfft2"])
    x = capi.irfftn(y.cuda(), 2).cpu()
    x_expected = torch.from_numpy(GOLD[name + "_irfft2"])
    assert torch.allclose(x_expected, x)                       # tests/test_dft.py:184, the reference's defaults (rtol 1e-5, atol 1e-8)
    assert O.rel_l2(x.numpy(), GOLD[name + "_irfft2_f64"]) <= FWD_TOL


@pytest.mark.parametrize("name", EXT)
def test_golden_extended(capi, name):
    nd = int(name.rsplit("_nd", 1)[1])
    x = torch.from_numpy(GOLD[name + "_x"])
    y = capi.rfftn(x.cuda(), nd).cpu().numpy()
    assert O.rel_l2(y, GOLD[name + "_rfftn_f64"]) <= FWD_TOL
    xi = capi.irfftn(torch.from_numpy(GOLD[name + "_rfftn"]).cuda(), nd, last=x.shape[-1]).cpu().numpy()
    assert O.rel_l2(xi, GOLD[name + "_irfftn_f64"]) <= FWD_TOL
    xz = capi.irfftn(torch.from_numpy(GOLD[name + "_z"]).cuda(), nd, last=x.shape[-1]).cpu().numpy()
    assert O.rel_l2(xz, GOLD[name + "_z_irfftn_f64"]) <= FWD_TOL


# ---- seeded shapes against the fp64 oracle ------------------------------------------------------
SHAPES = [
    # BASELINE configs 1-3 in full
    ((4, 32, 64), 1), ((16, 64, 128, 128), 2), ((8, 768, 90, 180), 2),
    # 1-D: a few short rows run one thread per row in registers (W = 4 .. 64, up to 8192 rows), one row more the engine
    ((5, 4), 1), ((7, 8), 1), ((100, 16), 1), ((33, 32), 1), ((8192, 64), 1), ((8193, 64), 1),
    # 1-D
    ((3, 2), 1), ((3, 6), 1), ((5, 30), 1), ((2, 9), 1), ((2, 15), 1), ((3, 1), 1), ((2, 360), 1),
    ((1000, 1024), 1), ((2, 2 * 7 * 11 * 13), 1), ((2, 34), 1), ((3, 19), 1), ((3, 4096), 1), ((2, 8192), 1),
    ((2, 50000), 1), ((2, 65536), 1), ((1203, 1440), 1),   # persistent two-step row kernels, last tile partial
    # prime factors above 31: generic prime-radix steps in shared memory (1009 and 4099 are prime)
    ((40, 74), 1), ((3, 2 * 37 * 41), 1), ((300, 1009), 1), ((5, 4099), 1), ((2, 8 * 53), 1), ((6, 43, 86), 2), ((2, 202, 74), 2),
    ((1, 3, 74, 6), 3), ((2, 2018, 40), 2),
    # 2-D
    ((7, 8, 8), 2), ((3, 6, 10), 2), ((2, 5, 7), 2), ((2, 12, 18), 2), ((2, 17, 6), 2), ((2, 6, 38), 2),
    ((700, 4, 4), 2), ((1, 1, 16), 2), ((2, 16, 1), 2), ((2, 3, 3), 2), ((300, 32, 32), 2), ((5, 64, 64), 2),
    ((7, 64, 128), 2), ((1030, 64, 128), 2), ((5, 128, 64), 2), ((1000, 128, 64), 2), ((2000, 64, 64), 2),   # fused rectangular FNO grids, looping and one-CTA-per-image grids
    ((3, 256, 256), 2), ((2, 180, 360), 2), ((2, 720, 1440), 2), ((4, 720, 1440), 2), ((1, 1024, 1024), 2), ((1, 2048, 4096), 2), ((2, 4096, 512), 2), ((1, 8192, 1024), 2), ((1, 5120, 320), 2),
    # 3-D
    ((2, 4, 6, 8), 3), ((2, 3, 5, 6), 3), ((1, 2, 3, 5, 7), 3), ((2, 1, 4, 4), 3), ((1, 17, 4, 6), 3),
    ((2, 32, 32, 32), 3), ((1, 20, 90, 180), 3),
]


@pytest.mark.parametrize("shape,nd", SHAPES)
def test_seeded_shapes_vs_fp64_oracle(capi, shape, nd):
    torch.manual_seed(1)                                      # tests/test_dft.py:135
    x = torch.randn(*shape)
    want = O.torch_rfftn(x, nd, dtype=torch.float64)
    xd = x.cuda()
    y = capi.rfftn(xd, nd)
    assert O.rel_l2(y.cpu().numpy(), want.numpy()) <= FWD_TOL
    # inverse of the oracle's fp32 spectrum
    y32 = want.float()
    s = shape[-nd:]
    want_i = torch.fft.irfftn(torch.view_as_complex(y32.double()), s=s, dim=dims_of(nd), norm="backward")
    xi = capi.irfftn(y32.cuda(), nd, last=shape[-1])
    assert O.rel_l2(xi.cpu().numpy(), want_i.numpy()) <= FWD_TOL
    # round trip through the CUDA path only
    rt = capi.irfftn(y, nd, last=shape[-1])
    assert O.rel_l2(rt.cpu().numpy(), x.numpy()) <= RT_TOL
    # input must not be modified (the reference const_casts it for cuFFT, dft_plugins.cpp:193)
    assert torch.equal(xd.cpu(), x)


@pytest.mark.parametrize("shape,nd", [((2, 3, 6, 10), 2), ((4, 32, 64), 1), ((2, 90, 180), 2), ((1, 4, 6, 8), 3)])
def test_c_oracle_agrees_with_gpu(capi, shape, nd):
    """Independent check against the plain-C naive DFT (no MKL / pocketfft in the loop)."""
    torch.manual_seed(1)
    x = torch.randn(*shape)
    y = capi.rfftn(x.cuda(), nd).cpu().numpy()
    assert O.rel_l2(y, O.c_rfftn(x.numpy(), nd)) <= FWD_TOL
    xi = capi.irfftn(torch.from_numpy(y).cuda(), nd).cpu().numpy()
    assert O.rel_l2(xi, O.c_irfftn(y, nd)) <= FWD_TOL


def test_linearity_and_parseval_at_full_size(capi):
    """Size-independent properties at cfg3's full size (no CPU oracle in the loop)."""
    torch.manual_seed(1)
    shape = (8, 768, 90, 180)
    a = torch.randn(*shape, device="cuda")
    b = torch.randn(*shape, device="cuda")
    fa, fb = capi.rfftn(a, 2), capi.rfftn(b, 2)
    fab = capi.rfftn(2.0 * a - 3.0 * b, 2)
    lin = (fab - (2.0 * fa - 3.0 * fb)).double().norm() / fab.double().norm()
    assert lin.item() <= 1e-6
    # Parseval with one-sided weights: sum |x|^2 = (1/N) sum_k w_k |X_k|^2
    X = torch.view_as_complex(fa).abs().double() ** 2
    w = torch.full((91,), 2.0, dtype=torch.float64, device="cuda")
    w[0] = 1.0
    w[90] = 1.0
    lhs = (a.double() ** 2).sum()
    rhs = (X * w).sum() / (90 * 180)
    assert abs(lhs.item() - rhs.item()) / lhs.item() <= 1e-6
    rt = capi.irfftn(fa, 2)
    assert ((rt - a).double().norm() / a.double().norm()).item() <= RT_TOL


def test_empty_batch_and_errors(capi):
    p = capi.Plan(capi.FORWARD, (8, 8), 0)
    assert p.num_launches == 0
    p.exec_ptr(0, 0)           # no-op
    with pytest.raises(capi.DftError):
        capi.Plan(capi.FORWARD, (8, 0), 1)
    with pytest.raises(capi.DftError):
        capi.Plan(capi.FORWARD, (2, 2, 2, 2), 1)
    with pytest.raises(capi.DftError):
        capi.Plan(2, (8,), 1)
    q = capi.Plan(capi.INVERSE, (720, 1440), 1)
    assert q.workspace_bytes == 720 * 721 * 8
    with pytest.raises(capi.DftError):
        x = torch.zeros(720 * 721 * 2, device="cuda")
        o = torch.zeros(720 * 1440, device="cuda")
        q.exec_ptr(x.data_ptr(), o.data_ptr(), 0)     # workspace required


def test_misaligned_real_input_takes_the_scalar_path(capi):
    torch.manual_seed(1)
    base = torch.randn(2 * 6 * 10 + 1)
    x = base[1:].view(2, 6, 10)
    big = base.cuda()
    xd = big[1:].view(2, 6, 10)          # 4-byte aligned only
    out = torch.empty(2, 6, 6, 2, device="cuda")
    p = capi.Plan(capi.FORWARD, (6, 10), 2)
    p.exec_ptr(xd.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    want = O.torch_rfftn(x.contiguous(), 2, dtype=torch.float64).numpy()
    assert O.rel_l2(out.cpu().numpy(), want) <= FWD_TOL


def test_exec_host_and_streams(capi):
    torch.manual_seed(1)
    x = torch.randn(6, 90, 180).pin_memory()
    y = torch.empty(6, 90, 91, 2).pin_memory()
    p = capi.Plan(capi.FORWARD, (90, 180), 6)
    p.exec_host(x, y)
    p.exec_host(x, y)
    want = O.torch_rfftn(x, 2, dtype=torch.float64).numpy()
    assert O.rel_l2(y.numpy(), want) <= FWD_TOL
    s = torch.cuda.Stream()
    xd = x.cuda()
    out = torch.empty(6, 90, 91, 2, device="cuda")
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        p.exec(xd, out)
    s.synchronize()
    assert O.rel_l2(out.cpu().numpy(), want) <= FWD_TOL


@pytest.mark.parametrize("outer,n,inner", [(2, 90, 91), (3, 64, 5), (1, 720, 37), (2, 1, 8), (1, 17, 6), (1, 8192, 12), (2, 8192, 300), (1, 4096, 70),
                                           (4, 720, 721), (3, 720, 800)])   # persistent two-step column kernel (>= 148 tiles)
@pytest.mark.parametrize("direction", [0, 1])
def test_c2c_strided_axis_plan(capi, outer, n, inner, direction):
    """trt_dft_plan_create_c2c: the column pass on its own (used by the pencil decomposition)."""
    torch.manual_seed(1)
    z = torch.randn(outer, n, inner, 2)
    zc = torch.view_as_complex(z.double())
    want = torch.fft.fft(zc, dim=1) if direction == 0 else torch.fft.ifft(zc, dim=1, norm="forward")
    plan = capi.C2CPlan(direction, n, outer, inner, scale=0.5)
    zd = z.cuda()
    out = torch.empty_like(zd)
    plan.exec(zd, out)
    torch.cuda.synchronize()
    assert O.rel_l2(out.cpu().numpy(), 0.5 * torch.view_as_real(want).numpy()) <= FWD_TOL
    plan.exec(zd, zd)                                   # in place
    torch.cuda.synchronize()
    assert torch.equal(zd, out)


@pytest.mark.parametrize("shape,nd", [((300, 90, 180), 2), ((2, 2048, 4100), 2), ((64, 128, 128), 2), ((3000, 60, 120), 2), ((40000, 64), 1)])
def test_one_plan_on_two_streams_at_once(capi, shape, nd):
    """include/trt_dft.h: a plan is immutable and may be used from several streams at once.  Two streams run the same
    forward and inverse plans concurrently on different tensors (each with a workspace of its own), several rounds
    without synchronising in between; both results must be what a lone call gives."""
    batch = 1
    for v in shape[:-nd]:
        batch *= v
    dims = shape[-nd:]
    f = capi.Plan(capi.FORWARD, dims, batch)
    i = capi.Plan(capi.INVERSE, dims, batch)
    need = max(f.workspace_bytes, i.workspace_bytes, 16)
    g = torch.Generator().manual_seed(11)
    xs = [torch.randn(*shape, generator=g).cuda() for _ in range(2)]
    ys = [torch.empty(shape[:-1] + (dims[-1] // 2 + 1, 2), device="cuda") for _ in range(2)]
    xr = [torch.empty_like(xs[0]) for _ in range(2)]
    ws = [torch.empty(need, dtype=torch.uint8, device="cuda") for _ in range(2)]
    alone_y = capi.rfftn(xs[0], nd), capi.rfftn(xs[1], nd)
    alone_x = capi.irfftn(alone_y[0], nd, last=dims[-1]), capi.irfftn(alone_y[1], nd, last=dims[-1])
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for _ in range(4):
        for k in range(2):
            with torch.cuda.stream(streams[k]):
                f.exec(xs[k], ys[k], workspace=ws[k])
                i.exec(ys[k], xr[k], workspace=ws[k])
    torch.cuda.synchronize()
    for k in range(2):
        assert torch.equal(ys[k], alone_y[k])
        assert torch.equal(xr[k], alone_x[k])


def test_exec_is_capturable_in_a_cuda_graph(capi):
    """trt_dft_exec only launches kernels on the given stream, so a whole rfft2 -> irfft2 pair can be
    captured once and replayed (what TensorRT does around enqueue)."""
    torch.manual_seed(1)
    x = torch.randn(6, 90, 180, device="cuda")
    y = torch.empty(6, 90, 91, 2, device="cuda")
    xr = torch.empty_like(x)
    f = capi.Plan(capi.FORWARD, (90, 180), 6)
    i = capi.Plan(capi.INVERSE, (90, 180), 6)
    f.exec(x, y)
    i.exec(y, xr)                              # warm-up outside capture (module load, attributes)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        f.exec(x, y)
        i.exec(y, xr)
    x.copy_(torch.randn(6, 90, 180, device="cuda"))
    xr.zero_()
    g.replay()
    torch.cuda.synchronize()
    want = O.torch_rfftn(x.cpu(), 2, dtype=torch.float64).numpy()
    assert O.rel_l2(y.cpu().numpy(), want) <= FWD_TOL
    assert O.rel_l2(xr.cpu().numpy(), x.cpu().numpy()) <= RT_TOL


@pytest.mark.parametrize("nimg", [1, 2, 3, 5, 147, 148, 149, 296, 297, 445, 593])
def test_headline_kernels_at_every_batch_remainder(capi, nimg):
    """The persistent 90x180 kernels deal images round-robin to 148 CTAs x 2 ping-ponging groups x 3 stages:
    cover batches below, at and just above each of those periods (odd counts leave one group an image short),
    every image checked against fp64 torch.fft, and the inverse against the exact inverse of what it was fed."""
    g = torch.Generator().manual_seed(nimg)
    x = torch.randn(nimg, 90, 180, generator=g)
    want = O.torch_rfftn(x, 2, dtype=torch.float64)
    y = capi.rfftn(x.cuda(), 2)
    torch.cuda.synchronize()
    err = (y.cpu().double() - want).flatten(1).norm(dim=1) / want.flatten(1).norm(dim=1)
    assert float(err.max()) <= FWD_TOL, f"worst image {int(err.argmax())}"
    # a spectrum that is NOT Hermitian-consistent in its DC / Nyquist columns, as a network would produce
    s = torch.randn(nimg, 90, 91, 2, generator=g)
    want_x = O.torch_irfftn(s, 2, dtype=torch.float64)
    xr = capi.irfftn(s.cuda(), 2)
    torch.cuda.synchronize()
    err = (xr.cpu().double() - want_x).flatten(1).norm(dim=1) / want_x.flatten(1).norm(dim=1)
    assert float(err.max()) <= FWD_TOL, f"worst image {int(err.argmax())}"


@pytest.mark.parametrize("W", [1440, 512, 90])
def test_column_split_output_matches_plain_rows(capi, W):
    """trt_dft_plan_set_column_split after plan creation: the row pass must write every image's spectrum split by
    destination rank whichever kernel the plan had picked for plain rows (1440-point rows run a specialised
    persistent kernel that only writes plain rows, so the plan has to switch kernels)."""
    import ctypes
    hn, nimg = 720, 2
    Mo = W // 2 + 1
    torch.manual_seed(3)
    x = torch.randn(nimg * hn, W, device="cuda")
    plain = capi.rfftn(x, 1)                                   # [rows, Mo, 2]
    plan = capi.Plan(capi.FORWARD, (W,), nimg * hn)
    starts = [0, Mo // 3, Mo // 3, Mo - 5, Mo]                  # uneven, one empty part
    L = capi.lib()
    L.trt_dft_plan_set_column_split.restype = ctypes.c_int
    L.trt_dft_plan_set_column_split.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int64]
    arr = (ctypes.c_int64 * len(starts))(*starts)
    assert L.trt_dft_plan_set_column_split(plan.handle, len(starts) - 1, arr, hn) == 0
    out = torch.full((nimg * hn * Mo, 2), float("nan"), device="cuda")
    plan.exec(x, out)
    torch.cuda.synchronize()
    want = []
    p3 = plain.view(nimg, hn, Mo, 2)
    for b in range(nimg):
        for j in range(len(starts) - 1):
            want.append(p3[b, :, starts[j]:starts[j + 1], :].reshape(-1, 2))
    want = torch.cat(want)
    # (two different kernels may have produced `plain` and `out`: same values up to fp32 rounding, every element placed)
    assert not torch.isnan(out).any()
    assert O.rel_l2(out.cpu().numpy(), want.cpu().numpy()) <= RT_TOL
    # and back to plain rows
    assert L.trt_dft_plan_set_column_split(plan.handle, 0, None, 0) == 0
    out2 = torch.empty_like(plain)
    plan.exec(x, out2)
    torch.cuda.synchronize()
    assert torch.equal(out2, plain)


@pytest.mark.parametrize("shape,nd,off_floats", [((300, 90, 180), 2, 1), ((300, 90, 180), 2, 2), ((40, 128, 128), 2, 1),
                                                  ((1300, 1440), 1, 1), ((1300, 1440), 1, 2), ((2, 720, 1440), 2, 2)])
def test_specialised_shapes_with_unaligned_buffers(capi, shape, nd, off_floats):
    """The shape-specialised kernels move whole images with 16-byte bulk copies; buffers that are only 4- or
    8-byte aligned (a view into a larger tensor) must take another kernel and still give the same transform."""
    torch.manual_seed(2)
    n = 1
    for s in shape:
        n *= s
    base = torch.randn(n + off_floats)
    x = base[off_floats:].view(*shape)
    xd = base.cuda()[off_floats:].view(*shape)
    batch, dims = 1, shape[-nd:]
    for s in shape[:-nd]:
        batch *= s
    oshape = shape[:-1] + (shape[-1] // 2 + 1, 2)
    obase = torch.full((n // shape[-1] * oshape[-2] * 2 + 2,), float("nan"), device="cuda")
    out = obase[2:].view(*oshape)                               # complex side: 8-byte aligned, not 16
    f = capi.Plan(capi.FORWARD, dims, batch)
    ws = torch.empty(max(f.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
    f.exec_ptr(xd.data_ptr(), out.data_ptr(), ws.data_ptr())
    torch.cuda.synchronize()
    want = O.torch_rfftn(x.contiguous(), nd, dtype=torch.float64).numpy()
    assert O.rel_l2(out.cpu().numpy(), want) <= FWD_TOL
    if off_floats % 2 == 0:                                     # real output rows are written as float2 pairs
        i = capi.Plan(capi.INVERSE, dims, batch)
        ws = torch.empty(max(i.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
        back = torch.full((n + off_floats,), float("nan"), device="cuda")
        i.exec_ptr(out.data_ptr(), back[off_floats:].data_ptr(), ws.data_ptr())
        torch.cuda.synchronize()
        assert O.rel_l2(back[off_floats:].view(*shape).cpu().numpy(), x.numpy()) <= RT_TOL


# ---- BASELINE configs 4 and 5 at sizes that cross the 32-bit boundaries -------------------------------------
def _check_images(got_dev, want_fn, which, tol):
    for b in which:
        want = want_fn(b)
        err = O.rel_l2(got_dev[b].cpu().numpy(), want.numpy())
        assert err <= tol, f"image {b}: rel-L2 {err:.3e}"


def test_cfg4_shape_with_offsets_beyond_2_31_bytes(capi):
    """BASELINE config 4's image size (720 x 1440) at a batch whose last images start beyond 2^31 BYTES and beyond
    2^31 ELEMENTS/4 of both tensors (600 images: 2.49 GB per side; SURVEY Q12 -- the reference's batch is an
    int32).  First, a middle and the LAST image are checked against fp64 torch.fft in both directions; the rest is
    covered by a checksum identity (sum over the batch of the transforms = transform of the sum, linearity)."""
    n, H, W = 600, 720, 1440
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(n, H, W, device="cuda", generator=g)
    assert x.numel() * 4 > 2**31 and (n - 1) * H * W * 4 > 2**31
    f = capi.Plan(capi.FORWARD, (H, W), n)
    y = torch.full((n, H, W // 2 + 1, 2), float("nan"), device="cuda")
    wsf = torch.empty(max(f.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
    f.exec(x, y, workspace=wsf)
    torch.cuda.synchronize()
    which = (0, 1, n // 2, n - 2, n - 1)
    _check_images(y, lambda b: O.torch_rfftn(x[b].cpu(), 2, dtype=torch.float64), which, FWD_TOL)
    assert not torch.isnan(y).any()
    # linearity over the whole batch: catches any image that was skipped, duplicated or misplaced
    ysum = y.double().sum(dim=0)
    fsum = capi.rfftn(x.double().sum(dim=0).float().unsqueeze(0), 2)[0].double()
    assert ((ysum - fsum).norm() / fsum.norm()).item() <= 1e-5
    del ysum, fsum, wsf
    i = capi.Plan(capi.INVERSE, (H, W), n)
    ws = torch.empty(max(i.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
    xr = torch.full((n, H, W), float("nan"), device="cuda")
    y_before = y[n - 1].clone()
    i.exec(y, xr, workspace=ws)
    torch.cuda.synchronize()
    assert torch.equal(y[n - 1], y_before)                           # input never written (dft_plugins.cpp:193)
    _check_images(xr, lambda b: O.torch_irfftn(y[b].cpu(), 2, dtype=torch.float64), which, FWD_TOL)
    per_img = (xr - x).flatten(1).double().norm(dim=1) / x.flatten(1).double().norm(dim=1)
    assert float(per_img.max()) <= RT_TOL, f"worst image {int(per_img.argmax())}"
    del x, y, xr, ws
    torch.cuda.empty_cache()


def test_cfg5_full_size_both_directions(capi):
    """BASELINE config 5 on one GPU, [4, 8192, 8192]: every image's round trip, one image (the last) in full
    against fp64 torch.fft in both directions."""
    B, H, W = 4, 8192, 8192
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(B, H, W, device="cuda", generator=g)
    f = capi.Plan(capi.FORWARD, (H, W), B)
    i = capi.Plan(capi.INVERSE, (H, W), B)
    ws = torch.empty(max(f.workspace_bytes, i.workspace_bytes, 16), dtype=torch.uint8, device="cuda")
    y = torch.full((B, H, W // 2 + 1, 2), float("nan"), device="cuda")
    f.exec(x, y, workspace=ws)
    torch.cuda.synchronize()
    want = O.torch_rfftn(x[B - 1].cpu(), 2, dtype=torch.float64)
    assert O.rel_l2(y[B - 1].cpu().numpy(), want.numpy()) <= FWD_TOL
    assert not torch.isnan(y).any()
    xr = torch.full((B, H, W), float("nan"), device="cuda")
    i.exec(y, xr, workspace=ws)
    torch.cuda.synchronize()
    want_x = O.torch_irfftn(y[B - 1].cpu(), 2, dtype=torch.float64)
    assert O.rel_l2(xr[B - 1].cpu().numpy(), want_x.numpy()) <= FWD_TOL
    per_img = (xr - x).flatten(1).double().norm(dim=1) / x.flatten(1).double().norm(dim=1)
    assert float(per_img.max()) <= RT_TOL, f"worst image {int(per_img.argmax())}"
    del x, y, xr, ws
    torch.cuda.empty_cache()


def test_reference_irfft2_cases_at_the_reference_defaults(capi):
    """tests/test_dft.py:184 asserts torch.allclose at the defaults (rtol 1e-5, atol 1e-8), and so does the parametrised
    test above: all 8 cases pass them on B200.  fp32 cannot promise an absolute 1e-8 on elements that cancel to ~0 in
    general, so this test also pins the fallback statement: any element that ever misses the defaults must be off by less
    than 1e-6 absolute (a few fp32 ulps of the O(1) intermediates)."""
    passed = 0
    for name in REF:
        y = torch.from_numpy(GOLD[name + "_rfft2"])
        x = capi.irfftn(y.cuda(), 2).cpu()
        want = torch.from_numpy(GOLD[name + "_irfft2"])
        ok = torch.isclose(want, x)                                  # the reference's own criterion
        passed += int(bool(ok.all()))
        bad = ~ok
        assert ((want - x)[bad].abs() < 1e-6).all(), name
    print(f"irfft2 reference cases passing torch.allclose defaults: {passed} of {len(REF)}")


def test_in_place_c2c_on_a_length_outside_the_radix_set(capi):
    """Lengths with a prime factor above 8191 run the naive column kernel, which must not run in place: the plan goes
    through its workspace instead (many blocks, so a race would show)."""
    torch.manual_seed(7)
    outer, n, inner = 1, 8209, 64                                    # 8209 is prime
    z = torch.randn(outer, n, inner, 2)
    want = torch.view_as_real(torch.fft.fft(torch.view_as_complex(z.double()), dim=1))
    plan = capi.C2CPlan(capi.FORWARD, n, outer, inner)
    assert plan.workspace_bytes == outer * n * inner * 8
    zd = z.cuda()
    plan.exec(zd, zd)
    torch.cuda.synchronize()
    assert O.rel_l2(zd.cpu().numpy(), want.numpy()) <= FWD_TOL
    # the strided 8192-point axis (four-step) also needs its workspace: NULL must be refused, not dereferenced
    p2 = capi.C2CPlan(capi.FORWARD, 8192, 1, 12)
    assert p2.workspace_bytes > 0
    w = torch.zeros(8192 * 12 * 2, device="cuda")
    with pytest.raises(capi.DftError):
        p2.exec_ptr(w.data_ptr(), w.data_ptr(), 0)


def test_exec_host_with_a_c2c_plan(capi):
    """trt_dft_exec_host on a c2c plan large enough for the chunked host pipeline (>= 64 MiB): the batch of a c2c plan
    is not a contiguous leading dimension, so it must run as one piece."""
    torch.manual_seed(8)
    outer, n, inner = 3, 720, 2100                                   # 2 x 36 MB
    z = torch.randn(outer, n, inner, 2)
    out = torch.empty_like(z)
    plan = capi.C2CPlan(capi.FORWARD, n, outer, inner)
    plan.exec_host(z, out)
    want = torch.view_as_real(torch.fft.fft(torch.view_as_complex(z[1].double()), dim=0))
    assert O.rel_l2(out[1].numpy(), want.numpy()) <= FWD_TOL


def test_exec_refuses_a_plan_from_another_device(capi):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    p = capi.Plan(capi.FORWARD, (90, 180), 2)                        # built on device 0
    with torch.cuda.device(1):
        x = torch.zeros(2, 90, 180, device="cuda:1")
        y = torch.zeros(2, 90, 91, 2, device="cuda:1")
        with pytest.raises(capi.DftError):
            p.exec_ptr(x.data_ptr(), y.data_ptr())
