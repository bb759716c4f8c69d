"""The Rfft / Irfft plugin shell driven by a mock TensorRT host (tests/mock_trt).

TensorRT is absent from this image, so the reference's tests (export -> parse -> build -> execute,
/root/reference/tests/test_dft.py:73-184) cannot run.  The mock provides getPluginRegistry() and walks
the same IPluginV2DynamicExt lifecycle, which pins the boundary of SURVEY.md 8(b): registration by
name/version/namespace, creator fields, attribute validation, shape inference, format negotiation,
serialisation, clone, workspace and enqueue.  CPU tests cover everything up to enqueue; the gpu tests
run enqueue and compare with the oracle exactly as the reference tests compare with torch.fft.
"""
import ctypes
import struct
import subprocess
import sys
from ctypes import c_char_p, c_int, c_int32, c_size_t, c_void_p
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
MOCK_DIR = Path(__file__).parent / "mock_trt"


@pytest.fixture(scope="module")
def mock():
    subprocess.check_call(["make", "-C", str(MOCK_DIR), "-s"])
    m = ctypes.CDLL(str(MOCK_DIR / "libmock_nvinfer.so"), mode=ctypes.RTLD_GLOBAL)
    from trt_dft_plugins import load_plugins
    lib = load_plugins()
    if m.mock_registry_count() == 0:
        # the plugin library was loaded earlier in this process (before the mock existed): hand the mock
        # its creators the way IPluginRegistry::loadLibrary would
        m.mock_register_from.argtypes = [c_void_p]
        m.mock_register_from(ctypes.cast(lib.getPluginCreators, c_void_p))
    m.mock_registry_entry.argtypes = [c_int, c_char_p, c_size_t]
    m.mock_try_create.argtypes = [c_char_p, c_int32, c_int32, c_int32]
    m.mock_output_dims.argtypes = [c_char_p, c_int32, c_int32, ctypes.POINTER(c_int32), ctypes.POINTER(c_int32)]
    m.mock_format_ok.argtypes = [c_char_p, c_int32, c_int32, c_int32]
    m.mock_format_pair_ok.argtypes = [c_char_p, c_int32, c_int32]
    m.mock_set_tensor_type.argtypes = [c_int32]
    m.mock_set_tensor_type.restype = None
    m.mock_serialize_roundtrip.argtypes = [c_char_p, c_int32, ctypes.c_char_p, c_size_t]
    m.mock_workspace_size.argtypes = [c_char_p, c_int32, c_int32, ctypes.POINTER(c_int32)]
    m.mock_workspace_size.restype = c_size_t
    m.mock_lifecycle.argtypes = [c_char_p, c_int32, c_int32, ctypes.POINTER(c_int32), c_void_p, c_void_p, c_void_p,
                                 c_int, ctypes.POINTER(c_int32), ctypes.POINTER(c_int32)]
    m.mock_capture_clone.argtypes = [c_char_p, c_int32, c_int32, ctypes.POINTER(c_int32), c_void_p, c_void_p, c_int32]
    return m


def arr(vals):
    return (c_int32 * len(vals))(*vals)


def test_static_registration_in_a_fresh_process():
    """tests/test_dft.py:118-121: after load_plugins(), "Rfft" and "Irfft" are in the registry."""
    subprocess.check_call(["make", "-C", str(MOCK_DIR), "-s"])
    code = f"""
import ctypes, sys
sys.path.insert(0, r"{ROOT / 'tensorrt-dft-plugins_b200'}")
m = ctypes.CDLL(r"{MOCK_DIR / 'libmock_nvinfer.so'}", mode=ctypes.RTLD_GLOBAL)   # stands in for libnvinfer
assert m.mock_registry_count() == 0
from trt_dft_plugins import load_plugins
load_plugins()                                   # static initialisers register the creators
names = set()
for i in range(m.mock_registry_count()):
    b = ctypes.create_string_buffer(256); m.mock_registry_entry(i, b, 256); names.add(b.value.decode().split("|")[0])
assert names == {{"Rfft", "Irfft"}}, names
print("ok")
"""
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr


def test_library_loads_without_tensorrt():
    """No libnvinfer in the process: the library must still load (weak getPluginRegistry)."""
    code = f"""
import sys
sys.path.insert(0, r"{ROOT / 'tensorrt-dft-plugins_b200'}")
from trt_dft_plugins import load_plugins
load_plugins(); print("ok")
"""
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr


def test_creators_names_versions_fields(mock):
    entries = {}
    for i in range(mock.mock_registry_count()):
        b = ctypes.create_string_buffer(256)
        assert mock.mock_registry_entry(i, b, 256) == 0
        name, version, ns, fields = b.value.decode().split("|")
        entries[name] = (version, ns, fields)
    assert set(entries) == {"Rfft", "Irfft"}                       # dft_plugins.cpp:385, :475
    for version, ns, fields in entries.values():
        assert version == "1" and ns == ""                          # :43, registrar namespace ""
        assert fields == "normalized:5,onesided:5,signal_ndim:5"    # order and kINT32, :500-507


@pytest.mark.parametrize("op", [b"Rfft", b"Irfft"])
def test_attribute_validation(mock, op):
    for nd in (1, 2, 3):
        assert mock.mock_try_create(op, 0, 1, nd) == 0
    # the accepted set of dft_plugins.cpp:55-57; the reference aborts, we return nullptr
    assert mock.mock_try_create(op, 1, 1, 2) == 1
    assert mock.mock_try_create(op, 0, 0, 2) == 1
    assert mock.mock_try_create(op, 0, 1, 0) == 1
    assert mock.mock_try_create(op, 0, 1, 4) == 1
    assert mock.mock_try_create(b"Fft", 0, 1, 2) == 2


@pytest.mark.parametrize("dims,nd,want", [
    ((2, 3, 2, 4), 2, (2, 3, 2, 3, 2)),            # reference test shape: W=4 -> 3 bins
    ((4, 32, 64), 1, (4, 32, 33, 2)), ((8, 768, 90, 180), 2, (8, 768, 90, 91, 2)),
    ((5, 7), 1, (5, 4, 2)), ((2, 4, 6, 8), 3, (2, 4, 6, 5, 2)), ((9,), 1, (5, 2)),
])
def test_rfft_shape_inference(mock, dims, nd, want):
    out = (c_int32 * 8)()
    n = mock.mock_output_dims(b"Rfft", nd, len(dims), arr(dims), out)
    assert tuple(out[:n]) == want                                    # dft_plugins.cpp:361-382


@pytest.mark.parametrize("dims,nd,want", [
    ((2, 3, 2, 3, 2), 2, (2, 3, 2, 4)), ((4, 32, 33, 2), 1, (4, 32, 64)), ((8, 768, 90, 91, 2), 2, (8, 768, 90, 180)),
    ((5, 3, 2), 1, (5, 4)),                         # always even: m=3 -> 4 (odd lengths unrepresentable, :415-436)
])
def test_irfft_shape_inference(mock, dims, nd, want):
    out = (c_int32 * 8)()
    n = mock.mock_output_dims(b"Irfft", nd, len(dims), arr(dims), out)
    assert tuple(out[:n]) == want


@pytest.mark.parametrize("op", [b"Rfft", b"Irfft"])
def test_format_negotiation(mock, op):
    FLOAT, HALF, INT32, LINEAR, CHW4 = 0, 1, 3, 0, 3
    for pos in (0, 1):
        assert mock.mock_format_ok(op, pos, FLOAT, LINEAR) == 1      # dft_plugins.cpp:93-105
        assert mock.mock_format_ok(op, pos, HALF, LINEAR) == 0
        assert mock.mock_format_ok(op, pos, INT32, LINEAR) == 0
        assert mock.mock_format_ok(op, pos, FLOAT, CHW4) == 0


@pytest.mark.parametrize("op", [b"Rfft", b"Irfft"])
def test_16_bit_tensors_are_offered_only_on_request(mock, op, monkeypatch):
    """Like the reference (dft_plugins.cpp:93-105) the plugins take fp32 only -- unless TRT_DFT_PLUGIN_16BIT_IO=1 is set
    while the engine is built; then fp16 on BOTH sides is a valid combination as well (mixed types never are)."""
    FLOAT, HALF, INT8 = 0, 1, 2
    monkeypatch.delenv("TRT_DFT_PLUGIN_16BIT_IO", raising=False)
    assert mock.mock_format_pair_ok(op, FLOAT, FLOAT) == 1
    assert mock.mock_format_pair_ok(op, HALF, HALF) == 0
    monkeypatch.setenv("TRT_DFT_PLUGIN_16BIT_IO", "1")
    assert mock.mock_format_pair_ok(op, FLOAT, FLOAT) == 1
    assert mock.mock_format_pair_ok(op, HALF, HALF) == 1
    assert mock.mock_format_pair_ok(op, HALF, FLOAT) == 0
    assert mock.mock_format_pair_ok(op, FLOAT, HALF) == 0
    assert mock.mock_format_pair_ok(op, INT8, INT8) == 0
    # the conversion route's fp32 copies of both tensors are part of the workspace TensorRT is asked for
    mock.mock_set_tensor_type(HALF)
    try:
        want = 4 * 60 * 120 * 4 + 4 * 60 * 61 * 2 * 4
        assert mock.mock_workspace_size(b"Rfft", 2, 3, arr((4, 60, 120))) >= want
    finally:
        mock.mock_set_tensor_type(FLOAT)


@pytest.mark.parametrize("op", [b"Rfft", b"Irfft"])
@pytest.mark.parametrize("nd", [1, 2, 3])
def test_serialisation_is_the_reference_wire_format(mock, op, nd):
    buf = ctypes.create_string_buffer(64)
    n = mock.mock_serialize_roundtrip(op, nd, buf, 64)
    assert n == 12                                                   # dft_plugins.cpp:201-204
    assert struct.unpack("<3i", buf.raw[:12]) == (0, 1, nd)          # normalized, onesided, signal_ndim (:212-214)


def test_workspace_is_derived_from_the_descriptors(mock):
    assert mock.mock_workspace_size(b"Rfft", 2, 4, arr((8, 768, 90, 180))) == 0
    assert mock.mock_workspace_size(b"Irfft", 2, 5, arr((8, 768, 90, 91, 2))) == 0          # fused
    assert mock.mock_workspace_size(b"Rfft", 2, 3, arr((4, 720, 1440))) == 0                # columns run in place
    assert mock.mock_workspace_size(b"Irfft", 2, 4, arr((4, 720, 721, 2))) == 4 * 720 * 721 * 8


# ---- enqueue through the mock lifecycle, on the GPU -------------------------------------------
def run_plugin(mock, op, nd, x, skip_configure=False):
    xd = x.cuda().contiguous()
    out_dims = (c_int32 * 8)()
    nb = c_int32(0)
    n = mock.mock_output_dims(op, nd, x.dim(), arr(tuple(x.shape)), out_dims)
    y = torch.full(tuple(out_dims[:n]), float("nan"), dtype=x.dtype, device="cuda")
    mock.mock_set_tensor_type({torch.float32: 0, torch.float16: 1}[x.dtype])
    try:
        rc = mock.mock_lifecycle(op, nd, x.dim(), arr(tuple(x.shape)), xd.data_ptr(), y.data_ptr(),
                                 torch.cuda.current_stream().cuda_stream, int(skip_configure), out_dims, ctypes.byref(nb))
    finally:
        mock.mock_set_tensor_type(0)
    assert rc == 0
    torch.cuda.synchronize()
    assert tuple(out_dims[:nb.value]) == tuple(y.shape)
    return y.cpu()


@pytest.mark.gpu
@pytest.mark.parametrize("dft_dim1", [1, 2])
@pytest.mark.parametrize("dft_dim2", [4])
@pytest.mark.parametrize("num_c", [1, 3])
@pytest.mark.parametrize("batch_size", [1, 2])
def test_rfft2_like_the_reference(mock, dft_dim1, dft_dim2, num_c, batch_size):
    """tests/test_dft.py:124-151 with the mock host in place of TensorRT."""
    from oracle import oracle as O
    torch.manual_seed(1)
    x = torch.randn(batch_size, num_c, dft_dim1, dft_dim2)
    y_expected = O.torch_rfftn(x, 2)
    y = run_plugin(mock, b"Rfft", 2, x)
    assert torch.allclose(y_expected, y)


@pytest.mark.gpu
@pytest.mark.parametrize("dft_dim1", [1, 2])
@pytest.mark.parametrize("dft_dim2", [4])
@pytest.mark.parametrize("num_c", [1, 3])
@pytest.mark.parametrize("batch_size", [1, 2])
def test_irfft2_like_the_reference(mock, dft_dim1, dft_dim2, num_c, batch_size):
    """tests/test_dft.py:154-184."""
    from oracle import oracle as O
    torch.manual_seed(1)
    x = torch.randn(batch_size, num_c, dft_dim1, dft_dim2)
    y = O.torch_rfftn(x, 2)
    x_expected = O.torch_irfftn(y, 2)
    x_actual = run_plugin(mock, b"Irfft", 2, y)
    assert torch.allclose(x_expected, x_actual)                # the reference's defaults, tests/test_dft.py:184


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nd", [((4, 32, 64), 1), ((2, 8, 90, 180), 2), ((1, 2, 180, 360), 2), ((2, 4, 6, 8), 3)])
@pytest.mark.parametrize("skip_configure", [False, True])
def test_plugin_pair_at_baseline_shapes(mock, shape, nd, skip_configure):
    """Both plugins at BASELINE-like shapes, including a runtime clone that never saw configurePlugin
    (unusable in the reference, dft_plugins.cpp:73-83; fine here because plans come from the descs)."""
    from oracle import oracle as O
    torch.manual_seed(1)
    x = torch.randn(*shape)
    want = O.torch_rfftn(x, nd, dtype=torch.float64)
    y = run_plugin(mock, b"Rfft", nd, x, skip_configure)
    assert O.rel_l2(y.numpy(), want.numpy()) <= 1e-5
    xr = run_plugin(mock, b"Irfft", nd, y, skip_configure)
    assert O.rel_l2(xr.numpy(), x.numpy()) <= 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nd", [((2, 9, 90, 180), 2), ((3, 64, 64), 2), ((4, 32, 64), 1)])
def test_plugin_pair_with_fp16_tensors(mock, shape, nd, monkeypatch):
    """The whole plugin lifecycle with kHALF on both sides (TRT_DFT_PLUGIN_16BIT_IO=1): fp32 arithmetic inside, so what
    separates the result from the fp64 oracle on the same fp16 input is the rounding of the output (tolerance as in
    tests/test_gpu_typed_io.py).  Without the switch the same descriptors are refused."""
    torch.manual_seed(1)
    x = torch.randn(*shape).to(torch.float16)
    dim = tuple(range(-nd, 0))
    monkeypatch.setenv("TRT_DFT_PLUGIN_16BIT_IO", "1")
    y = run_plugin(mock, b"Rfft", nd, x)
    want = torch.view_as_real(torch.fft.rfftn(x.double(), dim=dim))
    assert float((y.double() - want).norm() / want.norm()) <= 3e-4
    xr = run_plugin(mock, b"Irfft", nd, y)
    want_x = torch.fft.irfftn(torch.view_as_complex(y.double().contiguous()), s=shape[-nd:], dim=dim)
    assert float((xr.double() - want_x).norm() / want_x.norm()) <= 3e-4
    monkeypatch.delenv("TRT_DFT_PLUGIN_16BIT_IO")
    with pytest.raises(AssertionError):
        run_plugin(mock, b"Rfft", nd, x)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nd", [((3, 5, 90, 180), 2), ((2, 180, 360), 2), ((6, 64), 1)])
def test_unconfigured_clone_enqueues_inside_a_stream_capture(mock, shape, nd):
    """The execution context's clone never sees configurePlugin (dft_plugins.cpp:73-83 makes such a clone unusable in
    the reference; TensorRT configures it again).  Here the plan built when the ORIGINAL was configured
    (dft_plugins.cpp:131-178 is where the reference builds its cuFFT plan) is shared, so the clone's first enqueue is
    launch-only and can be captured in a CUDA graph; a shape nobody planned is refused inside the capture instead
    of allocating there."""
    from oracle import oracle as O
    torch.manual_seed(1)
    x = torch.randn(*shape)
    want = O.torch_rfftn(x, nd, dtype=torch.float64)
    for op, src, ref, tol in ((b"Rfft", x, want, 1e-5), (b"Irfft", want.float(), x, 1e-5)):
        xd = src.cuda().contiguous()
        y = torch.full(tuple(ref.shape), float("nan"), device="cuda")
        rc = mock.mock_capture_clone(op, nd, xd.dim(), arr(tuple(xd.shape)), xd.data_ptr(), y.data_ptr(), 2)
        assert rc == 0, f"step {rc}"
        torch.cuda.synchronize()
        assert O.rel_l2(y.cpu().numpy(), ref.numpy()) <= tol


def test_dynamic_profile_workspace_bounds_every_smaller_shape(mock):
    """A profile whose largest shape runs fused (no workspace) still has smaller shapes that need one (34 x 68 is
    un-fused); once configured as dynamic, getWorkspaceSize returns a bound that covers them."""
    m = mock
    m.mock_dynamic_workspace.argtypes = [c_char_p, c_int32, c_int32, ctypes.POINTER(c_int32), ctypes.POINTER(c_int32)]
    m.mock_dynamic_workspace.restype = c_size_t
    lo, hi = (1, 2, 3, 2), (4, 128, 65, 2)
    static_need = m.mock_workspace_size(b"Irfft", 2, 4, arr(hi))
    dyn_need = m.mock_dynamic_workspace(b"Irfft", 2, 4, arr(lo), arr(hi))
    assert static_need == 0                                   # fused at the largest shape
    assert dyn_need >= 2 * 4 * 128 * 65 * 8                   # two complex tensors of the largest shape
