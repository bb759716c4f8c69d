"""Real-TensorRT conformance: the reference's own end-to-end tests (export -> parse -> build -> execute -> compare with
CPU torch.fft; /root/reference/tests/test_dft.py:118-184) on the TensorRT 10 / torch 2.x APIs, widened to
signal_ndim 1 and 3 and to the BASELINE shapes.

TensorRT and onnx are absent from the image this repository is built in, so every test here skips itself; the file
is what a maintainer runs the day a TensorRT install exists:
    TENSORRT_INCLUDE_DIR=/usr/include/x86_64-linux-gnu python tensorrt-dft-plugins_b200/build.py --force
    python -m pytest tests/test_trt_conformance.py -m gpu
(the library must be rebuilt against the real NvInfer.h first: a build against the stand-in header does not register
its creators with a real TensorRT).  The ONNX-export half runs wherever `onnx` is importable, GPU or not.
"""
import pytest
import torch

from trt_dft_plugins import onnx_ops


def _oracle_rfft(x, nd):
    return torch.view_as_real(torch.fft.rfftn(x, dim=tuple(range(-nd, 0)), norm="backward"))


@pytest.mark.parametrize("nd,shape", [(1, (2, 3, 8)), (2, (2, 3, 4, 6)), (3, (2, 4, 4, 6))])
def test_eager_ops_match_torch_fft(nd, shape):
    """The eager halves of the export helpers are the oracle expressions themselves (tests/test_dft.py:40, :52-54)."""
    torch.manual_seed(1)
    x = torch.randn(*shape)
    y = onnx_ops.rfft(x, nd)
    assert torch.equal(y, _oracle_rfft(x, nd))
    assert torch.allclose(onnx_ops.irfft(y, nd), x, atol=1e-6)
    with pytest.raises(ValueError):
        onnx_ops.rfft(x, 4)


@pytest.mark.parametrize("nd", [1, 2, 3])
def test_export_emits_the_contrib_ops(nd):
    onnx = pytest.importorskip("onnx")
    x = torch.randn(2, 4, 4, 8)
    m = onnx.load_from_string(onnx_ops.export_onnx(onnx_ops.Rfft(nd), x))
    nodes = [(n.domain, n.op_type, {a.name: a.i for a in n.attribute}) for n in m.graph.node]
    assert ("com.microsoft", "Rfft", {"normalized": 0, "onesided": 1, "signal_ndim": nd}) in nodes
    y = onnx_ops.rfft(x, nd)
    m = onnx.load_from_string(onnx_ops.export_onnx(onnx_ops.Irfft(nd), y))
    nodes = [(n.domain, n.op_type, {a.name: a.i for a in n.attribute}) for n in m.graph.node]
    assert ("com.microsoft", "Irfft", {"normalized": 0, "onesided": 1, "signal_ndim": nd}) in nodes


@pytest.mark.parametrize("nd", [1, 2, 3])
def test_symbolics_without_the_onnx_package(nd):
    """Same check on the exporter's graph before serialisation (works in this image, which has no `onnx`)."""
    try:
        fwd = onnx_ops.traced_graph_text(onnx_ops.Rfft(nd), torch.randn(2, 4, 4, 8))
        inv = onnx_ops.traced_graph_text(onnx_ops.Irfft(nd), torch.randn(2, 4, 4, 5, 2))
    except (ImportError, AttributeError) as e:
        pytest.skip(f"torch moved its exporter internals: {e}")
    assert f"com.microsoft::Rfft[normalized=0, onesided=1, signal_ndim={nd}]" in fwd       # tests/test_dft.py:44-46
    assert f"com.microsoft::Irfft[normalized=0, onesided=1, signal_ndim={nd}]" in inv      # tests/test_dft.py:58-60


@pytest.fixture(scope="module")
def trt_env():
    trt = pytest.importorskip("tensorrt")
    pytest.importorskip("onnx")
    from trt_dft_plugins import load_plugins
    load_plugins()
    return trt


@pytest.mark.gpu
def test_plugins_load(trt_env):
    names = {c.name for c in trt_env.get_plugin_registry().plugin_creator_list}      # tests/test_dft.py:118-121
    assert {"Rfft", "Irfft"} <= names


REF_GRID = [(b, c, h, 4) for h in (1, 2) for c in (1, 3) for b in (1, 2)]            # tests/test_dft.py:124-127
WIDE = [((4, 32, 64), 1), ((2, 8, 128, 128), 2), ((2, 8, 90, 180), 2), ((1, 2, 720, 1440), 2), ((2, 4, 6, 8), 3)]


@pytest.mark.gpu
@pytest.mark.parametrize("shape", REF_GRID)
def test_rfft2_through_tensorrt(trt_env, shape):
    torch.manual_seed(1)
    x = torch.randn(*shape)
    want = onnx_ops.rfft(x, 2)
    plan = onnx_ops.build_engine(onnx_ops.export_onnx(onnx_ops.Rfft(2), x))
    (got,) = onnx_ops.run_engine(plan, [x], [want.shape])
    assert torch.allclose(want, got.cpu())                                          # tests/test_dft.py:151


@pytest.mark.gpu
@pytest.mark.parametrize("shape", REF_GRID)
def test_irfft2_through_tensorrt(trt_env, shape):
    torch.manual_seed(1)
    y = onnx_ops.rfft(torch.randn(*shape), 2)
    want = onnx_ops.irfft(y, 2)
    plan = onnx_ops.build_engine(onnx_ops.export_onnx(onnx_ops.Irfft(2), y))
    (got,) = onnx_ops.run_engine(plan, [y], [want.shape])
    assert torch.allclose(want, got.cpu())                                          # tests/test_dft.py:184


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nd", WIDE)
def test_wider_shapes_through_tensorrt(trt_env, shape, nd):
    torch.manual_seed(1)
    x = torch.randn(*shape)
    want = _oracle_rfft(x.double(), nd)
    plan = onnx_ops.build_engine(onnx_ops.export_onnx(onnx_ops.Rfft(nd), x))
    (y,) = onnx_ops.run_engine(plan, [x], [want.shape])
    assert ((y.cpu().double() - want).norm() / want.norm()).item() <= 1e-5
    plan = onnx_ops.build_engine(onnx_ops.export_onnx(onnx_ops.Irfft(nd), y.cpu()))
    (xr,) = onnx_ops.run_engine(plan, [y], [x.shape])
    assert ((xr.cpu() - x).double().norm() / x.double().norm()).item() <= 1e-6


@pytest.mark.gpu
def test_engine_survives_serialisation_and_a_second_context(trt_env):
    """deserializePlugin + clone per execution context (dft_plugins.cpp:545-556, :73-83)."""
    torch.manual_seed(1)
    x = torch.randn(2, 3, 90, 180)
    want = onnx_ops.rfft(x, 2)
    plan = onnx_ops.build_engine(onnx_ops.export_onnx(onnx_ops.Rfft(2), x))
    for _ in range(2):
        (got,) = onnx_ops.run_engine(plan, [x], [want.shape])
        assert ((got.cpu().double() - want.double()).norm() / want.double().norm()).item() <= 1e-5
