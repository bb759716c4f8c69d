"""ONNX export helpers for the com.microsoft Rfft / Irfft contrib ops that the plugins implement, for
signal_ndim 1, 2 and 3, plus TensorRT build / run helpers written against the TensorRT 10 Python API
(falling back to the 8.x calls where they still exist).

The reference keeps equivalents of these inside its test file, for signal_ndim = 2 only and on the TensorRT 8 /
torch 1.13 APIs (/root/reference/tests/test_dft.py:37-60 the autograd Functions with their ONNX symbolics, :73-86
export, :89-101 engine build, :104-115 execution).  They are library code here so that a user can export a model that
contains the ops without copying test helpers around.

Nothing in this module is on the hot path, and importing it needs neither `onnx` nor `tensorrt`: those are imported
inside the functions that use them.  TensorRT is absent from the image this repository is developed in, so the
TensorRT functions are exercised only by tests/test_trt_conformance.py, which skips itself without TensorRT.
"""
import io
from typing import Sequence

import torch
from torch.autograd import Function

ONNX_DOMAIN = "com.microsoft"
OPSET = 17


def _signal_dims(nd: int):
    if nd not in (1, 2, 3):
        raise ValueError("signal_ndim must be 1, 2 or 3 (the contrib ops define nothing else)")
    return tuple(range(-nd, 0))


class _Rfft(Function):
    """Eager: torch.fft.rfftn over the trailing signal_ndim dims, complex as a trailing dim of 2 ("backward" norm, i.e.
    unnormalised).  ONNX: one com.microsoft::Rfft node with the attribute values the plugin creator accepts."""

    @staticmethod
    def forward(ctx, x, signal_ndim):
        return torch.view_as_real(torch.fft.rfftn(x, dim=_signal_dims(signal_ndim), norm="backward"))

    @staticmethod
    def symbolic(g, x, signal_ndim):
        return g.op(f"{ONNX_DOMAIN}::Rfft", x, normalized_i=0, onesided_i=1, signal_ndim_i=int(signal_ndim))


class _Irfft(Function):
    """Eager: torch.fft.irfftn (scaled by 1/prod(signal dims)); the real length of the last dim is (m - 1) * 2, the
    plugin's shape rule (dft_plugins.cpp:415-436).  ONNX: one com.microsoft::Irfft node."""

    @staticmethod
    def forward(ctx, y, signal_ndim):
        return torch.fft.irfftn(torch.view_as_complex(y.contiguous()), dim=_signal_dims(signal_ndim), norm="backward")

    @staticmethod
    def symbolic(g, y, signal_ndim):
        return g.op(f"{ONNX_DOMAIN}::Irfft", y, normalized_i=0, onesided_i=1, signal_ndim_i=int(signal_ndim))


def rfft(x: torch.Tensor, signal_ndim: int = 2) -> torch.Tensor:
    """[..., d1..dk] real -> [..., d1.., dk//2+1, 2].  Traces to com.microsoft::Rfft on ONNX export."""
    return _Rfft.apply(x, signal_ndim)


def irfft(y: torch.Tensor, signal_ndim: int = 2) -> torch.Tensor:
    """[..., d1.., m, 2] -> [..., d1.., (m-1)*2] real.  Traces to com.microsoft::Irfft on ONNX export."""
    return _Irfft.apply(y, signal_ndim)


class Rfft(torch.nn.Module):
    def __init__(self, signal_ndim: int = 2):
        super().__init__()
        self.signal_ndim = signal_ndim

    def forward(self, x):
        return rfft(x, self.signal_ndim)


class Irfft(torch.nn.Module):
    def __init__(self, signal_ndim: int = 2):
        super().__init__()
        self.signal_ndim = signal_ndim

    def forward(self, y):
        return irfft(y, self.signal_ndim)


def export_onnx(model: torch.nn.Module, example: torch.Tensor, verbose: bool = False) -> bytes:
    """TorchScript-based exporter (the path that honours Function.symbolic); torch >= 2.5 needs dynamo=False for it."""
    buf = io.BytesIO()
    kwargs = dict(opset_version=OPSET, verbose=verbose, custom_opsets={ONNX_DOMAIN: 1})
    try:
        torch.onnx.export(model, (example,), buf, dynamo=False, **kwargs)
    except TypeError:                      # older torch: no `dynamo` keyword
        torch.onnx.export(model, (example,), buf, **kwargs)
    return buf.getvalue()


def traced_graph_text(model: torch.nn.Module, example: torch.Tensor) -> str:
    """The ONNX graph the exporter would serialise, as text, without needing the `onnx` package (which the exporter
    only needs for its very last step).  Uses torch-internal entry points, so callers should be ready for ImportError /
    AttributeError on torch versions that moved them.  For tests and for eyeballing what a model exports to."""
    from torch.onnx._internal.torchscript_exporter import utils as U
    from torch.onnx._internal.torchscript_exporter._globals import GLOBALS
    GLOBALS.export_onnx_opset_version = OPSET
    with U.exporter_context(model, torch.onnx.TrainingMode.EVAL, False):
        graph, _, _ = U._model_to_graph(model, (example,))
    return str(graph)


def build_engine(onnx_bytes: bytes, logger=None, workspace_bytes: int = 0) -> bytes:
    """ONNX -> serialized TensorRT engine.  load_plugins() must have run so that the parser finds the creators."""
    import tensorrt as trt
    logger = logger or trt.Logger(trt.Logger.WARNING)
    builder = trt.Builder(logger)
    flags = 0
    if hasattr(trt.NetworkDefinitionCreationFlag, "EXPLICIT_BATCH") and int(trt.__version__.split(".")[0]) < 10:
        flags = 1 << int(trt.NetworkDefinitionCreationFlag.EXPLICIT_BATCH)      # implicit in TensorRT 10
    network = builder.create_network(flags)
    parser = trt.OnnxParser(network, logger)
    if not parser.parse(onnx_bytes):
        raise RuntimeError("ONNX parse failed:\n" + "\n".join(str(parser.get_error(i)) for i in range(parser.num_errors)))
    config = builder.create_builder_config()
    if workspace_bytes:
        config.set_memory_pool_limit(trt.MemoryPoolType.WORKSPACE, workspace_bytes)
    plan = builder.build_serialized_network(network, config)
    if plan is None:
        raise RuntimeError("TensorRT could not build the engine")
    return bytes(plan)


def run_engine(engine_bytes: bytes, inputs: Sequence[torch.Tensor], output_shapes: Sequence[Sequence[int]], logger=None):
    """Deserialize and run once on the current CUDA stream; returns the outputs as CUDA tensors."""
    import tensorrt as trt
    logger = logger or trt.Logger(trt.Logger.WARNING)
    engine = trt.Runtime(logger).deserialize_cuda_engine(engine_bytes)
    context = engine.create_execution_context()
    ins = [t.cuda().contiguous() for t in inputs]
    outs = [torch.empty(tuple(s), dtype=torch.float32, device="cuda") for s in output_shapes]
    stream = torch.cuda.current_stream()
    if hasattr(context, "execute_async_v3"):                     # TensorRT >= 8.5 name-based I/O
        names = [engine.get_tensor_name(i) for i in range(engine.num_io_tensors)]
        in_names = [n for n in names if engine.get_tensor_mode(n) == trt.TensorIOMode.INPUT]
        out_names = [n for n in names if engine.get_tensor_mode(n) == trt.TensorIOMode.OUTPUT]
        for n, t in zip(in_names, ins):
            context.set_input_shape(n, tuple(t.shape))
            context.set_tensor_address(n, t.data_ptr())
        for n, t in zip(out_names, outs):
            context.set_tensor_address(n, t.data_ptr())
        if not context.execute_async_v3(stream.cuda_stream):
            raise RuntimeError("execute_async_v3 failed")
    else:
        if not context.execute_v2([t.data_ptr() for t in ins + outs]):
            raise RuntimeError("execute_v2 failed")
    stream.synchronize()
    return outs
