"""ctypes binding of include/trt_dft.h plus thin torch-tensor helpers.

torch is used only for device memory and streams; all arithmetic happens in the CUDA kernels of
libtrt_dft_plugins.so.  Every call raises on a non-zero status -- nothing here computes on the CPU.
"""
import ctypes
from ctypes import c_char_p, c_int, c_int64, c_size_t, c_void_p

from . import load_plugins

FORWARD, INVERSE = 0, 1

SYMBOLS = [
    "trt_dft_version", "trt_dft_error_string", "trt_dft_plan_create", "trt_dft_workspace_bytes",
    "trt_dft_exec", "trt_dft_plan_destroy", "trt_dft_exec_host", "trt_dft_plan_describe",
    "trt_dft_plan_input_floats", "trt_dft_plan_output_floats", "trt_dft_plan_create_c2c", "trt_dft_plan_set_column_split",
    "trt_dft_query_workspace_bytes", "trt_dft_exec_static_ws", "trt_dft_plan_create_truncated",
    "trt_dft_plan_set_peer_column_split", "trt_dft_plan_set_peer_row_split", "trt_dft_plan_set_peer_column_gather",
    "trt_dft_plan_create_typed", "trt_dft_plan_io_type",
]

_bound = None


class DftError(RuntimeError):
    pass


def lib():
    global _bound
    if _bound is None:
        L = load_plugins()
        L.trt_dft_version.restype = c_char_p
        L.trt_dft_version.argtypes = []
        L.trt_dft_error_string.restype = c_char_p
        L.trt_dft_error_string.argtypes = [c_int]
        L.trt_dft_plan_create.restype = c_int
        L.trt_dft_plan_create.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int, ctypes.POINTER(c_int64), c_int64]
        L.trt_dft_workspace_bytes.restype = c_size_t
        L.trt_dft_workspace_bytes.argtypes = [c_void_p]
        L.trt_dft_exec.restype = c_int
        L.trt_dft_exec.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
        L.trt_dft_plan_destroy.restype = c_int
        L.trt_dft_plan_destroy.argtypes = [c_void_p]
        L.trt_dft_exec_host.restype = c_int
        L.trt_dft_exec_host.argtypes = [c_void_p, c_void_p, c_void_p]
        L.trt_dft_plan_describe.restype = c_int
        L.trt_dft_plan_describe.argtypes = [c_void_p, c_char_p, c_size_t]
        L.trt_dft_plan_input_floats.restype = c_int64
        L.trt_dft_plan_input_floats.argtypes = [c_void_p]
        L.trt_dft_plan_output_floats.restype = c_int64
        L.trt_dft_plan_output_floats.argtypes = [c_void_p]
        _bound = L
    return _bound


def _check(status, what):
    if status != 0:
        raise DftError(f"{what}: {lib().trt_dft_error_string(status).decode()} (status {status})")


def version() -> str:
    return lib().trt_dft_version().decode()


class Plan:
    """One (direction, signal dims, batch) plan; the analogue of the cuFFT handle the reference
    creates in configurePlugin (dft_plugins.cpp:161-176)."""

    def __init__(self, direction: int, dims, batch: int):
        self.direction = int(direction)
        self.dims = tuple(int(d) for d in dims)
        self.batch = int(batch)
        self._h = c_void_p()
        arr = (c_int64 * len(self.dims))(*self.dims)
        _check(lib().trt_dft_plan_create(ctypes.byref(self._h), self.direction, len(self.dims), arr, self.batch),
               "trt_dft_plan_create")
        self._ws = None

    @property
    def handle(self):
        return self._h

    @property
    def workspace_bytes(self) -> int:
        return int(lib().trt_dft_workspace_bytes(self._h))

    @property
    def input_floats(self) -> int:
        return int(lib().trt_dft_plan_input_floats(self._h))

    @property
    def output_floats(self) -> int:
        return int(lib().trt_dft_plan_output_floats(self._h))

    def describe(self):
        buf = ctypes.create_string_buffer(1024)
        n = lib().trt_dft_plan_describe(self._h, buf, 1024)
        return n, buf.value.decode()

    @property
    def num_launches(self) -> int:
        return self.describe()[0]

    def exec_ptr(self, in_ptr: int, out_ptr: int, ws_ptr: int = 0, stream: int = 0):
        _check(lib().trt_dft_exec(self._h, c_void_p(in_ptr), c_void_p(out_ptr), c_void_p(ws_ptr or None),
                                  c_void_p(stream or None)), "trt_dft_exec")

    def exec(self, x, out, workspace=None, stream=None):
        """x, out: contiguous CUDA float32 tensors; runs on `stream` (default: torch's current stream)."""
        import torch
        want = getattr(self, "torch_dtype", None) or torch.float32
        assert x.is_cuda and out.is_cuda and x.dtype == want and out.dtype == want
        assert x.is_contiguous() and out.is_contiguous()
        assert x.numel() == self.input_floats and out.numel() == self.output_floats
        need = self.workspace_bytes
        if need and workspace is None:
            if self._ws is None or self._ws.numel() < need or self._ws.device != x.device:
                self._ws = torch.empty(need, dtype=torch.uint8, device=x.device)
            workspace = self._ws
        st = torch.cuda.current_stream(x.device).cuda_stream if stream is None else int(stream)
        self.exec_ptr(x.data_ptr(), out.data_ptr(), workspace.data_ptr() if workspace is not None else 0, st)
        return out

    def exec_host(self, x_host, out_host):
        """x_host, out_host: contiguous CPU float32 tensors or numpy arrays (pinned or pageable)."""
        def ptr(t):
            return t.data_ptr() if hasattr(t, "data_ptr") else t.ctypes.data
        _check(lib().trt_dft_exec_host(self._h, c_void_p(ptr(x_host)), c_void_p(ptr(out_host))), "trt_dft_exec_host")
        return out_host

    def destroy(self):
        if self._h:
            lib().trt_dft_plan_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def _split(shape, nd):
    lead = 1
    for s in shape[:-nd]:
        lead *= int(s)
    return lead, tuple(int(s) for s in shape[-nd:])


def rfftn(x, signal_ndim: int, out=None):
    """Rfft plugin semantics (dft_plugins.cpp:361-382): [..., d1..dk] -> [..., d1.., dk//2+1, 2]."""
    import torch
    batch, dims = _split(x.shape, signal_ndim)
    if out is None:
        out = torch.empty(tuple(x.shape[:-1]) + (dims[-1] // 2 + 1, 2), dtype=torch.float32, device=x.device)
    plan = Plan(FORWARD, dims, batch)
    try:
        plan.exec(x.contiguous(), out)
    finally:
        torch.cuda.current_stream(x.device).synchronize()
        plan.destroy()
    return out


def irfftn(y, signal_ndim: int, out=None, last=None):
    """Irfft plugin semantics (dft_plugins.cpp:415-436): [..., d1.., m, 2] -> [..., d1.., (m-1)*2].
    `last` overrides the real length (odd lengths; reachable through the C ABI only)."""
    import torch
    assert y.shape[-1] == 2
    m = int(y.shape[-2])
    last = (m - 1) * 2 if last is None else int(last)
    assert last // 2 + 1 == m
    rshape = tuple(y.shape[:-2]) + (last,)
    batch, dims = _split(rshape, signal_ndim)
    if out is None:
        out = torch.empty(rshape, dtype=torch.float32, device=y.device)
    plan = Plan(INVERSE, dims, batch)
    try:
        plan.exec(y.contiguous(), out)
    finally:
        torch.cuda.current_stream(y.device).synchronize()
        plan.destroy()
    return out


class C2CPlan(Plan):
    """Complex FFT of length n along the middle axis of data[outer][n][inner] (trt_dft_plan_create_c2c).
    Used by the pencil-decomposed multi-GPU rfft2 for the column pass after the all-to-all."""

    def __init__(self, direction: int, n: int, outer: int, inner: int, scale: float = 1.0):
        L = lib()
        L.trt_dft_plan_create_c2c.restype = c_int
        L.trt_dft_plan_create_c2c.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int64, c_int64, c_int64, ctypes.c_float]
        self.direction = int(direction)
        self.dims = (int(n),)
        self.batch = int(outer) * int(inner)
        self._h = c_void_p()
        self._ws = None
        _check(L.trt_dft_plan_create_c2c(ctypes.byref(self._h), self.direction, int(n), int(outer), int(inner),
                                         float(scale)), "trt_dft_plan_create_c2c")


class TruncatedPlan(Plan):
    """rfft2 / irfft2 that moves only the kept low modes (trt_dft_plan_create_truncated): rows [0, keep_h) and
    [H - keep_h, H), columns [0, keep_w) of the one-sided spectrum, as a dense [batch, 2*keep_h, keep_w, 2] tensor."""

    def __init__(self, direction: int, dims, batch: int, keep_h: int, keep_w: int):
        L = lib()
        L.trt_dft_plan_create_truncated.restype = c_int
        L.trt_dft_plan_create_truncated.argtypes = [ctypes.POINTER(c_void_p), c_int, ctypes.POINTER(c_int64), c_int64, c_int64, c_int64]
        self.direction = int(direction)
        self.dims = tuple(int(d) for d in dims)
        assert len(self.dims) == 2
        self.batch = int(batch)
        self.keep_h, self.keep_w = int(keep_h), int(keep_w)
        self._h = c_void_p()
        self._ws = None
        arr = (c_int64 * 2)(*self.dims)
        _check(L.trt_dft_plan_create_truncated(ctypes.byref(self._h), self.direction, arr, self.batch, self.keep_h, self.keep_w),
               "trt_dft_plan_create_truncated")


IO_F32, IO_F16, IO_BF16 = 0, 1, 2


class TypedPlan(Plan):
    """rfftN / irfftN over 16-bit tensors with fp32 arithmetic (trt_dft_plan_create_typed).  `io` is IO_F16 or IO_BF16
    (or a torch dtype); both the real and the complex tensor hold that type."""

    def __init__(self, direction: int, dims, batch: int, io):
        import torch
        io = {torch.float32: IO_F32, torch.float16: IO_F16, torch.bfloat16: IO_BF16}.get(io, io)
        L = lib()
        L.trt_dft_plan_create_typed.restype = c_int
        L.trt_dft_plan_create_typed.argtypes = [ctypes.POINTER(c_void_p), c_int, c_int, ctypes.POINTER(c_int64), c_int64, c_int]
        L.trt_dft_plan_io_type.restype = c_int
        L.trt_dft_plan_io_type.argtypes = [c_void_p]
        self.direction = int(direction)
        self.dims = tuple(int(d) for d in dims)
        self.batch = int(batch)
        self.io = int(io)
        self.torch_dtype = {IO_F32: torch.float32, IO_F16: torch.float16, IO_BF16: torch.bfloat16}.get(self.io)
        self._h = c_void_p()
        self._ws = None
        arr = (c_int64 * len(self.dims))(*self.dims)
        _check(L.trt_dft_plan_create_typed(ctypes.byref(self._h), self.direction, len(self.dims), arr, self.batch, self.io),
               "trt_dft_plan_create_typed")

    @property
    def io_type(self) -> int:
        return int(lib().trt_dft_plan_io_type(self._h))


def kept_rows(H: int, keep_h: int):
    """Row indices of the kept modes, in the order the truncated plans store them."""
    return list(range(keep_h)) + list(range(H - keep_h, H))
