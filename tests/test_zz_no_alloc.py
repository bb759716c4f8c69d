"""trt_dft_exec_static_ws: the entry point for callers that cannot allocate a workspace.

A plan for a large inverse transform needs scratch memory between its column pass and its row
pass.  Normally the caller allocates that buffer and passes it to trt_dft_exec.  Code that runs
where allocation is off limits (a plugin's enqueue, a serving loop that cannot afford cudaMalloc's
device synchronisation, an op interface with no workspace argument) calls
trt_dft_exec_static_ws instead, which lends the plan a fixed region the library holds on the
GPU (from this entry point's first call there).  This file checks, through the C ABI, that the result is the same as with a
caller-owned workspace -- twice in a row, with different inputs, because the region is shared and
keeps the previous call's contents -- and that a plan too large for the region is refused.

Named test_zz_* so pytest collects it after every other test file.
"""
import ctypes
from ctypes import c_int, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5          # the parity tolerance used throughout tests/test_gpu_parity.py
STATIC_WS_BYTES = 32 * 1024 * 1024


@pytest.fixture(scope="module")
def api():
    from trt_dft_plugins import capi
    lib = capi.lib()
    fn = lib.trt_dft_exec_static_ws
    fn.restype = c_int
    fn.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p]
    return capi, fn


def rel_l2(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


def test_static_workspace_matches_caller_workspace_twice(api):
    capi, exec_static = api
    batch, H, W = 3, 180, 360                       # too big to fuse: column pass -> workspace -> row pass
    plan = capi.Plan(capi.INVERSE, (H, W), batch)
    need = plan.workspace_bytes
    assert 0 < need <= STATIC_WS_BYTES
    stream = torch.cuda.current_stream().cuda_stream
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    for seed in (1, 2):                             # two calls in a row, different data
        torch.manual_seed(seed)
        x = torch.randn(batch, H, W // 2 + 1, 2, device="cuda")
        want = torch.full((batch, H, W), float("nan"), device="cuda")
        got = torch.full((batch, H, W), float("nan"), device="cuda")
        plan.exec_ptr(x.data_ptr(), want.data_ptr(), ws.data_ptr(), stream)           # existing entry point
        rc = exec_static(plan.handle, c_void_p(x.data_ptr()), c_void_p(got.data_ptr()), c_void_p(stream))
        assert rc == 0
        torch.cuda.synchronize()
        assert not torch.isnan(got).any()
        assert rel_l2(got, want) <= TOL
        assert torch.equal(got, want)               # same kernels, same order: bit-identical in fact
        # and both agree with the fp64 oracle expression (tests/test_dft.py:52-54)
        ref = torch.fft.irfft2(torch.view_as_complex(x.double().cpu()), dim=(-2, -1), norm="backward")
        assert rel_l2(got.cpu(), ref) <= TOL


def test_plans_without_workspace_run_unchanged(api):
    capi, exec_static = api
    plan = capi.Plan(capi.FORWARD, (90, 180), 4)
    assert plan.workspace_bytes == 0
    torch.manual_seed(1)
    x = torch.randn(4, 90, 180, device="cuda")
    want = torch.empty(4, 90, 91, 2, device="cuda")
    got = torch.empty(4, 90, 91, 2, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    plan.exec_ptr(x.data_ptr(), want.data_ptr(), 0, stream)
    assert exec_static(plan.handle, c_void_p(x.data_ptr()), c_void_p(got.data_ptr()), c_void_p(stream)) == 0
    torch.cuda.synchronize()
    assert torch.equal(got, want)


def test_oversized_plan_is_refused(api):
    capi, exec_static = api
    plan = capi.Plan(capi.INVERSE, (720, 1440), 16)          # 16 * 720 * 721 * 8 B = 66 MB > 32 MiB
    assert plan.workspace_bytes > STATIC_WS_BYTES
    x = torch.zeros(16, 720, 721, 2, device="cuda")
    out = torch.zeros(16, 720, 1440, device="cuda")
    rc = exec_static(plan.handle, c_void_p(x.data_ptr()), c_void_p(out.data_ptr()), c_void_p(0))
    assert rc == 5                                            # TRT_DFT_ERR_UNSUPPORTED
    assert exec_static(None, c_void_p(x.data_ptr()), c_void_p(out.data_ptr()), c_void_p(0)) == 1
