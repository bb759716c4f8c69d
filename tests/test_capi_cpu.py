"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol that
include/trt_dft.h declares, and the allocation-free / argument-checking paths behave without a GPU.
No compute call is made here."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    import importlib.util
    spec = importlib.util.spec_from_file_location("_dft_build", ROOT / "tensorrt-dft-plugins_b200" / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build()
    from trt_dft_plugins import load_plugins
    return load_plugins()


def declared_symbols():
    text = (ROOT / "include" / "trt_dft.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(trt_dft_[a-z0-9_]+)\s*\(", text)))


def test_library_is_in_the_package_dir_under_the_reference_name(lib):
    # reference: CMakeLists.txt:26 + src/trt_dft_plugins/__init__.py:28-30
    from trt_dft_plugins import library_path
    assert library_path().name == "libtrt_dft_plugins.so"
    assert library_path().parent.name == "trt_dft_plugins"


def test_every_declared_symbol_is_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/trt_dft.h but not exported"


def test_load_plugins_is_idempotent(lib):
    from trt_dft_plugins import load_plugins
    assert load_plugins() is lib


def test_version_and_error_strings(lib):
    from trt_dft_plugins import capi
    assert "sm_100a" in capi.version()
    L = capi.lib()
    assert L.trt_dft_error_string(0) == b"ok"
    for code in range(1, 6):
        assert L.trt_dft_error_string(code) not in (b"", b"ok")


def test_argument_validation_without_a_device(lib):
    from trt_dft_plugins import capi
    for bad in [dict(direction=2, dims=(8,), batch=1), dict(direction=0, dims=(0,), batch=1),
                dict(direction=0, dims=(2, 2, 2, 2), batch=1), dict(direction=1, dims=(8, 8), batch=-1)]:
        with pytest.raises(capi.DftError):
            capi.Plan(bad["direction"], bad["dims"], bad["batch"])
    # an empty batch needs no device at all
    p = capi.Plan(capi.FORWARD, (90, 180), 0)
    assert p.num_launches == 0 and p.workspace_bytes == 0
    assert p.input_floats == 0 and p.output_floats == 0
    p.exec_ptr(0, 0)


def test_typed_plan_argument_validation_without_a_device(lib):
    """trt_dft_plan_create_typed: the element type is checked before anything touches CUDA, an empty batch needs no device,
    and a 16-bit plan of a real size fails loudly without a GPU like every other plan."""
    import torch
    from trt_dft_plugins import capi
    with pytest.raises(capi.DftError):
        capi.TypedPlan(capi.FORWARD, (8, 8), 1, 3)
    with pytest.raises(capi.DftError):
        capi.TypedPlan(capi.FORWARD, (8, 8), 1, -1)
    with pytest.raises(capi.DftError):
        capi.TypedPlan(2, (8, 8), 1, capi.IO_F16)
    e = capi.TypedPlan(capi.INVERSE, (90, 180), 0, capi.IO_BF16)
    assert e.io_type == capi.IO_BF16 and e.num_launches == 0 and e.workspace_bytes == 0
    e.exec_ptr(0, 0)
    assert capi.TypedPlan(capi.FORWARD, (90, 180), 0, capi.IO_F32).io_type == capi.IO_F32
    if not torch.cuda.is_available():
        with pytest.raises(capi.DftError):
            capi.TypedPlan(capi.FORWARD, (8, 8), 1, capi.IO_F16)


def test_no_cpu_route_without_a_gpu(lib):
    """The product must fail loudly, not fall back, when there is no CUDA device."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from trt_dft_plugins import capi
    with pytest.raises(capi.DftError):
        capi.Plan(capi.FORWARD, (8, 8), 1)


def test_product_sources_never_reference_the_oracle():
    pkg = ROOT / "tensorrt-dft-plugins_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.h")) + list(pkg.rglob("*.cpp")):
        txt = f.read_text()
        assert "oracle" not in txt.lower(), f"{f} mentions the oracle"
