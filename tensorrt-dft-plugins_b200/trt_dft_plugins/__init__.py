"""trt_dft_plugins -- loader for the B200-native Rfft / Irfft TensorRT plugin library.

Same module name, function name and behaviour as the reference loader
(/root/reference/src/trt_dft_plugins/__init__.py:26-32): ``load_plugins()`` dlopens
``libtrt_dft_plugins.so`` from this package directory with RTLD_GLOBAL; the library's static
initialisers register the "Rfft" and "Irfft" plugin creators with TensorRT's global registry when
libnvinfer is present in the process.

Beyond the reference, the library exports the C ABI of include/trt_dft.h; ``trt_dft_plugins.capi``
binds it with ctypes for TensorRT-less use and for the parity tests.
"""
import ctypes
from pathlib import Path

_LIB_NAME = "libtrt_dft_plugins.so"
_lib = None


def library_path() -> Path:
    return Path(__file__).parent / _LIB_NAME


def load_plugins():
    """Loads the plugin library (idempotent) and returns the ctypes handle.

    Raises OSError when the library has not been built -- there is no Python or CPU fallback.
    """
    global _lib
    if _lib is None:
        path = library_path()
        if not path.exists():
            raise OSError(
                f"{path} is missing: build it with `python tensorrt-dft-plugins_b200/build.py` "
                "(nvcc, sm_100a). There is no CPU fallback.")
        _lib = ctypes.CDLL(str(path), mode=ctypes.RTLD_GLOBAL)
    assert _lib is not None
    return _lib
