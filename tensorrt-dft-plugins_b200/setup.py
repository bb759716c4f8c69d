"""pip / setuptools entry for the B200-native trt_dft_plugins package.

    pip install --no-build-isolation -e tensorrt-dft-plugins_b200      # or: python setup.py build_ext --inplace

`build_ext` compiles csrc/ with nvcc for sm_100a through build.py and leaves libtrt_dft_plugins.so inside the
trt_dft_plugins package directory, next to __init__.py -- where load_plugins() looks for it and where the reference's
build puts its library too (/root/reference/CMakeLists.txt:26, setup.py:30-42; the reference drives cmake, this drives
nvcc directly: there is no CMake project here).  Set TENSORRT_INCLUDE_DIR to compile the plugin shell against a real
NvInfer.h.
"""
import importlib.util
import shutil
from pathlib import Path

from setuptools import Extension, setup
from setuptools.command.build_ext import build_ext

HERE = Path(__file__).resolve().parent


def _build_module():
    spec = importlib.util.spec_from_file_location("_trt_dft_build", HERE / "build.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class NvccBuild(build_ext):
    """One pseudo-extension: the shared library is built by build.py, not by distutils' compiler."""

    def run(self):
        so = _build_module().build(force=False)
        if not self.inplace:
            # wheel / non-editable install: the library must travel inside the package
            dest = Path(self.build_lib) / "trt_dft_plugins" / so.name
            dest.parent.mkdir(parents=True, exist_ok=True)
            shutil.copy2(so, dest)

    def get_outputs(self):
        return [str(Path(self.build_lib) / "trt_dft_plugins" / "libtrt_dft_plugins.so")]


setup(
    ext_modules=[Extension(name="trt_dft_plugins.libtrt_dft_plugins", sources=[])],
    cmdclass={"build_ext": NvccBuild},
)
