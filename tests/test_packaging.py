"""Packaging parity with the reference (pyproject.toml:26-56, setup.py:30-48, CMakeLists.txt:26): `build_ext` builds the
library and leaves it inside the trt_dft_plugins package directory; an installed copy of the package carries it and
load_plugins() finds it there."""
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "tensorrt-dft-plugins_b200"


def test_build_ext_inplace_leaves_the_library_next_to_the_loader():
    r = subprocess.run([sys.executable, "setup.py", "-q", "build_ext", "--inplace"], cwd=PKG, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    assert (PKG / "trt_dft_plugins" / "libtrt_dft_plugins.so").exists()


def test_installed_package_carries_the_library(tmp_path):
    target = tmp_path / "site"
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "-q",
                        "--target", str(target), str(PKG)], capture_output=True, text=True)
    for scratch in (PKG / "build", PKG / "trt_dft_plugins_b200.egg-info"):      # pip builds in the source tree
        shutil.rmtree(scratch, ignore_errors=True)
    assert r.returncode == 0, r.stderr[-2000:]
    assert (target / "trt_dft_plugins" / "libtrt_dft_plugins.so").exists()
    code = ("import sys; sys.path.insert(0, r'%s'); import trt_dft_plugins, os; "
            "assert os.path.dirname(trt_dft_plugins.__file__).startswith(r'%s'); "
            "from trt_dft_plugins import capi; print(capi.version())") % (target, target)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0 and "b200-dft" in r.stdout, r.stderr[-2000:]
