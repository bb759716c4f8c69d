import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
# the package directory name has a '-' in it, so it goes on sys.path instead of being imported
sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
