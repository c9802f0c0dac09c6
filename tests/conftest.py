import os
import sys

import pytest

# test switch of the library (b2fwn_pipeline.cu): 16-KiB staging pieces, so that even the small test grids travel through the
# pinned ring in several pieces per chunk; bench.py runs with the default 32-MiB pieces
os.environ.setdefault("B2FWN_PIECE_FLOATS", "4096")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import meshlib_b200
        return meshlib_b200.load_library().b2fwn_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
