import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        from orb_slam2_map_b200 import capi
        return capi.lib().orbx_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # gpu tests are selected with -m gpu on the GPU box; if someone runs the whole suite on a
    # CPU-only machine they are skipped rather than failed.
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
