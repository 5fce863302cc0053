import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # A fresh checkout has no built library (it is git-ignored): build it once, like __graft_entry__.build() does.
    # nvcc cross-compiles for sm_100a without a GPU; the CPU tests only load it (exports, host logic), never compute.
    try:
        from sdrx_b200 import _lib
        if not os.path.exists(_lib.LIB_PATH):
            from sdrx_b200 import build
            build.build()
    except Exception as exc:       # the tests that need the library will say so themselves
        sys.stderr.write(f"conftest: could not build libsdrx_b200.so: {exc}\n")


def _have_gpu() -> bool:
    try:
        from sdrx_b200 import _lib
        return _lib.load().sdrx_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
