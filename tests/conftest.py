import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _cuda_ok():
    try:
        from signals_b200 import _lib
        return _lib.load().sgb_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if any("gpu" in it.keywords for it in items) and not _cuda_ok():
        skip = pytest.mark.skip(reason="no CUDA device visible (GPU tests run on the B200 box)")
        for it in items:
            if "gpu" in it.keywords:
                it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import cpu_oracle
    cpu_oracle.build()
    cpu_oracle.lib()
    return cpu_oracle
