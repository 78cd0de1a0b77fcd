import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device; run with -m gpu on the GPU box")


def _gpu_count():
    try:
        from lambdagap_b200 import capi
        return capi.device_count()
    except Exception:
        return 0


@pytest.fixture(scope="session")
def gpu():
    n = _gpu_count()
    if n < 1:
        pytest.fail("no sm_100 device visible: the -m gpu tests have no CPU fallback")
    return n
