import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "slow: takes more than ~20 s (still part of the gpu suite)")


def _has_gpu():
    try:
        from robosoccer_motion_planning_b200 import capi
        return os.path.exists(capi.LIB_PATH) and capi.lib().krrt_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu_required():
    """GPU tests must fail loudly (not skip) when the CUDA library or device is missing."""
    from robosoccer_motion_planning_b200 import capi
    assert os.path.exists(capi.LIB_PATH), "libkrrt.so not built: run __graft_entry__.build()"
    assert capi.lib().krrt_device_count() > 0, "no CUDA device visible"
    return True
