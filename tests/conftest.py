import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def rt():
    import sdb200  # noqa: F401
    from sdb200 import runtime
    return runtime


@pytest.fixture(autouse=True)
def _settle_device_between_tests():
    """Each GPU test starts from a settled device: garbage of the previous test (captured graphs with reference cycles,
    their private memory pools) is collected and pending work has finished before the next test starts capturing."""
    yield
    try:
        import torch
    except ImportError:   # pragma: no cover
        return
    if torch.cuda.is_available():
        import gc
        gc.collect()
        torch.cuda.synchronize()
