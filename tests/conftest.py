import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def cuda_dev():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("-m gpu tests need a CUDA device")
    return torch.device("cuda:0")


@pytest.fixture
def cta_group():
    """Select the single-CTA (1) or CTA-pair (2) tcgen05 kernels for one test; the default (2) is restored after."""
    from kubernetes_cloud_b200 import _lib
    yield lambda n: _lib.set_option("cta_group", n)
    _lib.set_option("cta_group", 2)
