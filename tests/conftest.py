import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_quantisers():
    import torch
    return torch.load(os.path.join(GOLDEN, "quantisers.pt"), weights_only=False)


@pytest.fixture(scope="session")
def golden_layers():
    import torch
    return torch.load(os.path.join(GOLDEN, "layers.pt"), weights_only=False)


@pytest.fixture(scope="session")
def golden_models():
    import torch
    return torch.load(os.path.join(GOLDEN, "models.pt"), weights_only=False)
