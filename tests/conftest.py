import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture
def backend_kernels(monkeypatch):
    """Real CUDA kernels on a GPU box; the oracle-backed stand-in (tests/fake_kernels.py) elsewhere.

    Only host logic is being tested through the stand-in; kernel parity lives in the
    ``-m gpu`` tests.
    """
    import torch

    if torch.cuda.is_available():
        from thermoextrap_b200 import kernels

        return kernels
    sys.path.insert(0, str(Path(__file__).parent))
    import fake_kernels

    import thermoextrap_b200.backend as backend

    monkeypatch.setattr(backend, "K", fake_kernels)
    backend.clear_device_cache()
    return fake_kernels
