import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_SCRIPT_CANDIDATES = [
    "/root/reference/Vaihingen PyTorch 2 (кластер).py",
    os.path.join(ROOT, "baseline", "_ref", "Vaihingen PyTorch 2 (кластер).py"),
]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 through gpurun)")
    config.addinivalue_line("markers", "multigpu: needs more than one CUDA device")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def reference_script():
    for p in REFERENCE_SCRIPT_CANDIDATES:
        if os.path.exists(p):
            return p
    pytest.skip("reference script not available")
