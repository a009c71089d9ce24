import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_PARENT = os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")
for p in (ROOT, PKG_PARENT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs at least 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        ndev = torch.cuda.device_count()
        skip_multi = pytest.mark.skip(reason="needs >= 2 GPUs")
        for item in items:
            if "multigpu" in item.keywords and ndev < 2:
                item.add_marker(skip_multi)
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """The C-ABI library must exist for every test session (host-only entry points run on CPU too)."""
    from b200ddp import _build
    _build.build()
    yield
