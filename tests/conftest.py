import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
importlib.import_module("3dgs_hierarchical_training_b200").install()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


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


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The CUDA library and the oracle must exist before any test runs (build() is idempotent)."""
    import __graft_entry__ as g
    lib = os.path.join(ROOT, "3dgs_hierarchical_training_b200", "libhtgs.so")
    orc = os.path.join(ROOT, "oracle", "liborc.so")
    stale = importlib.import_module("3dgs_hierarchical_training_b200.build").needs_build()
    if stale or not (os.path.exists(lib) and os.path.exists(orc)):
        g.build()
    yield
