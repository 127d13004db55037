import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


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
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _fresh_tie_break_default():
    """tree_module.set_tie_break(seed) pins the seed every Roots object created afterwards starts with - a process-wide
    default by design (the module-level API has no driver object to carry it).  Tests that pin it must not leak the pin
    into the tests that run after them in the same process."""
    yield
    try:
        from lightzero_b200.ctree import _tree_api
        _tree_api.set_tie_break(None)
    except Exception:
        pass
