import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# the staged copy of the reference's callers (see __graft_entry__.stage_reference_callers) is test
# INPUT: it is imported by tests/test_gpu_ref_callers.py, never collected
collect_ignore = ["_ref_callers"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def ops():
    """The product: the in-tree torch_sputnik extension (fails loudly if not built)."""
    import torch_sputnik_b200
    return torch_sputnik_b200.torch_sputnik


@pytest.fixture(scope="session")
def dev():
    import torch
    return torch.device("cuda:0")
