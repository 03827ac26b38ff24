import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def _have_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # GPU tests must never silently pass on a CPU box: without -m gpu they are deselected by the
    # driver; if someone runs them anyway without a device they fail inside bwk_init (no fallback).
    pass


@pytest.fixture(scope="session")
def oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py

    oracle_py.build()
    return oracle_py


@pytest.fixture(scope="session")
def bwk():
    import __graft_entry__ as ge

    ge.build(oracle=False)
    import codesign_kernels_b200 as m

    return m
