import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def engine():
    import dnn_b200
    return dnn_b200


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def cuda_engine(engine):
    """The engine with its CUDA library loaded on a real device; GPU tests must not
    silently pass without native code."""
    import torch
    assert torch.cuda.is_available(), "GPU test collected on a box without CUDA"
    assert os.path.exists(engine._lib.LIB_PATH), "libb200pipe.so missing: run __graft_entry__.build()"
    engine._lib.lib()
    return engine
