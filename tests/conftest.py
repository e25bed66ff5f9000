import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def native():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pytorch_distributed_examples_b200 import _native
    _native.set_instance(None)
    nat = _native.get()          # raises if libddp_step.so is missing: no silent fallback
    nat.bind_device(0)
    return nat
