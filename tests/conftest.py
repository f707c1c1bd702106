import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pytorch_r2d2_b200 import _lib
    _lib.load()  # fails loudly if the extension is missing on a GPU box
    return torch.device("cuda:0")


@pytest.fixture(scope="session", autouse=True)
def _oracle_warm_up():
    """See oracle.warm_up: the first torch-CPU forward of a process is occasionally 3.5e-5 off."""
    import oracle
    oracle.warm_up()
