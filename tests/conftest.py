import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def engine():
    """The product engine bound to cuda:0; building the library if needed."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from desi_mcmc_b200 import build, engine as eng
    build.build()
    eng.init(0)
    return eng


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (test infrastructure)."""
    from oracle import ref_cpu
    return ref_cpu
