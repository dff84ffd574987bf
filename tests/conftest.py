import os
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle_abi():
    from oracle import openblas_ref
    return openblas_ref.abi()


@pytest.fixture(scope="session")
def product_abi():
    """libla_b200.so bound through the reference's ABI.  Fails (never skips) when missing."""
    import linalg_b200
    return linalg_b200.abi()


@pytest.fixture
def rng():
    return np.random.default_rng(20261106)
