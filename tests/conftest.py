import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box: pytest -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """CPU restatement of the reference (oracle/), test infrastructure only."""
    from tests.oracle_lib import load_oracle
    return load_oracle()


@pytest.fixture(scope="session")
def emu():
    """Host build of the product's shape compiler + program evaluator (tests/emu), to check them without a GPU."""
    from tests.oracle_lib import load_emu
    return load_emu()


@pytest.fixture(scope="session")
def gpu_lib():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import implisolid_b200
    return implisolid_b200.load_library()
