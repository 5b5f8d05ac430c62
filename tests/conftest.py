import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    import oracle_binding
    return oracle_binding.load()


@pytest.fixture(scope="session")
def engine():
    """One engine (log_blowup 1) shared by the GPU tests. Fails loudly if the library or the GPU is missing."""
    from openvm_b200 import B2Engine, FriParameters
    e = B2Engine.new(FriParameters(log_blowup=1))
    yield e
    e.close()


@pytest.fixture(scope="session")
def engine_b2():
    from openvm_b200 import B2Engine, FriParameters
    e = B2Engine.new(FriParameters.standard_with_100_bits_conjectured_security(2))
    yield e
    e.close()
