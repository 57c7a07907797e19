import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    from loader import load_package
    return load_package()


@pytest.fixture(scope="session")
def scenes_dir():
    return os.path.join(ROOT, "scenes")


@pytest.fixture(scope="session")
def oracle():
    """The CPU checkers (test infrastructure): builds the C restatement if needed."""
    from oracle import oracle_py
    oracle_py.build()
    return oracle_py
