import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def rt():
    import __graft_entry__ as g

    m = g.load_package()
    m.load()  # fails loudly if the CUDA extension has not been built
    return m


@pytest.fixture(scope="session")
def golden_dir():
    return ROOT / "tests" / "golden"
