import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import _pkg
    return _pkg.load()


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure)."""
    import _pkg
    o = _pkg.load_oracle()
    o.build()
    return o


@pytest.fixture(scope="session")
def oracle(orc):
    return orc.Oracle()
