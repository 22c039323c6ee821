import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built_lib():
    """liborbx.so built in-tree (nvcc cross-compiles without a GPU)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_orbx_build", ROOT / "orb-slam3_b200" / "build.py")
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    return b.build()


@pytest.fixture(scope="session")
def pkg(built_lib):
    import _pkg
    return _pkg.load()


@pytest.fixture(scope="session")
def pkg_synth():
    """Synthetic-input generator only (no CUDA library involved)."""
    import _pkg
    return _pkg.load_synth_only().synth


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
