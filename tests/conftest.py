import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """Every test session starts from a built libtaxi_xf.so and oracle helper (nvcc/gcc work
    without a GPU); prebuilt files are reused when up to date."""
    import __graft_entry__ as entry
    entry.build()
