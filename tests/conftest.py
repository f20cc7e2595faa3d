import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built_lib():
    """Make sure the C-ABI library exists (builds it if the .so is missing)."""
    lib = os.path.join(ROOT, "makani_b200", "libmakani_b200.so")
    if not os.path.exists(lib):
        import __graft_entry__ as g
        g.build()
    return lib
