import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Make sure the product library and the oracle are built (the driver runs build() first; this is a safety net)."""
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "ngstrefftz_b200", "libngst_b200.so")) or not os.path.exists(
            os.path.join(ROOT, "oracle", "liboracle.so")):
        g.build()
    return True
