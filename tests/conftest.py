import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: full BASELINE.json sizes")


@pytest.fixture(scope="session")
def oracle():
    import oracle as _o

    _o.build()
    return _o.Oracle()


@pytest.fixture(scope="session")
def eigen_ref():
    import oracle as _o

    if not _o.EigenRef.available():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    return _o.EigenRef()
