import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.cpu_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle.cpu_lib import Reference, reference_available
    if not reference_available():
        pytest.skip("oracle/_ref/libpafprocess_ref.so not built and /root/reference absent")
    return Reference()
